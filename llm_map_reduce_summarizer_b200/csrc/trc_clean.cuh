// K1a: clean_text as one sequential pass per segment.
// Replaces /root/reference/preprocessor.py:69-89 (and the blank-segment filter at :38):
//   pass 1  ' '.join(text.split())                      :80   Python isspace set, 29 code points
//   pass 2  re.sub(r'\b(\w+)( \1\b)+', r'\1', s)        :84   == local rule: a \w-run equal
//           to the \w-run that ends exactly one space before it is dropped together with that space
//   pass 3  re.sub(r'([.!?])([A-Za-z])', r'\1 \2', s)   :87   applied while emitting
// A word is emitted only when it ends (its bytes are contiguous in the input), so nothing ever has to
// be un-written and the count pass and the write pass run the identical state machine.
#pragma once
#include "trc_common.cuh"

struct TrcCleanOut {
  int64_t bytes;   // cleaned length in UTF-8 bytes (0 => segment is blank and is dropped)
  int64_t chars;   // cleaned length in code points (len(text) of the processed segment)
  int32_t spaces;  // U+0020 characters emitted (every one of them starts a pre-tokeniser match)
};

template <bool WRITE>
TRC_HD TrcCleanOut trc_clean_segment(const TrcTables& T, const uint8_t* in, int64_t n, uint8_t* out) {
  int64_t o = 0, nch = 0;
  int32_t nsp = 0;
  bool seen = false, pending = false, in_word = false, cand = false, w_pending = false;
  bool have_prev = false, gap0 = false;
  int64_t w_start = 0, w_chars = 0, p_start = 0, p_len = 0;
  uint8_t lastb = 0;

  auto end_word = [&](int64_t e) {
    int64_t len = e - w_start;
    bool dup = false;
    if (cand && len == p_len) {
      dup = true;
      for (int64_t k = 0; k < len; ++k)
        if (in[w_start + k] != in[p_start + k]) { dup = false; break; }
    }
    if (!dup) {
      if (w_pending) { if (WRITE) out[o] = ' '; ++o; ++nch; ++nsp; lastb = ' '; }
      uint8_t f = in[w_start];
      if ((lastb == '.' || lastb == '!' || lastb == '?') && ((f >= 'A' && f <= 'Z') || (f >= 'a' && f <= 'z'))) {
        if (WRITE) out[o] = ' ';
        ++o; ++nch; ++nsp;
      }
      if (WRITE) for (int64_t k = 0; k < len; ++k) out[o + k] = in[w_start + k];
      o += len; nch += w_chars;
      lastb = in[e - 1];
      have_prev = true; p_start = w_start; p_len = len;
    }
    gap0 = true;   // after a deletion the cursor sits right behind the previous (equal) word again
    in_word = false;
  };

  int64_t i = 0;
  while (i < n) {
    int len;
    uint32_t cp = trc_decode(in, i, n, &len);
    uint32_t pr = trc_prop(T, cp);
    if (pr & TRC_P_SPACE) {
      if (in_word) end_word(i);
      if (seen) pending = true;
    } else if (pr & TRC_P_WORD) {
      if (!in_word) {
        in_word = true; w_start = i; w_chars = 0;
        cand = have_prev && gap0 && pending;
        w_pending = pending; pending = false;
      }
      ++w_chars;
      seen = true;
    } else {
      if (in_word) end_word(i);
      if (pending) { if (WRITE) out[o] = ' '; ++o; ++nch; ++nsp; pending = false; }
      if (WRITE) for (int k = 0; k < len; ++k) out[o + k] = in[i + k];
      o += len; ++nch;
      lastb = in[i + len - 1];
      gap0 = false;
      seen = true;
    }
    i += len;
  }
  if (in_word) end_word(n);
  TrcCleanOut r; r.bytes = o; r.chars = nch; r.spaces = nsp;
  return r;
}
