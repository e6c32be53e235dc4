// clean_text as LOCAL rules over the bytes of a tile ("all three passes fuse into one
// local map + compaction").  Replaces the same reference lines as trc_clean.cuh
// (/root/reference/preprocessor.py:38,69-89) but every byte is decided independently, so a CTA can work on
// a tile with all lanes converged:
//   P1 cp_classify   per byte : Python isspace / \w flags (code point lookup for multi-byte characters)
//   P2 cp_mark_dup   per word : a \w-run equal to the \w-run that ends right before the whitespace in
//                               front of it is dropped together with the space that would separate them
//   P3 cp_code       per byte : 0 drop | 1 copy | 2 write ' ' (last character of a whitespace run that is
//                               neither leading nor trailing) | 3 write ' ' + byte ([.!?] then [A-Za-z])
// followed by a prefix sum of the output sizes and a scatter.  The sequential state machine in
// trc_clean.cuh stays as the definition used by the slow lane and by the tests (both must agree).
#pragma once
#include "trc_common.cuh"

#define CP_SPACE 1u
#define CP_WORD 2u
#define CP_CONT 4u
#define CP_DROP 0x80u        // in code[]: byte belongs to a deleted duplicate word
#define CP_DROPSPACE 0x40u   // in code[]: this whitespace character must not produce the joining space

// P1.  `in`/`cls`/`code` are tile-relative; [a,b) is the segment that contains byte k.
TRC_HD void cp_classify(const TrcTables& T, const uint8_t* in, uint8_t* cls, uint8_t* code, int32_t k, int32_t b) {
  uint8_t c = in[k];
  if (c < 0x80) { cls[k] = (uint8_t)(T.props[c] & 3u); code[k] = 0; return; }
  if (c < 0xC0) return;                       // continuation byte: written by its lead byte
  int len;
  uint32_t cp = trc_decode(in, k, b, &len);
  uint8_t f = (uint8_t)(T.props[cp] & 3u);
  cls[k] = f; code[k] = 0;
  for (int q = 1; q < len; ++q) { cls[k + q] = (uint8_t)(f | CP_CONT); code[k + q] = 0; }
}

// P2.  Called for every byte; acts only on the first byte of a \w-run.
TRC_HD void cp_mark_dup(const uint8_t* in, const uint8_t* cls, uint8_t* code, int32_t k, int32_t a, int32_t b) {
  uint8_t c = cls[k];
  if (!(c & CP_WORD) || (c & CP_CONT)) return;
  if (k == a || !(cls[k - 1] & CP_SPACE)) return;            // must follow whitespace
  int32_t p = k;
  while (p > a && (cls[p - 1] & CP_SPACE)) --p;              // start of the whitespace run
  if (p == a || !(cls[p - 1] & CP_WORD)) return;             // ... which must follow a \w character
  int32_t e = k;
  while (e < b && (cls[e] & CP_WORD)) ++e;
  int32_t L = e - k;
  if (p - L < a) return;
  if (p - L > a && (cls[p - L - 1] & CP_WORD)) return;       // previous \w-run is longer
  for (int32_t i = 0; i < L; ++i) if (in[p - L + i] != in[k + i]) return;
  for (int32_t i = k; i < e; ++i) code[i] = CP_DROP;
  int32_t q = k - 1;
  while (cls[q] & CP_CONT) --q;
  code[q] = CP_DROPSPACE;
}

// Cheap filter in front of cp_mark_dup: first byte of a \w-run that directly follows whitespace.
TRC_HD bool cp_dup_candidate(const uint8_t* cls, int32_t k, int32_t a) {
  uint8_t c = cls[k];
  return (c & CP_WORD) && !(c & CP_CONT) && k > a && (cls[k - 1] & CP_SPACE);
}

// First non-whitespace byte of segment [a,b) (== b for a blank segment): whitespace before it is leading.
TRC_HD int32_t cp_first_nonspace(const uint8_t* cls, int32_t a, int32_t b) {
  int32_t p = a;
  while (p < b && (cls[p] & CP_SPACE)) ++p;
  return p;
}

// P3.  Returns the output code of byte k (and stores it).  fns = cp_first_nonspace of the segment.
TRC_HD uint32_t cp_code(const uint8_t* in, const uint8_t* cls, uint8_t* code, int32_t k, int32_t a, int32_t b, int32_t fns) {
  uint8_t c = cls[k], d = code[k];
  uint32_t r;
  if (d == CP_DROP) r = 0;
  else if (c & CP_SPACE) {
    r = 0;
    if (!(c & CP_CONT) && d != CP_DROPSPACE) {
      int32_t kn = k + trc_u8len(in[k]);
      if (kn < b && !(cls[kn] & CP_SPACE) && k > fns) r = 2;
    }
  } else {
    uint8_t ch = in[k];
    r = 1;
    if (k > a && (((ch | 0x20u) - 'a') < 26u) && ch < 0x80) {
      uint8_t pv = in[k - 1];
      if (pv == '.' || pv == '!' || pv == '?') r = 3;
    }
  }
  code[k] = (uint8_t)r;
  return r;
}

TRC_HD int cp_outlen(uint32_t code) { return code == 3 ? 2 : (code ? 1 : 0); }
