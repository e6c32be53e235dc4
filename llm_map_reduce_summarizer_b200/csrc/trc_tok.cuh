// K2: token counting = cl100k pre-tokenisation (hand matcher with the regex's leftmost-first
// alternation semantics) + byte-pair merge by minimum rank, leftmost on ties.
// Replaces `len(self.tokenizer.encode(x))` at /root/reference/big_chunkeroosky.py:83,298,370,510, i.e. the
// Rust CoreBPE of tiktoken==0.5.1 (third party, not under /root/reference; algorithm restated from
// tiktoken/_educational.py:83-116 and the 0.5.1 pattern string, see oracle/tokenizer.py).
//
// A "unit" is a UTF-8 byte range with an optional VIRTUAL leading U+0020: every string the reference
// tokenises decomposes at guaranteed match boundaries into  "[ts]" + " speaker:" + " text"  (DESIGN.md
// §3), so the kernels count  " " + text  without ever materialising the concatenation.
#pragma once
#include "trc_common.cuh"

struct TrcStr {        // logical string = [' '] + p[0..n)
  const uint8_t* p;
  int64_t n;           // logical length (including the virtual space)
  int vs;              // 1 if a virtual leading space is present
  TRC_HD uint8_t at(int64_t i) const { return (vs && i == 0) ? (uint8_t)' ' : p[i - vs]; }
};

TRC_HD uint32_t trc_sdecode(const TrcStr& S, int64_t i, int* len) {
  if (S.vs && i == 0) { *len = 1; return ' '; }
  return trc_decode(S.p, i - S.vs, S.n - S.vs, len);
}

// End (exclusive, logical index) of the regex match that starts at logical index p.
// Alternatives in pattern order:
//  1 (?i:'s|'t|'re|'ve|'m|'ll|'d)   2 [^\r\n\p{L}\p{N}]?\p{L}+   3 \p{N}{1,3}
//  4  ?[^\s\p{L}\p{N}]+[\r\n]*      5 \s*[\r\n]+   6 \s+(?!\S)   7 \s+
TRC_HD uint64_t trc_swar_letters(uint64_t x) {          // bit 7 of every byte that is an ASCII letter
  const uint64_t H = 0x8080808080808080ULL, L = 0x0101010101010101ULL;
  const uint64_t t = ((x | 0x2020202020202020ULL) & ~H) | H;   // lower-cased, high bit forced
  const uint64_t ge_a = (t - L * 0x61) & H, ge_z1 = (t - L * 0x7b) & H;
  return ge_a & ~ge_z1 & ~(x & H);
}
TRC_HD int64_t trc_span_end(const TrcTables& T, const TrcFast& F, const TrcStr& S, int64_t p) {
  const int64_t N = S.n;
  int l0, l1 = 0, l2 = 0;
  uint32_t c0 = trc_sdecode(S, p, &l0);
  uint32_t k0 = trc_cls(T, F, c0);
  bool has1 = p + l0 < N;
  uint32_t c1 = 0, k1 = 9;
  if (has1) { c1 = trc_sdecode(S, p + l0, &l1); k1 = trc_cls(T, F, c1); }
  if (c0 == '\'' && has1) {
    uint32_t a = c1;
    if (a == 's' || a == 'S' || a == 't' || a == 'T' || a == 'm' || a == 'M' || a == 'd' || a == 'D' || a == 0x17F)
      return p + l0 + l1;
    if (p + l0 + l1 < N) {
      uint32_t b = trc_sdecode(S, p + l0 + l1, &l2);
      bool e = (b == 'e' || b == 'E');
      if (((a == 'r' || a == 'R') && e) || ((a == 'v' || a == 'V') && e) ||
          ((a == 'l' || a == 'L') && (b == 'l' || b == 'L')))
        return p + l0 + l1 + l2;
    }
  }
  {  // alt 2
    int64_t q = p;
    uint32_t kq = k0;
    if (k0 != TRC_RS_L && k0 != TRC_RS_N && c0 != '\r' && c0 != '\n' && has1 && k1 == TRC_RS_L) { q = p + l0; kq = k1; }
    if (kq == TRC_RS_L) {
      while (q < N) {
        if (q >= S.vs && q + 8 <= N) {         // eight ASCII letters at a time (a 64 KB run of letters paid a load latency per byte)
          const uint8_t* a8 = S.p + (q - S.vs);
          if (((uintptr_t)a8 & 7) == 0 && trc_swar_letters(*reinterpret_cast<const uint64_t*>(a8)) == 0x8080808080808080ULL) { q += 8; continue; }
        }
        int l; uint32_t c = trc_sdecode(S, q, &l);
        if (trc_cls(T, F, c) != TRC_RS_L) break;
        q += l;
      }
      return q;
    }
  }
  if (k0 == TRC_RS_N) {  // alt 3
    int64_t q = p + l0;
    int cnt = 1;
    while (q < N && cnt < 3) {
      int l; uint32_t c = trc_sdecode(S, q, &l);
      if (trc_cls(T, F, c) != TRC_RS_N) break;
      q += l; ++cnt;
    }
    return q;
  }
  {  // alt 4
    int64_t q = p;
    uint32_t kq = k0;
    if (c0 == ' ' && has1 && k1 == TRC_RS_O) { q = p + l0; kq = k1; }
    if (kq == TRC_RS_O) {
      while (q < N) {
        int l; uint32_t c = trc_sdecode(S, q, &l);
        if (trc_cls(T, F, c) != TRC_RS_O) break;
        q += l;
      }
      while (q < N) { uint8_t b = S.at(q); if (b != '\r' && b != '\n') break; ++q; }
      return q;
    }
  }
  // whitespace run [p, r)
  int64_t r = p, last_nl = -1, last_start = p;
  while (r < N) {
    int l; uint32_t c = trc_sdecode(S, r, &l);
    if (trc_cls(T, F, c) != TRC_RS_S) break;
    if (c == '\r' || c == '\n') last_nl = r;
    last_start = r;
    r += l;
  }
  if (last_nl >= 0) return last_nl + 1;      // alt 5: through the last newline of the run
  if (r == N) return r;                      // alt 6 at end of string
  if (last_start > p) return last_start;     // alt 6: leave the last whitespace char to the next match
  return r;                                  // alt 7
}

// Is the byte string a vocabulary entry?  (tiktoken: `if piece in ranks` => one token)
TRC_HD bool trc_is_token_generic(const TrcTables& T, const TrcStr& S, int64_t a, int n);

// Exact membership test with the <=16-byte key sets; longer spans (and spans holding a NUL byte) use the
// hashed byte-string table.
TRC_HD bool trc_is_token(const TrcTables& T, const TrcFast& F, const TrcStr& S, int64_t a, int n) {
  if ((uint32_t)n > T.tok_max_len) return false;
  if (n <= 16) {
    uint64_t lo = 0, hi = 0;
    bool nul = false;
    for (int i = 0; i < n; ++i) {
      uint64_t b = S.at(a + i);
      nul |= (b == 0);
      if (i < 8) lo |= b << (8 * i); else hi |= b << (8 * (i - 8));
    }
    if (!nul) {
      if (n <= 8) {
        if (F.hot8) { const uint32_t hs = trc_key8_slot(lo, F.hot8_mask); if (F.hot8[hs] == lo || F.hot8[hs ^ 1] == lo) return true; }
        uint32_t slot = trc_key8_slot(lo, T.tok8_mask);
        for (;;) {
          uint64_t e = T.tok8[slot];
          if (e == lo) return true;
          if (e == 0) return false;
          slot = (slot + 1) & T.tok8_mask;
        }
      }
      uint32_t slot = trc_key8_slot(lo ^ (hi * 0x9E3779B97F4A7C15ULL), T.tok16_mask);
      for (;;) {
        uint64_t e0 = T.tok16[2 * (size_t)slot], e1 = T.tok16[2 * (size_t)slot + 1];
        if (e0 == lo && e1 == hi) return true;
        if (e0 == 0 && e1 == 0) return false;
        slot = (slot + 1) & T.tok16_mask;
      }
    }
  }
  return trc_is_token_generic(T, S, a, n);
}

TRC_HD bool trc_is_token_generic(const TrcTables& T, const TrcStr& S, int64_t a, int n) {
  uint64_t h = trc_hash_bytes([&](int i) { return S.at(a + i); }, n);
  uint32_t fp = (uint32_t)(h >> 32);
  uint32_t slot = (uint32_t)h & T.tokhash_mask;
  for (;;) {
    uint64_t e = T.tokhash[slot];
    if (e == ~0ULL) return false;
    if ((uint32_t)(e >> 32) == fp) {
      uint32_t id = (uint32_t)e;
      uint32_t o = T.tok_off[id];
      if ((int)(T.tok_off[id + 1] - o) == n) {
        bool eq = true;
        for (int i = 0; i < n; ++i) if (T.tok_bytes[o + i] != S.at(a + i)) { eq = false; break; }
        if (eq) return true;
      }
    }
    slot = (slot + 1) & T.tokhash_mask;
  }
}

// Byte-pair merge of a short span held in thread-local arrays; returns the number of parts left.
TRC_HD int trc_bpe_short(const TrcTables& T, const uint32_t* byte_id, const TrcStr& S, int64_t a, int n) {
  uint32_t id[TRC_SHORT_SPAN];
  uint32_t rk[TRC_SHORT_SPAN];
  for (int i = 0; i < n; ++i) id[i] = byte_id[S.at(a + i)];
  int m = n;
  for (int i = 0; i + 1 < m; ++i) rk[i] = trc_pair_rank(T, id[i], id[i + 1]);
  for (;;) {
    uint32_t best = TRC_NO_RANK; int bi = -1;
    for (int i = 0; i + 1 < m; ++i) if (rk[i] < best) { best = rk[i]; bi = i; }
    if (bi < 0) break;
    id[bi] = best;                       // rank of the merged bytes == its token id
    for (int j = bi + 1; j + 1 < m; ++j) { id[j] = id[j + 1]; rk[j] = rk[j + 1]; }
    --m;
    if (bi + 1 < m) rk[bi] = trc_pair_rank(T, id[bi], id[bi + 1]); else rk[bi] = TRC_NO_RANK;
    if (bi > 0) rk[bi - 1] = trc_pair_rank(T, id[bi - 1], id[bi]);
  }
  return m;
}

// Index of a special-token literal contained in the unit's real bytes, or -1.
TRC_HD int trc_find_special(const TrcTables& T, const uint8_t* p, int64_t n) {
  for (int64_t i = 0; i < n; ++i) {
    if (T.sp_single < 0x100u && i + 8 <= n && (((uintptr_t)(p + i)) & 7) == 0) {
      // every literal starts with the same ASCII byte: eight bytes without it are skipped at once (a 64 KB unit was a load
      // latency per byte for the one thread that walks it)
      uint64_t x;
#if defined(__CUDA_ARCH__)
      x = *reinterpret_cast<const unsigned long long*>(p + i);
#else
      memcpy(&x, p + i, 8);
#endif
      const uint64_t t = x ^ (0x0101010101010101ULL * (uint64_t)T.sp_single);
      if (!((t - 0x0101010101010101ULL) & ~t & 0x8080808080808080ULL)) { i += 7; continue; }
    }
    uint8_t b = p[i];
    if (!((T.sp_first[b >> 6] >> (b & 63)) & 1ULL)) continue;
    for (uint32_t k = 0; k < T.n_sp; ++k) {
      uint32_t o = T.sp_off[k], L = T.sp_off[k + 1] - o;
      if (L == 0 || T.sp_bytes[o] != b || i + (int64_t)L > n) continue;
      bool eq = true;
      for (uint32_t j = 1; j < L; ++j) if (p[i + j] != T.sp_bytes[o + j]) { eq = false; break; }
      if (eq) return (int)k;
    }
  }
  return -1;
}

// ---- whole-word fast path on 16 packed bytes --------------------------------------------------------------
// A "word" here is a space (real or virtual) followed by non-space bytes, at most 16 bytes in total, given
// as two little-endian 64-bit words (zero padded).  Handles the shapes that make up ordinary cleaned text:
//   ' ' letters                      -> one span (alt 2)
//   ' ' letters + one ASCII "other"  -> alt 2 + alt 4 of length 1 (trailing , . ? ! ; : ' ...)
//   ' ' letters + contraction        -> alt 2 + alt 1 ('s 't 'm 'd 'll 've 're, any case)
// Letter spans must be vocabulary entries (exact-key sets); anything else returns false and the caller
// uses the generic matcher.  Exact: every decision is the regex's decision for these shapes.

TRC_HD bool trc_key_is_token(const TrcTables& T, const TrcFast& F, uint64_t lo, uint64_t hi, int n) {
  if (n <= 8) {
    if (F.hot8) { const uint32_t hs = trc_key8_slot(lo, F.hot8_mask); if (F.hot8[hs] == lo || F.hot8[hs ^ 1] == lo) return true; }
    uint32_t slot = trc_key8_slot(lo, T.tok8_mask);
    for (;;) { uint64_t e = T.tok8[slot]; if (e == lo) return true; if (e == 0) return false; slot = (slot + 1) & T.tok8_mask; }
  }
  uint32_t slot = trc_key8_slot(lo ^ (hi * 0x9E3779B97F4A7C15ULL), T.tok16_mask);
  for (;;) {
    uint64_t e0 = T.tok16[2 * (size_t)slot], e1 = T.tok16[2 * (size_t)slot + 1];
    if (e0 == lo && e1 == hi) return true;
    if (e0 == 0 && e1 == 0) return false;
    slot = (slot + 1) & T.tok16_mask;
  }
}

// byte i of the 16 packed bytes
TRC_HD uint8_t trc_b16(uint64_t lo, uint64_t hi, int i) { return i < 8 ? (uint8_t)(lo >> (8 * i)) : (uint8_t)(hi >> (8 * (i - 8))); }
// bytes [a, a+n) of the 16 packed bytes as a zero-padded key (n <= 16)
TRC_HD void trc_sub16(uint64_t lo, uint64_t hi, int a, int n, uint64_t* klo, uint64_t* khi) {
  uint64_t l = lo, h = hi;
  if (a >= 8) { l = hi >> (8 * (a - 8)); h = 0; }
  else if (a > 0) { l = (lo >> (8 * a)) | (hi << (8 * (8 - a))); h = hi >> (8 * a); }
  if (n < 8) { l &= (1ULL << (8 * n)) - 1; h = 0; } else if (n == 8) h = 0; else if (n < 16) h &= (1ULL << (8 * (n - 8))) - 1;
  *klo = l; *khi = h;
}

// Tokens of an all-ASCII word of at most 16 bytes whose first byte is the space that precedes it: the
// regex alternatives 1-4 evaluated on class bit masks (letters / digits / other), one vocabulary probe
// per span.  Returns false (caller falls back to the generic matcher) when a byte is not ASCII, a
// whitespace or control-class byte appears inside the word, or a multi-byte span is not a vocabulary
// entry (it then needs byte-pair merging).
TRC_HD bool trc_count_word16(const TrcTables& T, const TrcFast& F, uint64_t lo, uint64_t hi, int n, uint32_t* count) {
  if (n < 2 || n > 16 || (lo & 0xff) != ' ') return false;
  const uint64_t H = 0x8080808080808080ULL, L = 0x0101010101010101ULL;
  if ((lo | hi) & H) return false;                                         // non-ASCII byte
  const uint32_t all = (n == 16) ? 0xffffu : ((1u << n) - 1u);
  const uint32_t lm = sw_pack16(trc_swar_letters(lo), trc_swar_letters(hi)) & all;
  const uint64_t tl = lo | H, th = hi | H;
  const uint32_t dm = sw_pack16(((tl - L * 0x30) & H) & ~((tl - L * 0x3a) & H), ((th - L * 0x30) & H) & ~((th - L * 0x3a) & H)) & all;
  const uint32_t ctl = sw_pack16(~((tl - L * 0x21) & H) & H, ~((th - L * 0x21) & H) & H) & all & ~1u;   // bytes < 0x21 after the space
  if (ctl) return false;
  if (sw_pack16(~((tl - L * 0x7f) & H) & H, ~((th - L * 0x7f) & H) & H) != 0xffffu) return false;         // DEL
  uint32_t total = 0;
  int p = 0;
  while (p < n) {
    const uint32_t rest_l = lm >> p, rest_d = dm >> p;
    int len;
    const uint8_t c0 = trc_b16(lo, hi, p);
    if (rest_l & 1u) {                                                     // alt 2: letters
      len = (int)trc_ctz32(~rest_l);
    } else if (c0 == '\'' && p + 1 < n && (rest_l & 2u)) {                 // alt 1 before alt 2
      const uint8_t a1 = trc_b16(lo, hi, p + 1) | 0x20, a2 = (p + 2 < n) ? (uint8_t)(trc_b16(lo, hi, p + 2) | 0x20) : 0;
      if (a1 == 's' || a1 == 't' || a1 == 'm' || a1 == 'd') len = 2;
      else if ((a1 == 'l' && a2 == 'l') || (a1 == 'v' && a2 == 'e') || (a1 == 'r' && a2 == 'e')) len = ((lm >> (p + 2)) & 1u) ? 3 : 0;
      else len = 0;
      if (len == 0) len = 1 + (int)trc_ctz32(~(rest_l >> 1));              // alt 2: "'" + letters
    } else if (rest_d & 1u) {                                              // alt 3: up to three digits
      len = (int)trc_ctz32(~rest_d); if (len > 3) len = 3;
    } else if (rest_l & 2u) {                                              // alt 2: one other byte (or the space) + letters
      len = 1 + (int)trc_ctz32(~(rest_l >> 1));
    } else {                                                               // alt 4: [space] other+   /  alt 7: lone space
      const uint32_t om = ~(lm | dm) & all;                                // other bytes (the space at 0 is handled by q)
      int q = p;
      if (p == 0) { if (n > 1 && ((om >> 1) & 1u)) q = 1; else { total += 1; p = 1; continue; } }   // ' ' + other, or ' ' alone (before a digit)
      len = (q - p) + (int)trc_ctz32(~(om >> q));
    }
    if (len == 1) total += 1;
    else {
      uint64_t kl, kh;
      trc_sub16(lo, hi, p, len, &kl, &kh);
      if (!trc_key_is_token(T, F, kl, kh, len)) return false;
      total += 1;
    }
    p += len;
  }
  *count = total;
  return true;
}

// Fast paths for the three span shapes that make up ordinary text (all ASCII, exact).  Returns the end
// of the span that starts at q and adds its token count to *total, or -1 when the generic matcher must
// decide.  With MERGE=false a short letter span that is not a vocabulary entry also returns -1 (the
// caller queues the word for the generic path instead of merging inline).
template <bool MERGE>
TRC_HD int64_t trc_fast_span(const TrcTables& T, const TrcFast& F, const uint32_t* byte_id, const TrcStr& S, int64_t q, uint32_t* total) {
  const uint8_t c0 = S.at(q);
  const uint32_t k0 = c0 < 0x80 ? (uint32_t)F.acls[c0] : 9u;
  if (c0 == ' ' || k0 == TRC_RS_L) {
    // alt 2: optional ' ' then ASCII letters, at most 8 bytes, ended by an ASCII non-letter or the end
    uint64_t key = c0; int k = 1; uint8_t b = 0; bool more = false;
    while ((more = (q + k < S.n)) && k < 9) {
      b = S.at(q + k);
      if (b >= 0x80 || F.acls[b] != TRC_RS_L) break;
      if (k < 8) key |= (uint64_t)b << (8 * k);
      ++k;
    }
    const int nlet = k - (c0 == ' ' ? 1 : 0);
    if (nlet >= 1 && k <= 8 && (!more || b < 0x80)) {
      bool hit = k == 1;
      if (!hit) {
        uint32_t hs_ = F.hot8 ? trc_key8_slot(key, F.hot8_mask) : 0u;
        if (F.hot8 && (F.hot8[hs_] == key || F.hot8[hs_ ^ 1] == key)) hit = true;
        else {
          uint32_t slot = trc_key8_slot(key, T.tok8_mask);
          for (;;) { uint64_t e8 = T.tok8[slot]; if (e8 == key) { hit = true; break; } if (e8 == 0) break; slot = (slot + 1) & T.tok8_mask; }
        }
      }
      if (hit) *total += 1u;
      else if (MERGE) *total += (uint32_t)trc_bpe_short(T, byte_id, S, q, k);
      else return -1;
      return q + k;
    }
  } else if (k0 == TRC_RS_O) {
    // alt 4 of length 1: an ASCII punctuation byte followed by a digit, whitespace (not a newline) or the end
    bool single = q + 1 == S.n;
    if (!single) { uint8_t b = S.at(q + 1); single = b < 0x80 && b != '\r' && b != '\n' && (F.acls[b] == TRC_RS_N || F.acls[b] == TRC_RS_S); }
    if (single) { *total += 1; return q + 1; }
  } else if (k0 == TRC_RS_N) {
    // alt 3: up to three ASCII digits; the count comes from the digit-group table
    int k = 1, v = c0 - '0'; bool ok = true;
    while (k < 3 && q + k < S.n) {
      uint8_t b = S.at(q + k);
      if (b >= 0x80) { ok = false; break; }
      if (F.acls[b] != TRC_RS_N) break;
      v = v * 10 + (b - '0'); ++k;
    }
    if (ok && T.dcount_ready) { *total += T.dcount[v + (k == 1 ? 0 : (k == 2 ? 10 : 110))]; return q + k; }
  }
  return -1;
}

// A whole word through the fast paths only: false when any of its spans needs the generic matcher.
TRC_HD bool trc_count_word_fast(const TrcTables& T, const TrcFast& F, const uint8_t* p, int64_t nbytes, int vs, uint32_t* count) {
  TrcStr S; S.p = p; S.n = nbytes + vs; S.vs = vs;
  uint32_t total = 0;
  int64_t q = 0;
  while (q < S.n) {
    q = trc_fast_span<false>(T, F, nullptr, S, q, &total);
    if (q < 0) return false;
  }
  *count = total;
  return true;
}

// Token count of one unit.  Spans longer than TRC_SHORT_SPAN that are not vocabulary entries are handed
// to `long_span(logical_start, length)` which returns their count (device: deferred to the block-wide
// long-span kernel and returns 0; hostsim: computed inline).
template <class LongSpan>
TRC_HD uint32_t trc_count_unit(const TrcTables& T, const TrcFast& F, const uint32_t* byte_id, const uint8_t* p, int64_t nbytes,
                               int vs, LongSpan&& long_span) {
  TrcStr S; S.p = p; S.n = nbytes + vs; S.vs = vs;
  uint32_t total = 0;
  int64_t q = 0;
  while (q < S.n) {
    if (S.at(q) == ' ') {                                    // whole-word fast path (also the fused kernel's)
      uint64_t lo = ' ', hi = 0; int k = 1; bool fits = true;
      while (q + k < S.n) {
        const uint8_t b = S.at(q + k);
        if (b == ' ') break;
        if (k >= 16) { fits = false; break; }
        if (k < 8) lo |= (uint64_t)b << (8 * k); else hi |= (uint64_t)b << (8 * (k - 8));
        ++k;
      }
      uint32_t c = 0;
      if (fits && trc_count_word16(T, F, lo, hi, k, &c)) { total += c; q += k; continue; }
    }
    {
      int64_t nq = trc_fast_span<true>(T, F, byte_id, S, q, &total);
      if (nq >= 0) { q = nq; continue; }
    }
    int64_t e = trc_span_end(T, F, S, q);
    int64_t len = e - q;
    if (len == 1) total += 1;
    else if (len <= 0x7fffffff && trc_is_token(T, F, S, q, (int)len)) total += 1;
    else if (len <= TRC_SHORT_SPAN) total += (uint32_t)trc_bpe_short(T, byte_id, S, q, (int)len);
    else total += long_span(q, len);
    q = e;
  }
  return total;
}
