// clean_text on 32-byte CHUNKS: the local rules of trc_swar.cuh (/root/reference/preprocessor.py:38,69-89) with one
// 32-bit mask word per class and chunk, bit i of a word <-> byte i of the chunk.  One lane of the warp-tile kernel
// (trc_fused4.cuh) owns a chunk: a mask operation decides 32 bytes, where the 8-byte cells of trc_swar.cuh spent the
// same instruction on 8.  The routines are plain functions over mask arrays so that the host test build can run
// them tile by tile (hostsim_clean_w32 in trchunk.cu) against the oracle.
//
// Mask arrays are indexed by chunk, from -1 to nch inclusive; entries -1 and nch are zero ("nothing there").
//   sp  Python-whitespace bytes (every byte of a multi-byte whitespace character)      w   \w bytes (likewise)
//   ct  UTF-8 continuation bytes      ep  '.', '!' or '?'      lt  ASCII letters      bd  first byte of a segment
//   drop / ds   set by the duplicate-word rule: bytes of a deleted word / the whitespace character in front of it
#pragma once
#include "trc_common.cuh"

struct W32 {
  uint32_t* sp; uint32_t* w; uint32_t* ct; uint32_t* bd; uint32_t* ep; uint32_t* lt; uint32_t* drop; uint32_t* ds;
  int32_t nch;
  const uint32_t* nolead = nullptr;   // optional: segment starts whose leading whitespace is NOT stripped (the segment
                                      // continues an oversized segment that was cut in front of a whitespace run)
};

TRC_HD uint32_t w32_bit(const uint32_t* a, int32_t k) { return (a[k >> 5] >> (k & 31)) & 1u; }     // k >= -32

TRC_HD void w32_or(uint32_t* a, int32_t chunk, uint32_t bits) {
#if defined(__CUDA_ARCH__)
  atomicOr(a + chunk, bits);
#else
  a[chunk] |= bits;
#endif
}
TRC_HD void w32_or_bit(uint32_t* a, int32_t k) { w32_or(a, k >> 5, 1u << (k & 31)); }

// ---- classification table: one 64-bit entry per byte value, eight 8-bit planes.  A group of eight bytes is
// classified by acc = 2 * acc + entry[byte] from its last byte to its first: plane p of acc then holds, in bit i,
// class p of byte i.  Planes: 0 sp  1 w  2 lt  3 ep | 4 non-ASCII  5 continuation  6 first byte of a special-token
// literal  7 any whitespace byte that is not U+0020 (its emitted replacement differs from the byte itself).
#define W32_P_SP 0
#define W32_P_W 1
#define W32_P_LT 2
#define W32_P_EP 3
#define W32_P_NA 4
#define W32_P_CT 5
#define W32_P_SF 6
#define W32_P_NS 7

TRC_HD uint64_t w32_lut_entry(const TrcTables& T, uint32_t b) {
  uint64_t e = 0;
  if (b < 0x80u) {
    const uint32_t pr = T.props[b];
    const bool lt = (b >= 'A' && b <= 'Z') || (b >= 'a' && b <= 'z');
    if (pr & TRC_P_SPACE) { e |= 1ULL << (8 * W32_P_SP); if (b != ' ') e |= 1ULL << (8 * W32_P_NS); }
    if (pr & TRC_P_WORD) e |= 1ULL << (8 * W32_P_W);
    if (lt) e |= 1ULL << (8 * W32_P_LT);
    if (b == '.' || b == '!' || b == '?') e |= 1ULL << (8 * W32_P_EP);
  } else {
    e |= 1ULL << (8 * W32_P_NA);
    if (b < 0xC0u) e |= 1ULL << (8 * W32_P_CT);
  }
  if ((T.sp_first[b >> 6] >> (b & 63)) & 1ULL) e |= 1ULL << (8 * W32_P_SF);
  return e;
}

struct W32Cls { uint32_t sp, w, lt, ep, na, ct, sf, ns; };

// The 32 bytes of a chunk given as eight little-endian words.
TRC_HD W32Cls w32_classify(const uint64_t* lut, const uint32_t x[8]) {
  uint32_t alo[4], ahi[4];
#pragma unroll
  for (int g = 0; g < 4; ++g) {
    uint32_t lo = 0, hi = 0;
#pragma unroll
    for (int q = 7; q >= 0; --q) {
      const uint32_t b = (x[2 * g + (q >> 2)] >> (8 * (q & 3))) & 0xffu;
      const uint64_t e = lut[b];
      lo = lo * 2u + (uint32_t)e;
      hi = hi * 2u + (uint32_t)(e >> 32);
    }
    alo[g] = lo; ahi[g] = hi;
  }
  W32Cls c;
#define W32_GATHER(arr, p) ((((arr)[0] >> (8 * (p))) & 0xffu) | ((((arr)[1] >> (8 * (p))) & 0xffu) << 8) | ((((arr)[2] >> (8 * (p))) & 0xffu) << 16) | ((((arr)[3] >> (8 * (p))) & 0xffu) << 24))
  c.sp = W32_GATHER(alo, 0); c.w = W32_GATHER(alo, 1); c.lt = W32_GATHER(alo, 2); c.ep = W32_GATHER(alo, 3);
  c.na = W32_GATHER(ahi, 0); c.ct = W32_GATHER(ahi, 1); c.sf = W32_GATHER(ahi, 2); c.ns = W32_GATHER(ahi, 3);
#undef W32_GATHER
  return c;
}

// bytes of the tile that belong to its own segments, [lo, hi), as a mask of chunk c
TRC_HD uint32_t w32_valid(int32_t c, int32_t lo, int32_t hi) {
  const int32_t k0 = c << 5;
  uint32_t v = 0xffffffffu;
  if (k0 < lo) v &= (lo - k0 >= 32) ? 0u : (0xffffffffu << (lo - k0));
  if (k0 + 32 > hi) v &= (hi <= k0) ? 0u : (0xffffffffu >> (k0 + 32 - hi));
  return v;
}

// Multi-byte characters: the owner of the lead byte looks the code point up and sets sp / w on every byte of the
// character (they may spill into the next chunk).  `lead` = lead bytes of this chunk inside [lo, hi).
TRC_HD void w32_fix_multibyte(const TrcTables& T, const uint8_t* in, const W32& M, int32_t c, uint32_t lead, int32_t hi) {
  while (lead) {
    const int bit = (int)trc_ctz32(lead);
    lead &= lead - 1;
    const int32_t k = (c << 5) + bit;
    int len;
    const uint32_t cp = trc_decode(in, k, hi, &len);
    const uint32_t pr = T.props[cp];
    if (pr & (TRC_P_SPACE | TRC_P_WORD)) {
      uint32_t* dst = (pr & TRC_P_SPACE) ? M.sp : M.w;
      for (int q = 0; q < len; ++q) w32_or_bit(dst, k + q);
    }
  }
}

// First bytes of \w-runs that directly follow whitespace and do not start a segment.
TRC_HD uint32_t w32_dup_candidates(const W32& M, int32_t c, uint32_t valid) {
  const uint32_t w = M.w[c], sp = M.sp[c];
  const uint32_t wstart = w & ~((w << 1) | (M.w[c - 1] >> 31));
  const uint32_t after_ws = (sp << 1) | (M.sp[c - 1] >> 31);
  return wstart & after_ws & ~M.bd[c] & valid;
}

// Number of consecutive set bits of a mask array from bit k on, at most up to `limit`.
TRC_HD int32_t w32_run_up(const uint32_t* a, int32_t k, int32_t limit) {
  int32_t e = k;
  while (e < limit) {
    const uint32_t x = a[e >> 5] >> (e & 31);
    const int room = 32 - (e & 31);
    const uint32_t inv = ~x & (room == 32 ? 0xffffffffu : ((1u << room) - 1u));
    if (inv) { e += (int32_t)trc_ctz32(inv); break; }
    e += room;
  }
  return e < limit ? e : limit;
}

// Cheap necessary condition (see sw_dup_quick_reject): with a single one-byte separator and a run that ends within
// 24 bytes inside its segment, equal words end with the same byte.  true = the candidate can be discarded.
TRC_HD bool w32_dup_quick_reject(const uint8_t* in, const W32& M, int32_t k, int32_t lim) {
  if (k < 2 || w32_bit(M.sp, k - 2) || w32_bit(M.ct, k - 1)) return false;
  const int32_t c = k >> 5, bit = k & 31;
  const uint64_t wwin = (((uint64_t)M.w[c + 1] << 32) | (uint64_t)M.w[c]) >> bit;
  const uint32_t low = (uint32_t)wwin;
  if ((low & 0xffffffu) == 0xffffffu) return false;                        // a long run: the full test decides
  const uint32_t L = trc_ctz32(~low);
  if (k + (int32_t)L > lim) return false;
  const uint32_t bwin = (uint32_t)(((((uint64_t)M.bd[c + 1] << 32) | (uint64_t)M.bd[c]) >> bit));
  if (bwin & ((1u << L) - 1u) & ~1u) return false;                         // a segment starts inside the run
  return in[k + L - 1] != in[k - 2];
}

// The duplicate-word rule at byte k, which starts a \w-run after whitespace inside segment [a, b).
TRC_HD void w32_mark_dup(const uint8_t* in, const W32& M, int32_t k, int32_t a, int32_t b) {
  int32_t p = k;
  while (p > a && w32_bit(M.sp, p - 1)) --p;                 // start of the whitespace run
  if (p == a || !w32_bit(M.w, p - 1)) return;                // ... which must follow a \w character
  const int32_t e = w32_run_up(M.w, k, b);
  if (in[e - 1] != in[p - 1]) return;
  const int32_t L = e - k;
  if (p - L < a) return;
  if (p - L > a && w32_bit(M.w, p - L - 1)) return;          // the previous \w-run is longer
  for (int32_t i = 0; i < L; ++i) if (in[p - L + i] != in[k + i]) return;
  for (int32_t i = k; i < e; ) {                             // drop [k, e)
    const int32_t cc = i >> 5, b0 = i & 31;
    const int32_t n = (e - i) < (32 - b0) ? (e - i) : (32 - b0);
    w32_or(M.drop, cc, (n == 32 ? 0xffffffffu : ((1u << n) - 1u)) << b0);
    i += n;
  }
  int32_t q = k - 1;
  while (w32_bit(M.ct, q)) --q;
  w32_or_bit(M.ds, q);
}

// All duplicate-word candidates of chunk c.  The common shape - one single-byte whitespace character between two
// \w-runs of fewer than 24 bytes inside one segment - is decided from the chunk's own mask words and its two
// neighbours' (held in registers by the caller's lane) plus a few byte loads; anything else goes through
// w32_mark_dup.  seg_of(k) returns the index i of the segment that contains byte k (rel[i] <= k < rel[i + 1]).
TRC_HD uint32_t w32_clz64(uint64_t x) {
#if defined(__CUDA_ARCH__)
  return (uint32_t)__clzll((long long)x);
#else
  return x ? (uint32_t)__builtin_clzll(x) : 64u;
#endif
}
// 32 bits of the 64-bit value hi:lo starting at bit s (0 <= s < 32)
TRC_HD uint32_t w32_fsr(uint32_t lo, uint32_t hi, int s) {
#if defined(__CUDA_ARCH__)
  return __funnelshift_r(lo, hi, (unsigned)s);
#else
  return s ? ((lo >> s) | (hi << (32 - s))) : lo;
#endif
}
TRC_HD int w32_clz32(uint32_t x) {
#if defined(__CUDA_ARCH__)
  return __clz((int)x);
#else
  return x ? __builtin_clz(x) : 32;
#endif
}
// Returns the candidates that need the generic path (runs of 24 bytes or more, separators of 8 bytes or more, a
// segment boundary inside the candidate run, a run that reaches the end of the tile).  Per candidate byte k two kinds
// of 32-bit windows are cut from the chunk's mask words and its neighbours': "down" windows, bit 31 <-> byte k-1 and
// bit j <-> byte k-32+j, and "up" windows, bit 0 <-> byte k.
TRC_HD uint32_t w32_dup_chunk_simple(const uint8_t* in, const W32& M, int32_t c, uint32_t valid, int32_t lim) {
  const uint32_t w = M.w[c], sp = M.sp[c], bd = M.bd[c];
  const uint32_t wp = M.w[c - 1], spp = M.sp[c - 1];
  const uint32_t wstart = w & ~((w << 1) | (wp >> 31));
  uint32_t cand = wstart & ((sp << 1) | (spp >> 31)) & ~bd & valid;
  if (!cand) return 0u;
  const uint32_t wn = M.w[c + 1], bdn = M.bd[c + 1], bdp = M.bd[c - 1], ct = M.ct[c], ctp = M.ct[c - 1];
  uint32_t generic = 0;
  while (cand) {
    const int bit = (int)trc_ctz32(cand);
    cand &= cand - 1;
    const int32_t k = (c << 5) + bit;
    // the candidate run [k, k + L)
    const uint32_t wup = w32_fsr(w, wn, bit);
    if ((wup & 0xffffffu) == 0xffffffu) { generic |= 1u << bit; continue; }
    const int32_t L = (int32_t)trc_ctz32(~wup);
    if ((w32_fsr(bd, bdn, bit) & ((1u << L) - 1u) & ~1u) || k + L > lim) { generic |= 1u << bit; continue; }   // a segment starts inside the run
    // the whitespace run [p, k) in front of it
    const uint32_t ds = w32_fsr(spp, sp, bit);
    const int lws = w32_clz32(~ds);                                                  // whitespace bytes from k-1 downwards (>= 1)
    if (lws > 7) { generic |= 1u << bit; continue; }
    const uint32_t db = w32_fsr(bdp, bd, bit);
    if (db >> (32 - lws)) continue;                                                  // the run starts its segment (or spans a boundary): no previous word
    if (in[k + L - 1] != in[k - lws - 1]) continue;                                  // equal words end with the same byte: rejects most
    const uint32_t dw = w32_fsr(wp, w, bit);
    // the previous run ends at p - 1: it must be exactly L bytes long inside the same segment
    if (w32_clz32(~(dw << lws)) < L) continue;                                       // (a run of zero bytes when p-1 is not \w)
    const int pb = 32 - lws - L;                                                     // window position of P = p - L  (>= 32 - 7 - 23 = 2)
    if ((db >> (pb + 1)) & ((1u << (L - 1)) - 1u)) continue;                         // a segment starts inside (P, p-1]
    if (((dw >> (pb - 1)) & 1u) && !((db >> pb) & 1u)) continue;                     // the previous run is longer
    bool eq = true;
    for (int32_t i = 0; i + 1 < L; ++i) if (in[k - lws - L + i] != in[k + i]) { eq = false; break; }
    if (!eq) continue;
    const uint32_t run = (1u << L) - 1u;                                             // L < 24
    w32_or(M.drop, c, run << bit);
    if (bit + L > 32) w32_or(M.drop, c + 1, run >> (32 - bit));
    // the emitted space would come from the last whitespace character of the run: its lead byte
    const int nct = w32_clz32(~w32_fsr(ctp, ct, bit));                               // continuation bytes from k-1 downwards (< lws)
    const int qb = bit - 1 - nct;                                                    // relative to this chunk
    if (qb >= 0) w32_or(M.ds, c, 1u << qb); else w32_or(M.ds, c - 1, 1u << (32 + qb));
  }
  return generic;
}
template <class SegOf>
TRC_HD void w32_dup_chunk(const uint8_t* in, const W32& M, int32_t c, uint32_t valid, int32_t lim, const int32_t* rel, SegOf&& seg_of) {
  uint32_t generic = w32_dup_chunk_simple(in, M, c, valid, lim);
  while (generic) {
    const int bit = (int)trc_ctz32(generic);
    generic &= generic - 1;
    const int32_t k = (c << 5) + bit;
    if (w32_dup_quick_reject(in, M, k, lim)) continue;
    const int i = seg_of(k);
    w32_mark_dup(in, M, k, rel[i], rel[i + 1]);
  }
}

struct W32Codes { uint32_t copy, space, ins; };   // per byte: copy it / write ' ' instead of it / write ' ' before it

// Output codes of chunk c.  [lo, hi) = bytes of the tile that belong to its segments.
TRC_HD W32Codes w32_chunk_codes(const uint8_t* in, const W32& M, int32_t c, int32_t lo, int32_t hi) {
  const int32_t k0 = c << 5;
  const uint32_t valid = w32_valid(c, lo, hi);
  const uint32_t sp = M.sp[c] & valid, ct = M.ct[c], bd = M.bd[c];
  const uint32_t spn = M.sp[c + 1], bdn = M.bd[c + 1], spp = M.sp[c - 1], epp = M.ep[c - 1];
  const uint32_t nvalid = (valid >> 1) | ((k0 + 32 < hi) ? 0x80000000u : 0u);
  const uint32_t nxt_sp = (sp >> 1) | (spn << 31), nxt_bd = (bd >> 1) | (bdn << 31);
  // one-byte whitespace characters that end their run and are followed by a byte of the same segment
  uint32_t last_ws = sp & ~ct & ~nxt_sp & ~nxt_bd & nvalid;
  // multi-byte whitespace characters are decided at their lead byte (their continuation bytes carry sp | ct)
  {
    uint32_t t = sp & ~ct & nxt_sp;
    while (t) {
      const int bit = (int)trc_ctz32(t);
      t &= t - 1;
      const int32_t k = k0 + bit;
      const uint8_t ch = in[k];
      if (ch < 0xC0) continue;                               // ASCII whitespace followed by whitespace: not last
      const int32_t kn = k + trc_u8len(ch);
      if (kn < hi && !w32_bit(M.bd, kn) && !w32_bit(M.sp, kn)) last_ws |= 1u << bit;
    }
  }
  // leading whitespace: runs that start at a segment start
  uint32_t seeds = sp & bd & ~(M.nolead ? M.nolead[c] : 0u);
  if ((sp & 1u) && (spp >> 31) && !(bd & 1u)) {              // a run enters from the previous chunk: is it leading?
    int32_t p = k0;
    while (p > lo && w32_bit(M.sp, p - 1) && !w32_bit(M.bd, p)) --p;
    if (w32_bit(M.sp, p) && w32_bit(M.bd, p) && !(M.nolead && w32_bit(M.nolead, p))) seeds |= 1u;
  }
  uint32_t lead_ws = 0;
  while (seeds) {                                            // one flood per seed: adjacent seeds must not cancel
    const uint32_t low = seeds & (0u - seeds);
    const uint32_t f = ((sp + low) ^ sp) & sp;               // the run of ones of sp that starts at `low`
    lead_ws |= f;
    seeds &= ~f;
  }
  const uint32_t ep_prev = (M.ep[c] << 1) | (epp >> 31);
  W32Codes r;
  r.copy = ~sp & ~M.drop[c] & valid;
  r.space = last_ws & ~lead_ws & ~M.ds[c] & valid;
  r.ins = r.copy & M.lt[c] & ep_prev & ~bd;
  return r;
}
