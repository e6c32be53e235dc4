// clean_text on 8-byte cells (SWAR): the same local rules as trc_clean_par.cuh
// (/root/reference/preprocessor.py:38,69-89), but one lane decides 8 consecutive bytes at once from bit
// masks (bit i of a mask <-> byte i of the cell):
//   sp  Python-whitespace bytes (all bytes of a multi-byte whitespace character)
//   w   \w bytes (all bytes of a multi-byte \w character)
//   ct  UTF-8 continuation bytes          ep  '.', '!' or '?'          lt  ASCII letters
//   bd  first byte of a segment           drop / ds  set by the duplicate-word pass
// ASCII bytes are classified arithmetically; multi-byte characters are fixed up by the lane that owns
// their lead byte (code-point lookup).  Masks of the neighbouring cells give the one byte of context the
// rules need; whitespace runs that cross a cell edge take a short scalar walk.
#pragma once
#include "trc_common.cuh"

#define SW_H 0x8080808080808080ULL
#define SW_L 0x0101010101010101ULL

// bit 7 of byte i set <=> (byte i & 0x7f) >= n      (1 <= n <= 0x80)
TRC_HD uint64_t sw_ge(uint64_t xa_or_h, uint32_t n) { return (xa_or_h - SW_L * n) & SW_H; }
TRC_HD uint32_t sw_pack(uint64_t hbits) { return (uint32_t)(((hbits >> 7) * 0x0102040810204080ULL) >> 56); }

struct SwCell { uint32_t sp, w, ct, lead, lt, ep; };   // 8-bit masks (ASCII part only for sp / w)

TRC_HD SwCell sw_classify(uint64_t x) {
  const uint64_t na = x & SW_H;
  const uint64_t t = (x & ~SW_H) | SW_H;
  const uint64_t g09 = sw_ge(t, 0x09), g0e = sw_ge(t, 0x0e), g1c = sw_ge(t, 0x1c), g21 = sw_ge(t, 0x21), g22 = sw_ge(t, 0x22);
  const uint64_t g2e = sw_ge(t, 0x2e), g2f = sw_ge(t, 0x2f), g30 = sw_ge(t, 0x30), g3a = sw_ge(t, 0x3a), g3f = sw_ge(t, 0x3f);
  const uint64_t g40 = sw_ge(t, 0x40), g41 = sw_ge(t, 0x41), g5b = sw_ge(t, 0x5b), g5f = sw_ge(t, 0x5f), g60 = sw_ge(t, 0x60);
  const uint64_t g61 = sw_ge(t, 0x61), g7b = sw_ge(t, 0x7b);
  const uint64_t sp = ((g09 & ~g0e) | (g1c & ~g21)) & ~na;
  const uint64_t lt = ((g41 & ~g5b) | (g61 & ~g7b)) & ~na;
  const uint64_t w = (lt | (g30 & ~g3a) | (g5f & ~g60)) & ~na;
  const uint64_t ep = ((g2e & ~g2f) | (g21 & ~g22) | (g3f & ~g40)) & ~na;
  const uint64_t lead = na & ((x << 1) & SW_H);
  SwCell c;
  c.sp = sw_pack(sp); c.w = sw_pack(w); c.lt = sw_pack(lt); c.ep = sw_pack(ep);
  c.lead = sw_pack(lead); c.ct = sw_pack(na & ~lead);
  return c;
}

// Flood `seed` upwards through runs of ones in `mask` (both 8-bit in a 32-bit word).
TRC_HD uint32_t sw_flood_up(uint32_t seed, uint32_t mask) { seed &= mask; return ((mask + seed) ^ mask) & mask; }

struct SwCodes { uint32_t copy, space, ins; };   // per byte: copy it / write ' ' instead / write ' ' before it

// Output codes of one cell.  All arguments are 8-bit masks of this cell unless suffixed _p (previous cell)
// or _n (next cell); `valid` marks the bytes that belong to the tile's own segments.
// lead_ws: whitespace bytes of this cell known to be LEADING (run starts at a segment start), including
// runs that entered from the previous cell - the caller resolves cross-cell runs and passes them here.
// last_ws: lead bytes of whitespace characters that are the last character of their run and are followed
// by a byte of the same segment - likewise resolved by the caller for the cross-cell cases.
TRC_HD SwCodes sw_codes(uint32_t sp, uint32_t ct, uint32_t lt, uint32_t bd, uint32_t drop, uint32_t ds, uint32_t valid,
                        uint32_t ep_prev_byte /* mask: previous byte is .!? */, uint32_t lead_ws, uint32_t last_ws) {
  SwCodes r;
  r.copy = ~sp & ~drop & valid & 0xffu;
  r.space = last_ws & ~lead_ws & ~ds & valid & 0xffu;
  r.ins = r.copy & lt & ep_prev_byte & ~bd & 0xffu;
  (void)ct;
  return r;
}

// ---- mask arrays of a tile (one byte per 8-byte cell) ------------------------------------------------------
struct SwMasks {
  uint8_t* sp; uint8_t* w; uint8_t* ct; uint8_t* lt; uint8_t* ep; uint8_t* bd; uint8_t* drop; uint8_t* ds;
  int32_t ncell;
};

TRC_HD uint32_t sw_bit(const uint8_t* arr, int32_t k) { return (arr[k >> 3] >> (k & 7)) & 1u; }

TRC_HD void sw_or_bits(uint8_t* arr, int32_t cell, uint32_t bits) {
#if defined(__CUDA_ARCH__)
  unsigned int* wptr = reinterpret_cast<unsigned int*>(arr + (cell & ~3));
  atomicOr(wptr, bits << (8 * (cell & 3)));
#else
  arr[cell] |= (uint8_t)bits;
#endif
}
TRC_HD void sw_or_bit(uint8_t* arr, int32_t k) { sw_or_bits(arr, k >> 3, 1u << (k & 7)); }

// Multi-byte characters: the lane that owns the lead byte looks the code point up and sets sp / w on every
// byte of the character (they may spill into the next cell).  `in` is tile-relative, hi = end of the tile.
TRC_HD void sw_fix_multibyte(const TrcTables& T, const uint8_t* in, const SwMasks& M, int32_t cell, uint32_t lead, int32_t hi) {
  while (lead) {
    const int bit = (int)trc_ctz32(lead);
    lead &= lead - 1;
    int32_t k = cell * 8 + bit;
    int len;
    uint32_t cp = trc_decode(in, k, hi, &len);
    uint32_t pr = T.props[cp];
    if (pr & (TRC_P_SPACE | TRC_P_WORD)) {
      uint8_t* dst = (pr & TRC_P_SPACE) ? M.sp : M.w;
      for (int q = 0; q < len; ++q) sw_or_bit(dst, k + q);
    }
  }
}

// Duplicate-word candidates of a cell: first byte of a \w-run that directly follows whitespace and is not a
// segment start.
TRC_HD uint32_t sw_dup_candidates(const SwMasks& M, int32_t c, uint32_t valid) {
  uint32_t w = M.w[c], sp = M.sp[c];
  uint32_t wp = c > 0 ? M.w[c - 1] : 0u, spp = c > 0 ? M.sp[c - 1] : 0u;
  uint32_t wstart = w & ~((w << 1) | (wp >> 7));
  uint32_t after_ws = (sp << 1) | (spp >> 7);
  return wstart & after_ws & ~(uint32_t)M.bd[c] & valid & 0xffu;
}

// Cheap necessary condition for sw_mark_dup at candidate byte k (a \w-run start after whitespace): when the
// separator is a single whitespace byte and the run ends within 16 bytes inside its segment, equal words
// must end with the same byte.  Returns true when the candidate can be discarded.
TRC_HD bool sw_dup_quick_reject(const uint8_t* in, const SwMasks& M, int32_t k, int32_t lim) {
  const int32_t c = k >> 3, bit = k & 7;
  if (k < 2 || sw_bit(M.sp, k - 2) || sw_bit(M.ct, k - 1)) return false;        // longer / multi-byte separator: undecided
  const uint32_t wwin = ((uint32_t)M.w[c] | ((uint32_t)M.w[c + 1] << 8) | ((uint32_t)M.w[c + 2] << 16)) >> bit;
  const uint32_t L = trc_ctz32(~wwin);
  if (L >= 16 || k + (int32_t)L > lim) return false;                            // bytes past lim belong to the next tile
  const uint32_t bwin = ((uint32_t)M.bd[c] | ((uint32_t)M.bd[c + 1] << 8) | ((uint32_t)M.bd[c + 2] << 16)) >> bit;
  if (bwin & ((1u << L) - 1u) & ~1u) return false;                             // a segment starts inside the run
  return in[k + L - 1] != in[k - 2];
}

// Number of consecutive set bits of a mask array starting at bit k (a byte of the array at a time).
TRC_HD int32_t sw_run_up(const uint8_t* arr, int32_t k, int32_t limit) {
  int32_t e = k;
  while (e < limit) {
    const uint32_t x = (uint32_t)arr[e >> 3] >> (e & 7);
    const int room = 8 - (e & 7);
    const uint32_t inv = ~x & ((1u << room) - 1u);
    if (inv) { e += (int32_t)trc_ctz32(inv); break; }
    e += room;
  }
  return e < limit ? e : limit;
}

// The rule of cp_mark_dup on mask arrays: byte k starts a \w-run after whitespace inside segment [a,b).
TRC_HD void sw_mark_dup(const uint8_t* in, const SwMasks& M, int32_t k, int32_t a, int32_t b) {
  int32_t p = k;
  while (p > a && sw_bit(M.sp, p - 1)) --p;                 // start of the whitespace run
  if (p == a || !sw_bit(M.w, p - 1)) return;                // ... which must follow a \w character
  const int32_t e = sw_run_up(M.w, k, b);
  if (in[e - 1] != in[p - 1]) return;                       // equal words end with the same byte: rejects most
  int32_t L = e - k;
  if (p - L < a) return;
  if (p - L > a && sw_bit(M.w, p - L - 1)) return;          // previous \w-run is longer
  for (int32_t i = 0; i < L; ++i) if (in[p - L + i] != in[k + i]) return;
  for (int32_t i = k; i < e; ++i) sw_or_bit(M.drop, i);
  int32_t q = k - 1;
  while (sw_bit(M.ct, q)) --q;
  sw_or_bit(M.ds, q);
}

// Output codes of cell c.  [lo,hi) = bytes of the tile that belong to its segments.
TRC_HD SwCodes sw_cell_codes(const uint8_t* in, const SwMasks& M, int32_t c, int32_t lo, int32_t hi) {
  const int32_t k0 = c * 8;
  uint32_t valid = 0xffu;
  if (k0 < lo) valid &= (lo - k0 >= 8) ? 0u : (0xffu << (lo - k0));
  if (k0 + 8 > hi) valid &= (hi <= k0) ? 0u : (0xffu >> (k0 + 8 - hi));
  valid &= 0xffu;
  const uint32_t sp = M.sp[c] & valid, ct = M.ct[c], bd = M.bd[c];
  const bool hasn = c + 1 < M.ncell, hasp = c > 0;
  const uint32_t spn = hasn ? M.sp[c + 1] : 0u, bdn = hasn ? M.bd[c + 1] : 0u, spp = hasp ? M.sp[c - 1] : 0u;
  const uint32_t epp = hasp ? M.ep[c - 1] : 0u;
  const uint32_t nvalid = (valid >> 1) | ((k0 + 8 < hi) ? 0x80u : 0u);
  const uint32_t nxt_sp = (sp >> 1) | ((spn & 1u) << 7), nxt_bd = (bd >> 1) | ((bdn & 1u) << 7);
  // single-byte whitespace characters that end their run and are followed by a byte of the same segment
  uint32_t last_ws = sp & ~ct & ~nxt_sp & ~nxt_bd & nvalid;
  // multi-byte whitespace characters: decided at their lead byte (continuation bytes carry sp|ct)
  uint32_t mb = sp & ~ct & nxt_sp;          // candidates: a whitespace byte followed by another whitespace byte
  {
    uint32_t t = mb;
    while (t) {
      const int bit = (int)trc_ctz32(t);
      t &= t - 1;
      int32_t k = k0 + bit;
      uint8_t ch = in[k];
      if (ch < 0xC0) continue;                               // ASCII whitespace followed by whitespace: not last
      int32_t kn = k + trc_u8len(ch);
      if (kn < hi && !sw_bit(M.bd, kn) && !sw_bit(M.sp, kn)) last_ws |= 1u << bit;
    }
  }
  // leading whitespace: runs that start at a segment start
  uint32_t seeds = sp & bd;                                  // a segment that begins with whitespace
  if ((sp & 1u) && (spp >> 7) && !(bd & 1u)) {               // a run enters from the previous cell: is it leading?
    int32_t p = k0;
    while (p > lo && sw_bit(M.sp, p - 1) && !sw_bit(M.bd, p)) --p;
    if (sw_bit(M.sp, p) && sw_bit(M.bd, p)) seeds |= 1u;
  }
  uint32_t lead_ws = 0;
  while (seeds) {                                            // one flood per seed: adjacent seeds must not cancel
    const uint32_t low = seeds & (0u - seeds);
    const uint32_t f = sw_flood_up(low, sp);
    lead_ws |= f;
    seeds &= ~f;
  }
  const uint32_t ep_prev = (((uint32_t)M.ep[c] << 1) | (epp >> 7)) & 0xffu;
  return sw_codes(sp, ct, M.lt[c], bd, M.drop[c], M.ds[c], valid, ep_prev, lead_ws, last_ws);
}
