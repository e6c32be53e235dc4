// K1a + K2 fused, fourth generation: WARP-PRIVATE tiles.  (preprocessor.py:38,69-89 + big_chunkeroosky.py:83 via
// DESIGN.md §3; same results as k_clean_tok_fused3, which stays selectable with TRCHUNK_FUSED=3.)
//
// What changed against the third generation, and why (profiles/r1q_fused3_summary.txt: 7.1 warp instructions per raw
// byte, barrier stalls first, 62 % issue-active, 8-bit masks in 32-bit registers):
//  * a tile is F4_SEGS raw segments (about 2 KB) and belongs to ONE warp: its twelve phases are separated by
//    __syncwarp() instead of CTA-wide barriers, and no warp ever waits for another warp of its CTA;
//  * a lane owns a 32-byte CHUNK, so every mask operation decides 32 bytes (trc_w32.cuh) instead of 8, and the
//    per-chunk prefixes are one warp shuffle scan per KB;
//  * bytes are classified through a 256-entry table in shared memory (two multiply-adds per byte give eight class
//    planes) instead of seventeen SWAR range tests per 8 bytes;
//  * the raw tile arrives by a bulk asynchronous copy (cp.async.bulk global -> shared, completion on an mbarrier,
//    issued by lane 0 for the warp's NEXT tile as soon as the current tile's raw bytes are dead), so no lane spends
//    issue slots on loads and the copy overlaps the token counting of the current tile.  Source and destination are
//    this GPU's own memory; no multicast, no cluster;
//  * buffers are sized for the worst case of a tile that fits (cleaned <= 1.5 x raw, words <= raw / 2 + segments), so
//    the only other path is the slow lane for tiles that do not fit (sequential per-segment routines; segments of
//    more than 4 KB arrive pre-cleaned from trc_giant.cuh as before).
#pragma once

#define F4_SEGS 32
#define F4_IN_CAP 2304
#define F4_NCH (F4_IN_CAP / 32)
#define F4_MAXSTEP ((F4_NCH + 31) / 32)
#define F4_OUT_CAP (F4_IN_CAP + F4_IN_CAP / 2)
#define F4_WORD_CAP (F4_IN_CAP / 2 + F4_SEGS + 8)
#define F4_SLOW_CAP 248
#define F4_WARPS 16
#define F4_T (F4_WARPS * 32)
#define F4_NM 14
#define F4_FULL 0xffffffffu

// m[] rows
#define F4_SP 0
#define F4_W 1
#define F4_CT 2
#define F4_BD 3
#define F4_EP 4
#define F4_LT 5
#define F4_DROP 6
#define F4_DS 7
#define F4_COPY 8
#define F4_SPACE 9
#define F4_INS 10
#define F4_PRE_OS 11     // exclusive prefix: output bytes | emitted spaces << 16
#define F4_PRE_TB 12     // exclusive prefix: copied continuation bytes | segment starts << 16
#define F4_NS 13         // whitespace bytes other than U+0020

struct F4Warp {
  alignas(16) uint8_t in[F4_IN_CAP + 32];
  alignas(16) uint8_t out[F4_OUT_CAP + 32];
  union alignas(16) {
    uint32_t m[F4_NM][F4_NCH + 2];                                 // mask rows, chunk c at [c + 1]
    struct { uint8_t cnt[F4_WORD_CAP + 8]; uint16_t miss[F4_WORD_CAP + 8]; uint16_t slow[F4_SLOW_CAP]; } t;   // after the scatter
  } u;
  uint16_t wstart[F4_WORD_CAP + 8];                                // position in out | first word of its segment << 15
  int32_t rel[F4_SEGS + 1];
  int32_t xtok[F4_SEGS];
  int32_t spmin[F4_SEGS];
  uint16_t loff[F4_SEGS + 1];
  uint16_t wfirst[F4_SEGS + 1];
  uint8_t kr[F4_SEGS + 4];                                         // by rank among the non-empty raw segments: kept segments up to it
  alignas(8) unsigned long long mbar;
};
static_assert(sizeof(((F4Warp*)0)->u.t) <= sizeof(((F4Warp*)0)->u.m), "token lists fit the mask rows");

struct F4Smem {
  uint64_t lut[256];
  uint64_t hot8[TRC_HOT8];
  F4Warp w[F4_WARPS];
};
static_assert(sizeof(F4Smem) <= 227 * 1024, "one CTA per SM");

// ---- bulk asynchronous copy + mbarrier (PTX) ---------------------------------------------------------------------
__device__ __forceinline__ uint32_t f4_saddr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void f4_mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void f4_mbar_expect(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void f4_bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst), "l"(src), "r"(bytes),
               "r"(bar)
               : "memory");
}
__device__ __forceinline__ void f4_mbar_wait(uint32_t bar, uint32_t parity) {
  uint32_t done;
  do {
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                 : "=r"(done)
                 : "r"(bar), "r"(parity)
                 : "memory");
  } while (!done);
}
__device__ __forceinline__ void f4_fence_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

__device__ __forceinline__ int32_t f4_excl_scan(int32_t v, int lane, int32_t* total) {
  int32_t x = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) { const int32_t y = __shfl_up_sync(F4_FULL, x, o); if (lane >= o) x += y; }
  *total = __shfl_sync(F4_FULL, x, 31);
  return x - v;
}


// ---- two-level look-back ------------------------------------------------------------------------------------------
// With a tile per warp there are ~2 400 tiles in flight and ~70 new ones per microsecond: a flat decoupled look-back
// has every tile walk back over thousands of aggregates (measured: 40 % of the kernel, L2 traffic of 19 x the text).
// So tiles are grouped by 32 and there is a second row of (status, value) words per GROUP.  A tile publishes its two
// aggregates (cleaned bytes, kept segments) as soon as it knows them.  To resolve its offsets it sums the aggregates
// of the tiles before it in its own group (one 32-wide round) and walks back over the group words; a group word that
// nobody has written yet is computed on the spot from the group's 32 tile aggregates - all of them belong to tiles
// that were claimed earlier and publish without waiting for anyone - and left behind for the next reader, as is the
// inclusive prefix of the previous group.  No tile ever waits for another tile's look-back.
#define F4_VAL ((1ULL << 62) - 1)
__device__ __forceinline__ void f4_publish(const FuseArgs& A, long long tile, long long agg_b, long long agg_k, int lane) {
  if (lane == 0) {
    st_relaxed(&A.desc_bytes[tile], (1ULL << 62) | (unsigned long long)agg_b);
    st_relaxed(&A.desc_kept[tile], (1ULL << 62) | (unsigned long long)agg_k);
  }
}
__device__ __forceinline__ unsigned long long f4_spin(const unsigned long long* p) {
  unsigned long long v = ld_relaxed(p);
  while ((v >> 62) == 0) v = ld_relaxed(p);
  return v;
}
__device__ __noinline__ void f4_resolve(const FuseArgs& A, long long tile, int64_t ntiles, int lane, long long* gbase, long long* kbase) {
  const long long g = tile >> 5;
  const int r = (int)(tile & 31);
  long long vb = 0, vk = 0;
  if (lane < r) {
    vb = (long long)(f4_spin(&A.desc_bytes[(g << 5) + lane]) & F4_VAL);
    vk = (long long)(f4_spin(&A.desc_kept[(g << 5) + lane]) & F4_VAL);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) { vb += __shfl_xor_sync(F4_FULL, vb, o); vk += __shfl_xor_sync(F4_FULL, vk, o); }
  long long eb = 0, ek = 0;                              // exclusive prefix of group g
  bool done_b = false, done_k = false;                   // the two rows are written independently: each has its own walk
  long long idx = g - 1;
  while (idx >= 0 && !(done_b && done_k)) {
    const long long my = idx - lane;
    unsigned long long db = (2ULL << 62), dk = (2ULL << 62);
    if (my >= 0) { db = ld_relaxed(&A.grp_bytes[my]); dk = ld_relaxed(&A.grp_kept[my]); }
    const unsigned sb = (unsigned)(db >> 62), sk = (unsigned)(dk >> 62);
    const unsigned b2 = __ballot_sync(F4_FULL, sb == 2), b0 = __ballot_sync(F4_FULL, sb == 0);
    const unsigned k2 = __ballot_sync(F4_FULL, sk == 2), k0 = __ballot_sync(F4_FULL, sk == 0);
    const int fb = b2 ? __ffs(b2) - 1 : 32, fk = k2 ? __ffs(k2) - 1 : 32;
    unsigned need0 = 0;
    if (!done_b) need0 |= b0 & (fb >= 31 ? 0xffffffffu : ((2u << fb) - 1u));
    if (!done_k) need0 |= k0 & (fk >= 31 ? 0xffffffffu : ((2u << fk) - 1u));
    if (need0) {                                         // the nearest group nobody has summed yet: do it here
      const long long h = idx - (__ffs(need0) - 1);
      const long long gsz = (ntiles - (h << 5)) < 32 ? (ntiles - (h << 5)) : 32;
      long long sb2 = 0, sk2 = 0;
      if (lane < gsz) { sb2 = (long long)(f4_spin(&A.desc_bytes[(h << 5) + lane]) & F4_VAL); sk2 = (long long)(f4_spin(&A.desc_kept[(h << 5) + lane]) & F4_VAL); }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) { sb2 += __shfl_xor_sync(F4_FULL, sb2, o); sk2 += __shfl_xor_sync(F4_FULL, sk2, o); }
      if (lane == 0) {                                   // only over an empty word: a later status never goes back
        if (ld_relaxed(&A.grp_bytes[h]) >> 62 == 0) atomicCAS(&A.grp_bytes[h], 0ULL, ((h == 0 ? 2ULL : 1ULL) << 62) | (unsigned long long)sb2);
        if (ld_relaxed(&A.grp_kept[h]) >> 62 == 0) atomicCAS(&A.grp_kept[h], 0ULL, ((h == 0 ? 2ULL : 1ULL) << 62) | (unsigned long long)sk2);
      }
      __syncwarp();
      continue;                                          // read the round again
    }
    if (!done_b) {
      long long cb = lane <= fb ? (long long)(db & F4_VAL) : 0;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) cb += __shfl_xor_sync(F4_FULL, cb, o);
      eb += cb; done_b = b2 != 0;
    }
    if (!done_k) {
      long long ck = lane <= fk ? (long long)(dk & F4_VAL) : 0;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) ck += __shfl_xor_sync(F4_FULL, ck, o);
      ek += ck; done_k = k2 != 0;
    }
    idx -= 32;
  }
  if (g > 0 && lane == 0) {                              // inclusive prefix of group g - 1, for the next reader
    st_relaxed(&A.grp_bytes[g - 1], (2ULL << 62) | (unsigned long long)eb);
    st_relaxed(&A.grp_kept[g - 1], (2ULL << 62) | (unsigned long long)ek);
  }
  *gbase = vb + eb; *kbase = vk + ek;
}

// what a warp knows about a tile before it becomes current
struct F4Tile {
  long long id;            // tile index, or >= ntiles
  int64_t so, so_end;      // seg_off[s0 + lane] (lanes <= ns), seg_off[s0 + ns]
  int64_t raw0;            // seg_off[s0]
  int ns;
  bool fast;
};

__device__ __forceinline__ void f4_describe(F4Tile& t, const TrcDev& D, int64_t ntiles, int tsegs, int lane) {
  t.ns = 0; t.fast = false; t.so = t.so_end = t.raw0 = 0;
  if (t.id >= ntiles) return;
  const int64_t s0 = t.id * tsegs;
  t.ns = (int)((D.n - s0) < tsegs ? (D.n - s0) : tsegs);
  t.so = D.seg_off[s0 + (lane <= t.ns ? lane : t.ns)];
  t.so_end = D.seg_off[s0 + t.ns];
  t.raw0 = __shfl_sync(F4_FULL, t.so, 0);
  t.fast = (t.so_end - (t.raw0 & ~(int64_t)15)) <= F4_IN_CAP;
}

// lane 0: start the bulk copy of a fast tile's raw bytes (whole 16-byte units inside the text buffer)
__device__ __forceinline__ uint32_t f4_issue(const F4Tile& t, const TrcDev& D, const FuseArgs& A, F4Warp& W, int lane) {
  const int64_t base = t.raw0 & ~(int64_t)15;
  const int64_t span = t.so_end - base;
  int64_t n16 = (span + 15) & ~(int64_t)15;
  const int64_t room = (A.text_bytes - base) & ~(int64_t)15;
  if (n16 > room) n16 = room;
  if (n16 <= 0) return 0u;
  if (lane == 0) {
    f4_fence_async();
    f4_mbar_expect(f4_saddr(&W.mbar), (uint32_t)n16);
    f4_bulk_g2s(f4_saddr(W.in), D.text + base, (uint32_t)n16, f4_saddr(&W.mbar));
  }
  return (uint32_t)n16;
}

// 32 contiguous output bytes (eight little-endian words) to an arbitrary shared-memory address
__device__ __forceinline__ void f4_store32(uint8_t* dst, const uint32_t x[8]) {
  const unsigned d = (unsigned)(f4_saddr(dst) & 3u);
  if (d == 0) {
    uint32_t* p = reinterpret_cast<uint32_t*>(dst);
#pragma unroll
    for (int j = 0; j < 8; ++j) p[j] = x[j];
    return;
  }
  uint32_t* p = reinterpret_cast<uint32_t*>(dst - d);                 // aligned word that holds the first byte
  const unsigned sh = 8u * (4u - d);
#pragma unroll
  for (int j = 1; j < 8; ++j) p[j] = __funnelshift_r(x[j - 1], x[j], sh);
  // head: bytes 0 .. 3-d of x[0] ; tail: bytes 4-d .. 3 of x[7]
  dst[0] = (uint8_t)x[0];
  if (d < 3) dst[1] = (uint8_t)(x[0] >> 8);
  if (d < 2) dst[2] = (uint8_t)(x[0] >> 16);
  uint8_t* tl = dst + 32 - d;
  tl[0] = (uint8_t)(x[7] >> (8 * (4 - d)));
  if (d > 1) tl[1] = (uint8_t)(x[7] >> (8 * (5 - d)));
  if (d > 2) tl[2] = (uint8_t)(x[7] >> 24);
}


// ---- out-of-line parts.  The warps of a CTA run independently, each somewhere else in the kernel, so what matters
// for the instruction cache is the size of the code the common tile touches: rare paths are separate functions and
// the per-step loops are not unrolled (the first version, fully inlined and unrolled, was 300 KB of SASS and spent
// 55 % of its issue slots waiting for instructions).
struct F4Seg { int32_t mylen, mykeep, mychars, loff, koff, mytok, mysp, tot_out, tot_kept; long long gbase, kbase; };

// generic duplicate-word candidates (separator of several bytes, long runs, segment edges)
__device__ __noinline__ void f4_dup_generic(const uint8_t* in, const W32& M, int32_t c, uint32_t generic, int32_t lim, const int32_t* rel, int ns) {
  while (generic) {
    const int bit = __ffs(generic) - 1;
    generic &= generic - 1;
    const int32_t k = (c << 5) + bit;
    if (w32_dup_quick_reject(in, M, k, lim)) continue;
    int lo = 0, hi = ns;
    while (hi - lo > 1) { const int mid = (lo + hi) >> 1; if (rel[mid] <= k) lo = mid; else hi = mid; }
    w32_mark_dup(in, M, k, rel[lo], rel[lo + 1]);
  }
}
__device__ __noinline__ void f4_fix_multibyte(const TrcTables& T, const uint8_t* in, const W32& M, int32_t c, uint32_t lead, int32_t hi) {
  w32_fix_multibyte(T, in, M, c, lead, hi);
}
// special-token literals in the cleaned tile (only when some byte of the tile can start one)
__device__ __noinline__ void f4_special_scan(const TrcTables& T, F4Warp& W, int ns, int32_t tot_out, int lane) {
  for (int32_t x = lane; x < tot_out; x += 32) {
    const uint8_t b = W.out[x];
    if ((T.sp_first[b >> 6] >> (b & 63)) & 1ULL) {
      int lo = 0, hi = ns;
      while (hi - lo > 1) { const int mid = (lo + hi) >> 1; if ((int32_t)W.loff[mid] <= x) lo = mid; else hi = mid; }
      const int32_t end = (int32_t)W.loff[lo + 1];
      for (uint32_t q = 0; q < T.n_sp; ++q) {
        const uint32_t o = T.sp_off[q], L = T.sp_off[q + 1] - o;
        if (L == 0 || T.sp_bytes[o] != b || x + (int32_t)L > end) continue;
        bool eq = true;
        for (uint32_t j = 1; j < L; ++j) if (W.out[x + j] != T.sp_bytes[o + j]) { eq = false; break; }
        if (eq) { atomicMin(&W.spmin[lo], (x << 4) | (int32_t)q); break; }
      }
    }
  }
}
// packed key of word w of the cleaned tile: up to 16 bytes, the segment's first word with its virtual space
__device__ __forceinline__ void f4_word_key(const F4Warp& W, int32_t w, uint64_t& lo, uint64_t& hi, int& Lw, int& vs, int32_t& st, int32_t& n) {
  const uint32_t e = W.wstart[w];
  st = (int32_t)(e & 0x7fffu); vs = (int)(e >> 15);
  n = (int32_t)(W.wstart[w + 1] & 0x7fffu) - st;
  Lw = n + vs;
  lo = hi = 0;
  if (Lw > 16) return;
  const uint32_t* p32 = reinterpret_cast<const uint32_t*>(W.out + (st & ~3));
  const unsigned sh = (unsigned)(st & 3) * 8u;
  const uint32_t w0 = p32[0], w1 = p32[1], w2 = p32[2], w3 = p32[3], w4 = p32[4];
  uint32_t k0 = __funnelshift_r(w0, w1, sh), k1 = __funnelshift_r(w1, w2, sh), k2 = __funnelshift_r(w2, w3, sh), k3 = __funnelshift_r(w3, w4, sh);
  if (vs) { k3 = (k3 << 8) | (k2 >> 24); k2 = (k2 << 8) | (k1 >> 24); k1 = (k1 << 8) | (k0 >> 24); k0 = (k0 << 8) | (uint32_t)' '; }
  const int r0 = Lw, r1 = Lw - 4, r2 = Lw - 8, r3 = Lw - 12;
  k0 &= r0 >= 4 ? 0xffffffffu : (r0 <= 0 ? 0u : ((1u << (8 * r0)) - 1u));
  k1 &= r1 >= 4 ? 0xffffffffu : (r1 <= 0 ? 0u : ((1u << (8 * r1)) - 1u));
  k2 &= r2 >= 4 ? 0xffffffffu : (r2 <= 0 ? 0u : ((1u << (8 * r2)) - 1u));
  k3 &= r3 >= 4 ? 0xffffffffu : (r3 <= 0 ? 0u : ((1u << (8 * r3)) - 1u));
  lo = (uint64_t)k0 | ((uint64_t)k1 << 32);
  hi = (uint64_t)k2 | ((uint64_t)k3 << 32);
}
// first sightings: the packed-word matcher (regex alternatives 1-4 on bit masks, exact-key vocabulary sets); what it
// cannot decide joins the tile's slow list.  Returns the new length of that list.
__device__ __noinline__ int32_t f4_match_words(const TrcTables& T, const FuseArgs& A, uint64_t* hot8, F4Warp& W, int32_t nmiss, int32_t nslow, int lane) {
  TrcFast F; F.acls = T.acls; F.hot8 = nullptr; F.hot8_mask = 0;
  uint8_t* const cnt = W.u.t.cnt; uint16_t* const miss = W.u.t.miss; uint16_t* const slow = W.u.t.slow;
  const uint32_t lt_mask = (1u << lane) - 1u;
  for (int32_t qb = 0; qb < nmiss; qb += 32) {
    const int32_t q = qb + lane;
    int res = -2;                                  // -2 nothing, -1 slow, >= 0 decided
    int32_t w = 0;
    if (q < nmiss) {
      w = miss[q];
      uint64_t lo, hi; int Lw, vs; int32_t st, n;
      f4_word_key(W, w, lo, hi, Lw, vs, st, n);
      uint32_t c = 0;
      bool done;
      if (Lw <= 16) {
        done = trc_count_word16(T, F, lo, hi, Lw, &c);
        if (Lw <= 7) hot8[trc_key8_slot(lo, TRC_HOT8 - 1)] = lo | ((uint64_t)(done ? c : (uint32_t)F3_MEMO_SLOW) << 56);
        if (Lw <= 15) f3_memo_put(A, lo, hi, done ? c : (uint32_t)F3_MEMO_SLOW);
      } else done = trc_count_word_fast(T, F, W.out + st, n, vs, &c) && c < 250u;
      if (done) { cnt[w] = (uint8_t)c; res = (int)c; } else res = -1;
    }
    const uint32_t bs = __ballot_sync(F4_FULL, res == -1);
    if (res == -1) { const int qq = nslow + __popc(bs & lt_mask); if (qq < F4_SLOW_CAP) slow[qq] = (uint16_t)w; else cnt[w] = 255; }
    nslow += __popc(bs);
  }
  __syncwarp();
  return nslow;
}
// words for the generic matcher go to k_count_slow_words (full occupancy); what does not fit its list, or this tile's
// own short list (cnt == 255), is counted here
__device__ __noinline__ void f4_slow_words(const TrcTables& T, const uint32_t* byte_id, const FuseArgs& A, F4Warp& W, int ns, int32_t nslow,
                                           int32_t tot_words, int32_t koff, long long gbase, long long kbase, int lane) {
  TrcFast F; F.acls = T.acls; F.hot8 = nullptr; F.hot8_mask = 0;
  uint8_t* const cnt = W.u.t.cnt; uint16_t* const slow = W.u.t.slow;
  const int32_t nl = nslow < F4_SLOW_CAP ? nslow : F4_SLOW_CAP;
  long long slbase = -1;
  if (lane == 0 && nl > 0) {
    const unsigned long long b0 = atomicAdd(A.sl_cnt, (unsigned long long)nl);
    slbase = ((int64_t)b0 + nl <= A.sl_cap) ? (long long)b0 : ((int64_t)b0 < A.sl_cap ? -2LL - (long long)b0 : -1LL);
  }
  slbase = __shfl_sync(F4_FULL, slbase, 0);
  if (slbase < -1) {
    SlowWord e; e.pos = 0; e.len_vs = 0; e.unit = 0;
    for (int64_t q = -2 - slbase + lane; q < A.sl_cap; q += 32) A.sl[q] = e;
  }
  auto seg_of_word = [&](int32_t w) { int lo = 0, hi = ns; while (hi - lo > 1) { const int mid = (lo + hi) >> 1; if ((int32_t)W.wfirst[mid] <= w) lo = mid; else hi = mid; } return lo; };
  bool any_here = nslow > nl || slbase < 0;
  for (int32_t qb = 0; qb < nl; qb += 32) {
    const int32_t q = qb + lane;
    const bool have = q < nl;
    const int32_t w = have ? (int32_t)slow[q] : 0;
    const int seg = have ? seg_of_word(w) : 0;
    const int32_t ko = __shfl_sync(F4_FULL, koff, seg);          // lane i holds the kept rank of segment i
    if (!have) continue;
    if (slbase >= 0) {
      const uint32_t e = W.wstart[w];
      const int32_t st = (int32_t)(e & 0x7fffu); const int vs = (int)(e >> 15);
      const int32_t n = (int32_t)(W.wstart[w + 1] & 0x7fffu) - st;
      SlowWord sw; sw.pos = gbase + st; sw.len_vs = n | (vs << 31); sw.unit = (int32_t)(kbase + ko);
      A.sl[slbase + q] = sw;
    } else cnt[w] = 255;
  }
  __syncwarp();
  if (!any_here) return;
  for (int32_t wb = 0; wb < tot_words; wb += 32) {
    const int32_t w = wb + lane;
    const bool here = w < tot_words && cnt[w] == 255;
    if (!__any_sync(F4_FULL, here)) continue;
    int seg = 0;
    if (here) seg = seg_of_word(w);
    const int32_t ko = __shfl_sync(F4_FULL, koff, seg);
    if (here) {
      const uint32_t e = W.wstart[w];
      const int32_t st = (int32_t)(e & 0x7fffu); const int vs = (int)(e >> 15);
      const int32_t n = (int32_t)(W.wstart[w + 1] & 0x7fffu) - st;
      const int32_t sl = (int32_t)W.loff[seg];
      const int64_t j = kbase + ko;
      const uint32_t c = trc_count_unit(T, F, byte_id, W.out + st, n, vs, [&](int64_t ls, int64_t len) -> uint32_t {
        const unsigned long long slot = atomicAdd(&A.wl_cnt[0], 1ULL);
        const unsigned long long off = atomicAdd(&A.wl_cnt[1], (unsigned long long)len);
        if ((int64_t)slot < A.wl_cap) { LongEntry le; le.unit = j; le.lstart = ls + (vs ? 0 : (st - sl + 1)); le.len = len; le.scratch_off = (int64_t)off; A.wl[slot] = le; }
        return 0u;
      });
      cnt[w] = 0;
      atomicAdd(&W.xtok[seg], (int32_t)c);
    }
  }
  __syncwarp();
}
// slow lane: the tile does not fit the buffers.  Sequential per-segment routines on global memory, a lane per segment;
// segments larger than 4 KB come pre-cleaned (trc_giant.cuh) and are copied.
__device__ __noinline__ void f4_slow_lane(const TrcDev& D, const TrcTables& T, const uint32_t* byte_id, const FuseArgs& A, long long tile,
                                          int64_t ntiles, int64_t s0, int ns, int64_t so, int64_t raw1, int lane, F4Seg& R) {
  TrcFast F; F.acls = T.acls; F.hot8 = nullptr; F.hot8_mask = 0;
  const uint32_t lt_mask = (1u << lane) - 1u;
  int32_t gq = -1;
  const int64_t so_next = __shfl_down_sync(F4_FULL, so, 1);       // lane ns holds seg_off[s0 + ns]
  const int64_t a = so, b = lane == 31 ? raw1 : so_next;
  if (lane < ns) {
    gq = A.G.of_seg ? A.G.of_seg[s0 + lane] : -1;
    if (gq >= 0) { R.mylen = (int32_t)(A.G.toff[gq + 1] - A.G.toff[gq]); R.mychars = A.G.chars[gq]; R.mykeep = R.mylen > 0; }
    else {
      const TrcCleanOut r = trc_clean_segment<false>(T, D.text + a, b - a, nullptr);
      R.mylen = (int32_t)r.bytes; R.mychars = (int32_t)r.chars; R.mykeep = r.bytes > 0;
    }
  }
  R.loff = f4_excl_scan(R.mylen, lane, &R.tot_out);
  {
    const uint32_t kb = __ballot_sync(F4_FULL, R.mykeep != 0);
    R.koff = __popc(kb & lt_mask); R.tot_kept = __popc(kb);
  }
  f4_publish(A, tile, R.tot_out, R.tot_kept, lane);
  f4_resolve(A, tile, ntiles, lane, &R.gbase, &R.kbase);
  if (lane < ns && R.mylen > 0 && gq < 0) trc_clean_segment<true>(T, D.text + a, b - a, D.C + R.gbase + R.loff);
  for (int i = 0; i < ns; ++i) {          // giants: the whole warp copies their cleaned text
    const int32_t q = __shfl_sync(F4_FULL, gq, i);
    if (q < 0) continue;
    const int32_t gl = __shfl_sync(F4_FULL, R.mylen, i), go = __shfl_sync(F4_FULL, R.loff, i);
    const uint8_t* gsrc = A.G.text + A.G.toff[q];
    uint8_t* gdst = D.C + R.gbase + go;
    for (int32_t k = lane; k < gl; k += 32) gdst[k] = gsrc[k];
  }
  __syncwarp();
  if (lane < ns && R.mylen > 0 && gq >= 0) {
    R.mytok = (int32_t)A.G.tok[gq];
    const long long v = A.G.spw[gq];
    R.mysp = v == 0x7fffffffffffffffLL ? -1 : (int32_t)(v & 15);
  } else if (lane < ns && R.mylen > 0) {
    const int64_t j = R.kbase + R.koff;
    const uint8_t* p = D.C + R.gbase + R.loff;
    const int32_t len0 = R.mylen;
    const uint32_t c = trc_count_unit(T, F, byte_id, p, len0, 1, [&](int64_t ls, int64_t len) -> uint32_t {
      const unsigned long long slot = atomicAdd(&A.wl_cnt[0], 1ULL);
      const unsigned long long off = atomicAdd(&A.wl_cnt[1], (unsigned long long)len);
      if ((int64_t)slot < A.wl_cap) { LongEntry le; le.unit = j; le.lstart = ls; le.len = len; le.scratch_off = (int64_t)off; A.wl[slot] = le; }
      return 0u;
    });
    R.mytok = (int32_t)c;
    R.mysp = trc_find_special(T, p, len0);
  }
}

__global__ void __launch_bounds__(F4_T, 1) k_clean_tok_w32(TrcDev D, TrcTables T, const uint32_t* __restrict__ byte_id, FuseArgs A,
                                                           int64_t ntiles, int tsegs) {
  extern __shared__ __align__(16) unsigned char fuse_raw[];
  F4Smem& S = *reinterpret_cast<F4Smem*>(fuse_raw);
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  F4Warp& W = S.w[wid];
  for (int k = tid; k < 256; k += F4_T) S.lut[k] = w32_lut_entry(T, (uint32_t)k);
  for (int k = tid; k < TRC_HOT8; k += F4_T) S.hot8[k] = 0ULL;
  if (lane == 0) f4_mbar_init(f4_saddr(&W.mbar), 1u);
  f4_fence_async();
  __syncthreads();
  W32 M;
  M.sp = W.u.m[F4_SP] + 1; M.w = W.u.m[F4_W] + 1; M.ct = W.u.m[F4_CT] + 1; M.bd = W.u.m[F4_BD] + 1; M.ep = W.u.m[F4_EP] + 1;
  M.lt = W.u.m[F4_LT] + 1; M.drop = W.u.m[F4_DROP] + 1; M.ds = W.u.m[F4_DS] + 1;
  uint32_t* const c_cp = W.u.m[F4_COPY] + 1; uint32_t* const c_se = W.u.m[F4_SPACE] + 1; uint32_t* const c_in = W.u.m[F4_INS] + 1;
  uint32_t* const p_os = W.u.m[F4_PRE_OS] + 1; uint32_t* const p_tb = W.u.m[F4_PRE_TB] + 1; uint32_t* const m_ns = W.u.m[F4_NS] + 1;
  uint32_t parity = 0;
  const uint32_t lt_mask = (1u << lane) - 1u;
  long long ph_t = clock64();
#define F4_PH(i) do { if (A.phase && lane == 0) { const long long now_ = clock64(); atomicAdd(&A.phase[i], (unsigned long long)(now_ - ph_t)); ph_t = now_; } } while (0)

  F4Tile cur, nxt;
  { long long id = 0; if (lane == 0) id = (long long)atomicAdd(A.tile_counter, 1ULL); cur.id = __shfl_sync(F4_FULL, id, 0); }
  f4_describe(cur, D, ntiles, tsegs, lane);
  uint32_t pending = cur.fast ? f4_issue(cur, D, A, W, lane) : 0u;   // bytes in flight into W.in for `cur`

#pragma unroll 1
  while (cur.id < ntiles) {
    const long long tile = cur.id;
    const int64_t s0 = tile * tsegs;
    const int ns = cur.ns;
    const int64_t raw0 = cur.raw0, raw1 = cur.so_end;
    const int al = (int)(raw0 & 15);
    const int64_t base = raw0 - al;
    nxt.id = ntiles;
    bool nxt_known = false;
    // document of the tile's first segment: one 32-ary search per tile
    int32_t doc0;
    {
      int64_t lo = 0, hi = D.n_doc;
      while (hi - lo > 1) {
        const int64_t step = (hi - lo + 31) >> 5;
        const int64_t pv = lo + step * (lane + 1);
        const bool le = pv < hi && D.doc_off[pv] <= s0;
        const int cnt = __popc(__ballot_sync(F4_FULL, le));
        const int64_t nhi = lo + step * (cnt + 1);
        lo += step * cnt;
        if (nhi < hi) hi = nhi;
      }
      doc0 = (int32_t)lo;
    }
    F4Seg R; R.mylen = R.mykeep = R.mychars = R.loff = R.koff = R.mytok = 0; R.mysp = -1; R.tot_out = R.tot_kept = 0; R.gbase = R.kbase = 0;

    if (cur.fast) {
      const int32_t span = (int32_t)(raw1 - base);
      const int nch = (span + 31) >> 5;
      const int nstep = (nch + 31) >> 5;
      // ---- P0: offsets, cleared rows
      if (lane <= ns) W.rel[lane] = (int32_t)(cur.so - base);
      if (lane == 0) W.rel[ns] = span;
      for (int k = lane; k < 3 * (F4_NCH + 2); k += 32) {               // rows that are built with atomic ORs
        const int r = k / (F4_NCH + 2), q = k - r * (F4_NCH + 2);
        W.u.m[r == 0 ? F4_BD : (r == 1 ? F4_DROP : F4_DS)][q] = 0u;
      }
      if (lane < F4_NM) { W.u.m[lane][0] = 0u; W.u.m[lane][nch + 1] = 0u; }   // "nothing there" on both sides of the tile
      W.xtok[lane] = 0; W.spmin[lane] = 0x7fffffff;
      if (pending) {
        f4_mbar_wait(f4_saddr(&W.mbar), parity);
        parity ^= 1u;
        for (int32_t k = (int32_t)pending + lane; k < span; k += 32) W.in[k] = D.text[base + k];     // the buffer's last bytes
      } else {
        for (int32_t k = lane; k < span; k += 32) W.in[k] = D.text[base + k];
      }
      __syncwarp();
      F4_PH(0);
      if (lane < ns) { const int32_t a = W.rel[lane], b = W.rel[lane + 1]; if (b > a) w32_or_bit(M.bd, a); }
      // ---- P1: classes of every chunk (lead bytes of multi-byte characters are parked in the COPY row)
      uint32_t any_lead = 0, sfany = 0;
#pragma unroll 1
      for (int st = 0; st < nstep; ++st) {
        const int c = (st << 5) + lane;
        if (c < nch) {
          const uint4 q0 = *reinterpret_cast<const uint4*>(W.in + 32 * c), q1 = *reinterpret_cast<const uint4*>(W.in + 32 * c + 16);
          const uint32_t x[8] = {q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, q1.z, q1.w};
          const W32Cls cl = w32_classify(S.lut, x);
          const uint32_t v = w32_valid(c, al, span);
          const uint32_t lead = cl.na & ~cl.ct & v;
          M.sp[c] = cl.sp & v; M.w[c] = cl.w & v; M.ct[c] = cl.ct & v; M.ep[c] = cl.ep & v; M.lt[c] = cl.lt & v; m_ns[c] = (cl.ns | lead) & v;
          c_cp[c] = lead;
          any_lead |= lead;
          sfany |= cl.sf & v;
        }
      }
      __syncwarp();
      if (__any_sync(F4_FULL, any_lead != 0)) {
#pragma unroll 1
        for (int st = 0; st < nstep; ++st) {
          const int c = (st << 5) + lane;
          const uint32_t lead = c < nch ? c_cp[c] : 0u;
          if (lead) f4_fix_multibyte(T, W.in, M, c, lead, span);
        }
        __syncwarp();
      }
      const bool has_sp = __any_sync(F4_FULL, sfany != 0) && T.n_sp > 0;
      F4_PH(1);
      // ---- P2: duplicate words
#pragma unroll 1
      for (int st = 0; st < nstep; ++st) {
        const int c = (st << 5) + lane;
        if (c >= nch) continue;
        const uint32_t generic = w32_dup_chunk_simple(W.in, M, c, w32_valid(c, al, span), span);
        if (generic) f4_dup_generic(W.in, M, c, generic, span, W.rel, ns);
      }
      __syncwarp();
      F4_PH(2);
      // ---- P3: output codes and their prefixes
      uint32_t base_os = 0, base_tb = 0;
#pragma unroll 1
      for (int st = 0; st < nstep; ++st) {
        const int c = (st << 5) + lane;
        W32Codes r; r.copy = r.space = r.ins = 0;
        uint32_t ctm = 0, bd = 0;
        if (c < nch) { r = w32_chunk_codes(W.in, M, c, al, span); ctm = M.ct[c]; bd = M.bd[c]; }
        const uint32_t s = __popc(r.space) + __popc(r.ins), o = __popc(r.copy) + s;
        const uint32_t va = o | (s << 16), vb = (uint32_t)__popc(r.copy & ctm) | ((uint32_t)__popc(bd) << 16);
        uint32_t xa = va, xb = vb;
#pragma unroll
        for (int q = 1; q < 32; q <<= 1) {
          const uint32_t ya = __shfl_up_sync(F4_FULL, xa, q), yb = __shfl_up_sync(F4_FULL, xb, q);
          if (lane >= q) { xa += ya; xb += yb; }
        }
        if (c < nch) { c_cp[c] = r.copy; c_se[c] = r.space; c_in[c] = r.ins; p_os[c] = base_os + xa - va; p_tb[c] = base_tb + xb - vb; }
        base_os += __shfl_sync(F4_FULL, xa, 31);
        base_tb += __shfl_sync(F4_FULL, xb, 31);
      }
      __syncwarp();
      F4_PH(3);
      // ---- P4: per segment (lane i <-> segment i): offsets, lengths, words, code points
      const int32_t tot_out = (int32_t)(base_os & 0xffffu);
      const int32_t tot_sp = (int32_t)(base_os >> 16), tot_ct = (int32_t)(base_tb & 0xffffu);
      int32_t po = tot_out, ps = tot_sp, pt = tot_ct, br = 0;
      bool nonempty = false;
      if (lane <= ns) {
        const int32_t k = W.rel[lane];
        if (k < span) {
          const int c = k >> 5;
          const uint32_t below = (1u << (k & 31)) - 1u;
          const uint32_t cp = c_cp[c], se = c_se[c], in = c_in[c];
          const uint32_t os = p_os[c], tb = p_tb[c];
          po = (int32_t)(os & 0xffffu) + __popc((cp | se) & below) + __popc(in & below);
          ps = (int32_t)(os >> 16) + __popc((se | in) & below);
          pt = (int32_t)(tb & 0xffffu) + __popc(cp & M.ct[c] & below);
          br = (int32_t)(tb >> 16) + __popc(M.bd[c] & below);
        }
        nonempty = lane < ns && W.rel[lane + 1] > k;
      }
      int32_t po_n = __shfl_down_sync(F4_FULL, po, 1), ps_n = __shfl_down_sync(F4_FULL, ps, 1), pt_n = __shfl_down_sync(F4_FULL, pt, 1);
      if (lane == 31) { po_n = tot_out; ps_n = tot_sp; pt_n = tot_ct; }
      int32_t mynw = 0, wfirst = 0, tot_words = 0;
      if (lane < ns) {
        R.loff = po; R.mylen = po_n - po; R.mykeep = R.mylen > 0;
        mynw = R.mykeep ? ps_n - ps + 1 : 0;
        R.mychars = R.mylen - (pt_n - pt);
      }
      {
        const uint32_t kb = __ballot_sync(F4_FULL, R.mykeep != 0);
        R.koff = __popc(kb & lt_mask); R.tot_kept = __popc(kb);
      }
      R.tot_out = tot_out;
      wfirst = f4_excl_scan(mynw, lane, &tot_words);
      f4_publish(A, tile, tot_out, R.tot_kept, lane);
      if (lane <= ns) { W.loff[lane] = (uint16_t)(lane < ns ? R.loff : tot_out); W.wfirst[lane] = (uint16_t)wfirst; }
      if (lane == 31 && ns == 32) { W.loff[32] = (uint16_t)tot_out; W.wfirst[32] = (uint16_t)tot_words; }
      if (nonempty) W.kr[br] = (uint8_t)(R.koff + (R.mykeep ? 1 : 0));
      if (R.mykeep) W.wstart[wfirst] = (uint16_t)(R.loff | 0x8000);
      if (lane == 0) W.wstart[tot_words] = (uint16_t)tot_out;
      __syncwarp();
      F4_PH(4);
      // ---- P5: scatter into the cleaned tile, word starts
#pragma unroll 1
      for (int st = 0; st < nstep; ++st) {
        const int c = (st << 5) + lane;
        if (c >= nch) continue;
        const uint32_t cp = c_cp[c], se = c_se[c], in = c_in[c];
        if (!(cp | se)) continue;
        const uint32_t os = p_os[c];
        int32_t pos = (int32_t)(os & 0xffffu);
        uint32_t sm = se | in;
        if (sm) {                                   // a word starts at every emitted space
          const int32_t sbase = (int32_t)(os >> 16);
          const uint32_t bdm = M.bd[c];
          const int32_t brb = (int32_t)(p_tb[c] >> 16);
          while (sm) {
            const int bit = __ffs(sm) - 1;
            sm &= sm - 1;
            const uint32_t below = (1u << bit) - 1u;
            const int32_t wpos = pos + __popc((cp | se) & below) + __popc(in & below);
            const int32_t rank = brb + __popc(bdm & (below | (1u << bit))) - 1;       // segment whose start is at or before this byte
            const int32_t w = sbase + __popc((se | in) & below) + W.kr[rank];
            W.wstart[w] = (uint16_t)wpos;
          }
        }
        const uint4 q0 = *reinterpret_cast<const uint4*>(W.in + 32 * c), q1 = *reinterpret_cast<const uint4*>(W.in + 32 * c + 16);
        const uint32_t x[8] = {q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, q1.z, q1.w};
        if (in == 0u && (cp | se) == 0xffffffffu && !(se & m_ns[c])) { f4_store32(W.out + pos, x); continue; }
        uint8_t* dst = W.out;
#pragma unroll
        for (int j = 0; j < 8; ++j) {               // a word of four bytes at a time; untouched words are common
          const uint32_t c4 = (cp >> (4 * j)) & 15u, s4 = (se >> (4 * j)) & 15u, i4 = (in >> (4 * j)) & 15u;
          if (!(c4 | s4)) continue;
          const uint32_t xv = x[j];
#pragma unroll
          for (int b = 0; b < 4; ++b) {
            if ((i4 >> b) & 1u) dst[pos++] = ' ';
            if ((c4 >> b) & 1u) dst[pos++] = (uint8_t)(xv >> (8 * b));
            else if ((s4 >> b) & 1u) dst[pos++] = ' ';
          }
        }
      }
      __syncwarp();
      pending = 0u;
      F4_PH(5);
      // ---- P6: special-token literals in the cleaned tile (skipped when no byte can start one)
      if (has_sp) f4_special_scan(T, W, ns, tot_out, lane);
      F4_PH(6);
      // ---- P7: tokens, word by word.  The mask rows are dead: counts, miss list and slow list live there.
      uint8_t* const cnt = W.u.t.cnt; uint16_t* const miss = W.u.t.miss; uint16_t* const slow = W.u.t.slow;
      int32_t nmiss = 0, nslow = 0;
#pragma unroll 1
      for (int32_t wb = 0; wb < tot_words; wb += 32) {
        const int32_t w = wb + lane;
        int mc = -2;                                   // -2 no word, -1 unknown, >= 0 count, F3_MEMO_SLOW: generic matcher
        if (w < tot_words) {
          uint64_t lo, hi; int Lw, vs; int32_t st, n;
          f4_word_key(W, w, lo, hi, Lw, vs, st, n);
          mc = -1;
          if (Lw <= 15) {
            uint32_t slot8 = 0xffffffffu;
            if (Lw <= 7) {
              slot8 = trc_key8_slot(lo, TRC_HOT8 - 1);
              const uint64_t ce = S.hot8[slot8];
              if ((ce & 0x00ffffffffffffffULL) == lo) mc = (int)(ce >> 56);
            }
            if (mc < 0 && A.memo != nullptr && lo != 0) {
              const ulonglong2 e = __ldcg(reinterpret_cast<const ulonglong2*>(A.memo) + f3_memo_slot(lo, hi, A.memo_mask));
              if (e.x == lo && (e.y & 0x00ffffffffffffffULL) == hi && (e.y >> 56)) {
                mc = (int)(e.y >> 56) - 1;
                if (slot8 != 0xffffffffu) S.hot8[slot8] = lo | ((uint64_t)mc << 56);
              }
            }
          }
          cnt[w] = (uint8_t)((mc >= 0 && mc != F3_MEMO_SLOW) ? mc : 0);
        }
        const uint32_t bm = __ballot_sync(F4_FULL, mc == -1), bs = __ballot_sync(F4_FULL, mc == F3_MEMO_SLOW);
        if (mc == -1) miss[nmiss + __popc(bm & lt_mask)] = (uint16_t)w;
        if (mc == F3_MEMO_SLOW) { const int q = nslow + __popc(bs & lt_mask); if (q < F4_SLOW_CAP) slow[q] = (uint16_t)w; else cnt[w] = 255; }
        nmiss += __popc(bm); nslow += __popc(bs);
      }
      __syncwarp();
      F4_PH(7);
      if (nmiss) nslow = f4_match_words(T, A, S.hot8, W, nmiss, nslow, lane);
      F4_PH(8);
      // ---- P8: global offsets (the predecessors published long ago)
      f4_resolve(A, tile, ntiles, lane, &R.gbase, &R.kbase);
      // the next tile is claimed only now: a tile that is claimed but not started holds back the look-back of every
      // later tile, so the claim sits behind this tile's last wait; its raw bytes arrive under the copy-out
      {
        long long id = 0;
        if (lane == 0) id = (long long)atomicAdd(A.tile_counter, 1ULL);
        nxt.id = __shfl_sync(F4_FULL, id, 0);
        f4_describe(nxt, D, ntiles, tsegs, lane);
        pending = nxt.fast ? f4_issue(nxt, D, A, W, lane) : 0u;
        nxt_known = true;
      }
      F4_PH(9);
      f3_copy_out(D.C + R.gbase, W.out, tot_out, lane, 32);     // the cleaned tile leaves shared memory once, coalesced
      F4_PH(10);
      if (nslow) f4_slow_words(T, byte_id, A, W, ns, nslow, tot_words, R.koff, R.gbase, R.kbase, lane);
      // per segment: sum of its words
      if (lane < ns && R.mykeep) {
        int32_t t = W.xtok[lane];
        const int32_t w1 = wfirst + mynw;
        for (int32_t w = wfirst; w < w1; ++w) t += cnt[w];
        R.mytok = t;
        const int32_t v = W.spmin[lane];
        R.mysp = v == 0x7fffffff ? -1 : (v & 15);
      }
    } else {
      f4_slow_lane(D, T, byte_id, A, tile, ntiles, s0, ns, cur.so, raw1, lane, R);
    }
    F4_PH(11);
    // ---- outputs
    if (lane < ns) {
      D.cchars[s0 + lane] = R.mychars;
      if (R.mykeep) {
        const int64_t j = R.kbase + R.koff;
        const int64_t i = s0 + lane;
        D.k2r[j] = (int32_t)i; D.coff[j] = R.gbase + R.loff; D.ctok[j] = (uint32_t)R.mytok; A.seg_special[j] = R.mysp;
        int64_t lo = doc0;                    // documents are rarely shorter than a few segments: walk, then search
        for (int q = 0; q < 4 && D.doc_off[lo + 1] <= i; ++q) ++lo;
        if (D.doc_off[lo + 1] <= i) {
          int64_t hi = D.n_doc;
          while (hi - lo > 1) { const int64_t mid = (lo + hi) >> 1; if (D.doc_off[mid] <= i) lo = mid; else hi = mid; }
        }
        D.sdoc[j] = (int32_t)lo;
      }
    }
    if (lane == 0 && s0 + ns == D.n) {
      D.coff[R.kbase + R.tot_kept] = R.gbase + R.tot_out;
      A.totals[0] = R.kbase + R.tot_kept; A.totals[1] = R.gbase + R.tot_out;
    }
    __syncwarp();
    F4_PH(12);
    // ---- next tile
    if (!nxt_known) {
      long long id = 0;
      if (lane == 0) id = (long long)atomicAdd(A.tile_counter, 1ULL);
      nxt.id = __shfl_sync(F4_FULL, id, 0);
      f4_describe(nxt, D, ntiles, tsegs, lane);
      pending = nxt.fast ? f4_issue(nxt, D, A, W, lane) : 0u;
    }
    cur = nxt;
  }
}
