// K1a + K2 fused, third generation: every lane owns an 8-byte CELL of the tile (one 64-bit load), the clean
// rules (preprocessor.py:38,69-89) are evaluated on bit masks (trc_swar.cuh) and only the rare cases
// (multi-byte characters, whitespace runs across a cell edge, duplicate-word candidates) take scalar side
// paths.  A warp covers 256 bytes per step.  Tile scheduling, look-back compaction, word-parallel token
// counting (big_chunkeroosky.py:83 via DESIGN.md §3) and the slow lane are those of trc_fused.cuh.
#pragma once

#define F3_T 512
#define F3_NW (F3_T / 32)
#define F3_SEGS 256
#define F3_IN_CAP 18432
#define F3_OUT_CAP 26624
#define F3_WORD_CAP 4608
#define F3_CAND_CAP 2560
#define F3_NCELL (F3_IN_CAP / 8)
#define F3_NSTEP (F3_IN_CAP / 256)
#define F3_MASKS 9
#define F3_MSTRIDE (F3_NCELL + 8)
#define F3_HINT_LOG2 4

struct F3Smem {
  union alignas(16) {                         // raw tile; after the scatter: sorted word list
    unsigned long long in64[F3_IN_CAP / 8 + 4];
    uint8_t in[F3_IN_CAP + 32];
    struct { uint16_t wstart[F3_WORD_CAP]; uint16_t wseg[F3_WORD_CAP]; } w;
  } y;
  union alignas(16) {                         // bit masks (sp w ct lt ep bd drop ds lead); after the word list: slow words
    uint8_t m[F3_MASKS][F3_MSTRIDE];
    uint16_t slow[2 * F3_WORD_CAP];           // slow words, then cache-miss words
  } z;
  alignas(16) uint8_t out[F3_OUT_CAP + 32];
  uint64_t hot8[TRC_HOT8];
  uint16_t cand[F3_CAND_CAP];
  uint16_t pre_o[F3_NCELL + 32], pre_s[F3_NCELL + 32];   // per cell: output bytes / spaces before it, inside its step
  uint16_t st_o[F3_NSTEP + 1], st_s[F3_NSTEP + 1], st_t[F3_NSTEP + 1];
  uint16_t blk_seg[F3_IN_CAP >> F3_HINT_LOG2];   // segment of the first byte of every 16-byte block
  uint16_t sg_o[F3_SEGS + 1], sg_s[F3_SEGS + 1], sg_t[F3_SEGS + 1];   // step-relative prefixes at segment starts
  uint16_t spb[F3_SEGS + 1], ctb[F3_SEGS + 1];
  int32_t rel[F3_SEGS + 1];
  int32_t loff[F3_SEGS + 1];
  int32_t len[F3_SEGS], koff[F3_SEGS], tok[F3_SEGS], spmin[F3_SEGS];
  int32_t wsum[3][F3_NW];
  int32_t tot[3];
  int32_t ncand, nslow, nmiss, has_sp;
  unsigned int tile;
  int32_t doc0;                             // document of the tile's first segment
  long long gbase, kbase, slbase;
};
static_assert(sizeof(F3Smem) <= 115200, "two CTAs per SM");

__device__ __forceinline__ void f3_scan3(F3Smem& S, int lane, int wid, int32_t v0, int32_t v1, int32_t v2, int32_t* e0, int32_t* e1, int32_t* e2) {
  int32_t x0 = v0, x1 = v1, x2 = v2;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    int32_t y0 = __shfl_up_sync(0xffffffffu, x0, o), y1 = __shfl_up_sync(0xffffffffu, x1, o), y2 = __shfl_up_sync(0xffffffffu, x2, o);
    if (lane >= o) { x0 += y0; x1 += y1; x2 += y2; }
  }
  __syncthreads();
  if (lane == 31) { S.wsum[0][wid] = x0; S.wsum[1][wid] = x1; S.wsum[2][wid] = x2; }
  __syncthreads();
  if (wid == 0) {
    int32_t w0 = lane < F3_NW ? S.wsum[0][lane] : 0, w1 = lane < F3_NW ? S.wsum[1][lane] : 0, w2 = lane < F3_NW ? S.wsum[2][lane] : 0;
#pragma unroll
    for (int o = 1; o < F3_NW; o <<= 1) {
      int32_t y0 = __shfl_up_sync(0xffffffffu, w0, o), y1 = __shfl_up_sync(0xffffffffu, w1, o), y2 = __shfl_up_sync(0xffffffffu, w2, o);
      if (lane >= o) { w0 += y0; w1 += y1; w2 += y2; }
    }
    if (lane < F3_NW) { S.wsum[0][lane] = w0; S.wsum[1][lane] = w1; S.wsum[2][lane] = w2; }
    if (lane == F3_NW - 1) { S.tot[0] = w0; S.tot[1] = w1; S.tot[2] = w2; }
  }
  __syncthreads();
  *e0 = (wid ? S.wsum[0][wid - 1] : 0) + x0 - v0;
  *e1 = (wid ? S.wsum[1][wid - 1] : 0) + x1 - v1;
  *e2 = (wid ? S.wsum[2][wid - 1] : 0) + x2 - v2;
}

// segment of tile byte k, starting the search at the first segment of k's step
__device__ __forceinline__ int f3_seg_of(const F3Smem& S, int ns, int32_t k) {
  int i = S.blk_seg[k >> F3_HINT_LOG2];
  while (i + 1 < ns && S.rel[i + 1] <= k) ++i;
  return i;
}

// Copies n bytes of the shared tile (src 16-byte aligned, readable a few words past n) to global memory with
// 16-byte stores aligned to the DESTINATION: five aligned shared words and four funnel shifts per store; the
// bytes before the first and after the last aligned 16 go one by one.  Threads t of nt cooperate.
__device__ __forceinline__ void f3_copy_out(uint8_t* dstg, const uint8_t* src, int32_t n, int t, int nt) {
  const int head = (int)((16 - ((uintptr_t)dstg & 15)) & 15);
  if (n < 64 + head) { for (int32_t k = t; k < n; k += nt) dstg[k] = src[k]; return; }
  if (t < head) dstg[t] = src[t];
  const int32_t nq = (n - head) >> 4;
  const uint32_t* sw = reinterpret_cast<const uint32_t*>(src) + (head >> 2);
  const unsigned sh = (unsigned)(head & 3) * 8u;
  uint4* dq = reinterpret_cast<uint4*>(dstg + head);
  for (int32_t i = t; i < nq; i += nt) {
    const uint32_t* w = sw + 4 * i;
    const uint32_t a = w[0], b = w[1], c = w[2], d = w[3], e = w[4];
    dq[i] = make_uint4(__funnelshift_r(a, b, sh), __funnelshift_r(b, c, sh), __funnelshift_r(c, d, sh), __funnelshift_r(d, e, sh));
  }
  const int32_t done = head + (nq << 4);
  if (t < n - done) dstg[done + t] = src[done + t];
}

#define F3_MATCH_T 128     // threads that run the packed-word matcher when only a few words are queued
#define F3_MEMO_SLOW 254   // count value that marks a word the packed matcher cannot decide (-> slow list)
// ---- word memo (FuseArgs::memo).  An entry only ever goes from {0, 0} to its final value, one aligned 64-bit
// store per half, so whatever mixture of old and new halves a reader sees, it matches only a fully written entry:
// the second half carries count+1 in its top byte (0 = not written) and keys whose first half is 0 stay out.
__device__ __forceinline__ uint32_t f3_memo_slot(uint64_t lo, uint64_t hi, uint32_t mask) {
  uint32_t x = (uint32_t)lo ^ ((uint32_t)(lo >> 32) * 0x9E3779B1u) ^ ((uint32_t)hi * 0x85EBCA77u) ^ ((uint32_t)(hi >> 32) * 0xC2B2AE3Du);
  x *= 0x85EBCA6Bu; x ^= x >> 15; x *= 0xC2B2AE35u; x ^= x >> 13;
  return x & mask;
}
__device__ __forceinline__ int f3_memo_get(const FuseArgs& A, uint64_t lo, uint64_t hi) {
  if (!A.memo || lo == 0) return -1;
  const ulonglong2 e = __ldcg(reinterpret_cast<const ulonglong2*>(A.memo) + f3_memo_slot(lo, hi, A.memo_mask));
  if (e.x == lo && (e.y & 0x00ffffffffffffffULL) == hi && (e.y >> 56)) return (int)(e.y >> 56) - 1;
  return -1;
}
__device__ __forceinline__ void f3_memo_put(const FuseArgs& A, uint64_t lo, uint64_t hi, uint32_t c) {
  if (!A.memo || lo == 0 || c > (uint32_t)F3_MEMO_SLOW) return;
  const uint32_t slot = f3_memo_slot(lo, hi, A.memo_mask);
  if (atomicCAS(&A.memo_lock[slot], 0u, 1u) != 0u) return;       // write-once: the first word to claim a slot keeps it
  st_relaxed(&A.memo[2 * (size_t)slot], lo);
  st_relaxed(&A.memo[2 * (size_t)slot + 1], hi | ((uint64_t)(c + 1u) << 56));
}

__global__ void __launch_bounds__(F3_T, 2) k_clean_tok_fused3(TrcDev D, TrcTables T, const uint32_t* __restrict__ byte_id, FuseArgs A,
                                                              int64_t ntiles) {
  extern __shared__ __align__(16) unsigned char fuse_raw[];
  F3Smem& S = *reinterpret_cast<F3Smem*>(fuse_raw);
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  // S.hot8 is used as a WORD cache here: packed bytes of a word of <= 7 bytes (space included) | tokens << 56.
  // Exact (the key is the whole word), filled on the fly, kept across tiles.
  for (int k = tid; k < TRC_HOT8; k += F3_T) S.hot8[k] = 0ULL;
  TrcFast F; F.acls = T.acls; F.hot8 = nullptr; F.hot8_mask = 0;
  SwMasks M;
  M.sp = S.z.m[0]; M.w = S.z.m[1]; M.ct = S.z.m[2]; M.lt = S.z.m[3]; M.ep = S.z.m[4]; M.bd = S.z.m[5]; M.drop = S.z.m[6]; M.ds = S.z.m[7];
  uint8_t* const mlead = S.z.m[8];
  uint8_t* const c_cp = M.drop; uint8_t* const c_se = M.ds; uint8_t* const c_in = M.lt;    // codes overwrite own-cell-only masks
  __syncthreads();
  long long ph_t = clock64();
  // (Claiming the next tile early, while the current one is processed, hides the atomic's round trip and gains 2 % on
  // uniform tiles - but a tile id held back behind a long slow-lane tile stalls the look-back of every later tile:
  // config 4 went from 1.3 to 5.8 ms.  Tiles are claimed when they are started.)
  for (;;) {
    if (tid == 0) S.tile = (unsigned int)atomicAdd(A.tile_counter, 1ULL);
    __syncthreads();
    const unsigned tile = S.tile;
    if ((int64_t)tile >= ntiles) break;
    const int64_t s0 = (int64_t)tile * F3_SEGS;
    const int ns = (int)((D.n - s0) < F3_SEGS ? (D.n - s0) : F3_SEGS);
    if (wid == F3_NW - 1) {                   // document of the first segment: one 32-ary search per tile, under the tile load
      int64_t lo = 0, hi = D.n_doc;           // doc_off[lo] <= s0 < doc_off[hi]
      while (hi - lo > 1) {
        const int64_t step = (hi - lo + 31) >> 5;
        const int64_t pv = lo + step * (lane + 1);
        const bool le = pv < hi && D.doc_off[pv] <= s0;
        const int cnt = __popc(__ballot_sync(0xffffffffu, le));       // doc_off is sorted: the lanes that say yes are a prefix
        const int64_t nhi = lo + step * (cnt + 1);
        lo += step * cnt;
        if (nhi < hi) hi = nhi;
      }
      if (lane == 0) S.doc0 = (int32_t)lo;
    }
    const int64_t raw0 = D.seg_off[s0], raw1 = D.seg_off[s0 + ns];
    const int al = (int)(raw0 & 15);
    const int64_t base = raw0 - al;
    const int32_t span = (int32_t)((raw1 - base) < (1 << 30) ? (raw1 - base) : (1 << 30));
    const bool fast_in = (raw1 - base) <= F3_IN_CAP;
    for (int k = tid; k <= ns; k += F3_T) S.rel[k] = (int32_t)(D.seg_off[s0 + k] - base);
    __syncthreads();
    int32_t mylen = 0, mynw = 0, mykeep = 0, mychars = 0, loff = 0, woff = 0, koff = 0;
    int32_t tot_out = 0, tot_words = 0, tot_kept = 0;
    long long gbase = 0, kbase = 0;
    bool fast_out = false;
    int32_t gq = -1;
    if (fast_in) {
      const int ncell = (span + 7) >> 3, nstep = (span + 255) >> 8;
      M.ncell = ncell;
      // ---- load the tile (16-byte loads), clear the masks that are built with atomics
      for (int32_t k = tid * 16; k < nstep * 256 + 16; k += F3_T * 16) {
        int4 v = make_int4(0, 0, 0, 0);
        if (k < span) {
          if (base + k + 16 <= A.text_bytes) v = *reinterpret_cast<const int4*>(D.text + base + k);
          else { __align__(16) uint8_t tmp[16]; for (int q = 0; q < 16; ++q) tmp[q] = base + k + q < A.text_bytes ? D.text[base + k + q] : 0; v = *reinterpret_cast<int4*>(tmp); }
        }
        *reinterpret_cast<int4*>(&S.y.in[k]) = v;
      }
      for (int k = tid; k < (F3_MSTRIDE * 3) / 4; k += F3_T) reinterpret_cast<uint32_t*>(S.z.m[5])[k] = 0u;   // bd, drop, ds
      if (tid == 0) { S.ncand = 0; S.has_sp = T.sp_single < 0x100u ? 0 : 1; }
      __syncthreads();
      FUSE_PH(0);
      if (tid < ns) {
        const int32_t a = S.rel[tid], b = S.rel[tid + 1];
        if (b > a) {
          sw_or_bit(M.bd, a);
          for (int32_t s = (a + (1 << F3_HINT_LOG2) - 1) >> F3_HINT_LOG2; (s << F3_HINT_LOG2) < b; ++s) S.blk_seg[s] = (uint16_t)tid;
        }
      }
      if (tid == 0) S.blk_seg[0] = 0;
      // ---- A: ASCII classes of every cell
      for (int st = wid; st < nstep; st += F3_NW) {
        const int c = (st << 5) + lane;
        const unsigned long long xx = S.y.in64[c];
        const SwCell cl = sw_classify(xx);
        if (T.sp_single < 0x100u) {
          const uint64_t tt = (xx & ~SW_H) | SW_H;
          if (sw_ge(tt, T.sp_single) & ~sw_ge(tt, T.sp_single + 1) & ~(xx & SW_H)) S.has_sp = 1;
        }
        M.sp[c] = (uint8_t)cl.sp; M.w[c] = (uint8_t)cl.w; M.ct[c] = (uint8_t)cl.ct; M.lt[c] = (uint8_t)cl.lt; M.ep[c] = (uint8_t)cl.ep;
        mlead[c] = (uint8_t)cl.lead;
      }
      if (tid < 8) { M.w[(nstep << 5) + tid] = 0; M.sp[(nstep << 5) + tid] = 0; }   // window reads may look two cells past the tile
      __syncthreads();
      // ---- A2: multi-byte characters (code point lookup by the owner of the lead byte)
      for (int st = wid; st < nstep; st += F3_NW) {
        const int c = (st << 5) + lane;
        const uint32_t ld = mlead[c];
        if (ld) sw_fix_multibyte(T, S.y.in, M, c, ld, span);
      }
      __syncthreads();
      FUSE_PH(1);
      // ---- B: duplicate-word candidates, compacted, then decided densely
      for (int st = wid; st < nstep; st += F3_NW) {
        const int c = (st << 5) + lane;
        const int32_t k0 = c << 3;
        uint32_t valid = 0xffu;
        if (k0 < al) valid &= (al - k0 >= 8) ? 0u : (0xffu << (al - k0));
        if (k0 + 8 > span) valid &= (span <= k0) ? 0u : (0xffu >> (k0 + 8 - span));
        uint32_t cand = (c < ncell) ? sw_dup_candidates(M, c, valid & 0xffu) : 0u;
        {                                        // most candidates fail a two-byte test and never reach the list
          uint32_t t = cand;
          while (t) {
            const int bit = __ffs(t) - 1;
            t &= t - 1;
            if (sw_dup_quick_reject(S.y.in, M, k0 + bit, span)) cand &= ~(1u << bit);
          }
        }
        const int cnt = __popc(cand);
        int x = cnt;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { int y = __shfl_up_sync(0xffffffffu, x, o); if (lane >= o) x += y; }
        int b0 = 0;
        if (lane == 31 && x) b0 = atomicAdd(&S.ncand, x);
        b0 = __shfl_sync(0xffffffffu, b0, 31);
        int slot = b0 + x - cnt;
        while (cand) {
          const int bit = __ffs(cand) - 1;
          cand &= cand - 1;
          if (slot < F3_CAND_CAP) S.cand[slot] = (uint16_t)(k0 + bit);
          ++slot;
        }
      }
      __syncthreads();
      {
        const int32_t nc = S.ncand;
        if (nc <= F3_CAND_CAP) {
          for (int32_t q = tid; q < nc; q += F3_T) {
            const int32_t k = S.cand[q];
            const int i = f3_seg_of(S, ns, k);
            sw_mark_dup(S.y.in, M, k, S.rel[i], S.rel[i + 1]);
          }
        } else {                                 // list overflow: every cell walks its own candidates
          for (int st = wid; st < nstep; st += F3_NW) {
            const int c = (st << 5) + lane;
            const int32_t k0 = c << 3;
            uint32_t valid = 0xffu;
            if (k0 < al) valid &= (al - k0 >= 8) ? 0u : (0xffu << (al - k0));
            if (k0 + 8 > span) valid &= (span <= k0) ? 0u : (0xffu >> (k0 + 8 - span));
            uint32_t cand = (c < ncell) ? sw_dup_candidates(M, c, valid & 0xffu) : 0u;
            while (cand) {
              const int bit = __ffs(cand) - 1;
              cand &= cand - 1;
              const int i = f3_seg_of(S, ns, k0 + bit);
              sw_mark_dup(S.y.in, M, k0 + bit, S.rel[i], S.rel[i + 1]);
            }
          }
        }
      }
      __syncthreads();
      FUSE_PH(2);
      // ---- C: output codes, per-cell prefixes inside the step, step totals, prefixes at segment starts
      for (int st = wid; st < nstep; st += F3_NW) {
        const int c = (st << 5) + lane;
        SwCodes r; r.copy = r.space = r.ins = 0;
        uint32_t ctm = 0, bd = 0;
        if (c < ncell) { r = sw_cell_codes(S.y.in, M, c, al, span); ctm = M.ct[c]; bd = M.bd[c]; }
        const uint32_t o = __popc(r.copy) + __popc(r.space) + __popc(r.ins), s = __popc(r.space) + __popc(r.ins), t = __popc(r.copy & ctm);
        uint32_t v = o | (s << 11) | (t << 21), x = v;
#pragma unroll
        for (int q = 1; q < 32; q <<= 1) { uint32_t y = __shfl_up_sync(0xffffffffu, x, q); if (lane >= q) x += y; }
        const uint32_t ex = x - v;
        if (c < ncell) { c_cp[c] = (uint8_t)r.copy; c_se[c] = (uint8_t)r.space; c_in[c] = (uint8_t)r.ins; }
        S.pre_o[c] = (uint16_t)(ex & 0x7ffu); S.pre_s[c] = (uint16_t)((ex >> 11) & 0x3ffu);
        if (lane == 31) { S.st_o[st] = (uint16_t)(x & 0x7ffu); S.st_s[st] = (uint16_t)((x >> 11) & 0x3ffu); S.st_t[st] = (uint16_t)(x >> 21); }
        const int32_t k0 = c << 3;
        if (k0 + 8 <= al) bd = 0; else if (k0 < al) bd &= 0xffu << (al - k0);
        while (bd) {                              // segments that start in this cell
          const int bit = __ffs(bd) - 1;
          bd &= bd - 1;
          const int32_t k = k0 + bit;
          if (k >= span) break;
          const uint32_t below = (1u << bit) - 1u;
          const uint16_t po = (uint16_t)((ex & 0x7ffu) + __popc((r.copy | r.space) & below) + __popc(r.ins & below));
          const uint16_t ps = (uint16_t)(((ex >> 11) & 0x3ffu) + __popc((r.space | r.ins) & below));
          const uint16_t pt = (uint16_t)((ex >> 21) + __popc(r.copy & ctm & below));
          const int i = f3_seg_of(S, ns, k);
          for (int ii = i; ii >= 0 && S.rel[ii] == k; --ii) { S.sg_o[ii] = po; S.sg_s[ii] = ps; S.sg_t[ii] = pt; }
        }
      }
      __syncthreads();
      if (wid == 0) {                       // exclusive scans of the step totals (at most 72 steps)
        int32_t c0 = 0, c1 = 0, c2 = 0;
        for (int b0 = 0; b0 < nstep; b0 += 32) {
          const bool in_r = b0 + lane < nstep;
          int32_t v0 = in_r ? S.st_o[b0 + lane] : 0, v1 = in_r ? S.st_s[b0 + lane] : 0, v2 = in_r ? S.st_t[b0 + lane] : 0;
          int32_t x0 = v0, x1 = v1, x2 = v2;
#pragma unroll
          for (int o = 1; o < 32; o <<= 1) {
            int32_t y0 = __shfl_up_sync(0xffffffffu, x0, o), y1 = __shfl_up_sync(0xffffffffu, x1, o), y2 = __shfl_up_sync(0xffffffffu, x2, o);
            if (lane >= o) { x0 += y0; x1 += y1; x2 += y2; }
          }
          if (in_r) { S.st_o[b0 + lane] = (uint16_t)(c0 + x0 - v0); S.st_s[b0 + lane] = (uint16_t)(c1 + x1 - v1); S.st_t[b0 + lane] = (uint16_t)(c2 + x2 - v2); }
          c0 += __shfl_sync(0xffffffffu, x0, 31); c1 += __shfl_sync(0xffffffffu, x1, 31); c2 += __shfl_sync(0xffffffffu, x2, 31);
        }
        if (lane == 0) { S.st_o[nstep] = (uint16_t)c0; S.st_s[nstep] = (uint16_t)c1; S.st_t[nstep] = (uint16_t)c2; }
      }
      __syncthreads();
      tot_out = S.st_o[nstep];
      const int32_t tot_sp = S.st_s[nstep], tot_ct = S.st_t[nstep];
      for (int i = tid; i <= ns; i += F3_T) {
        const int32_t k = S.rel[i];
        if (k >= span) { S.loff[i] = tot_out; S.spb[i] = (uint16_t)tot_sp; S.ctb[i] = (uint16_t)tot_ct; }
        else { const int s = k >> 8; S.loff[i] = S.st_o[s] + S.sg_o[i]; S.spb[i] = (uint16_t)(S.st_s[s] + S.sg_s[i]); S.ctb[i] = (uint16_t)(S.st_t[s] + S.sg_t[i]); }
      }
      __syncthreads();
      FUSE_PH(3);
      int32_t myspb = 0;
      if (tid < ns) {
        loff = S.loff[tid]; mylen = S.loff[tid + 1] - loff; mykeep = mylen > 0;
        myspb = S.spb[tid];
        mynw = mykeep ? (int32_t)S.spb[tid + 1] - myspb + 1 : 0;
        mychars = mylen - ((int32_t)S.ctb[tid + 1] - (int32_t)S.ctb[tid]);
      }
      int32_t d1;
      f3_scan3(S, lane, wid, mykeep, mynw, 0, &koff, &woff, &d1);
      tot_kept = S.tot[0]; tot_words = S.tot[1];
      if (tid < ns) { S.len[tid] = mylen; S.koff[tid] = koff; S.spmin[tid] = 0x7fffffff; }
      fast_out = tot_out <= F3_OUT_CAP;
      if (tid == 0) { fuse_publish(A.desc_bytes, tile, tot_out); fuse_publish(A.desc_kept, tile, tot_kept); }
      if (!fast_out) {                        // scatter straight to global memory needs the offsets now
        if (wid == 0) {
          long long gb = fuse_resolve(A.desc_bytes, tile, tot_out, lane);
          long long kb = fuse_resolve(A.desc_kept, tile, tot_kept, lane);
          if (lane == 0) { S.gbase = gb; S.kbase = kb; }
        }
        __syncthreads();
        gbase = S.gbase; kbase = S.kbase;
      }
      const bool resolved_early = !fast_out;
      FUSE_PH(4);
      // ---- D: scatter
      {
        uint8_t* dst = fast_out ? S.out : D.C + gbase;
        for (int st = wid; st < nstep; st += F3_NW) {
          const int c = (st << 5) + lane;
          if (c >= ncell) continue;
          uint32_t cp = c_cp[c], se = c_se[c], in = c_in[c];
          if (!(cp | se)) continue;
          int32_t pos = S.st_o[st] + S.pre_o[c];
          const unsigned long long x = S.y.in64[c];
          if (in == 0u && (cp | se) == 0xffu) {   // the common cell: every byte yields exactly one byte
            uint8_t* d8 = dst + pos;
#pragma unroll
            for (int bit = 0; bit < 8; ++bit) d8[bit] = ((se >> bit) & 1u) ? (uint8_t)' ' : (uint8_t)(x >> (8 * bit));
            continue;
          }
#pragma unroll
          for (int bit = 0; bit < 8; ++bit) {
            if ((in >> bit) & 1u) dst[pos++] = ' ';
            if ((cp >> bit) & 1u) dst[pos++] = (uint8_t)(x >> (8 * bit));
            else if ((se >> bit) & 1u) dst[pos++] = ' ';
          }
        }
      }
      __syncthreads();
      FUSE_PH(5);
      const bool out_in_smem = fast_out;
      fast_out = fast_out && tot_words <= F3_WORD_CAP;
      if (out_in_smem && !fast_out) {         // too many words for the list: resolve now, write the tile out, count per segment
        if (wid == 0) {
          long long gb = fuse_resolve(A.desc_bytes, tile, tot_out, lane);
          long long kb = fuse_resolve(A.desc_kept, tile, tot_kept, lane);
          if (lane == 0) { S.gbase = gb; S.kbase = kb; }
        }
        __syncthreads();
        gbase = S.gbase; kbase = S.kbase;
        uint8_t* dstg = D.C + gbase;
        f3_copy_out(dstg, S.out, tot_out, tid, F3_T);
      }
      (void)resolved_early;
      if (fast_out) {
        // sorted word list (the raw tile is dead): a word starts at every emitted space and at every segment start
        for (int st = wid; st < nstep; st += F3_NW) {
          const int c = (st << 5) + lane;
          if (c >= ncell) continue;
          const uint32_t cp = c_cp[c], se = c_se[c], in = c_in[c];
          uint32_t sm = se | in;
          if (!sm) continue;
          const int32_t pbase = S.st_o[st] + S.pre_o[c], sbase = S.st_s[st] + S.pre_s[c];
          const int32_t k0 = c << 3;
          while (sm) {
            const int bit = __ffs(sm) - 1;
            const uint32_t below = (1u << bit) - 1u;
            const int i = f3_seg_of(S, ns, k0 + bit);
            const int32_t pos = pbase + __popc((cp | se) & below) + __popc(in & below);
            const int32_t w = sbase + __popc((se | in) & below) + S.koff[i] + 1;
            S.y.w.wstart[w] = (uint16_t)pos; S.y.w.wseg[w] = (uint16_t)i;
            sm &= sm - 1;
          }
        }
        if (tid < ns && mylen > 0) { const int32_t w = myspb + koff; S.y.w.wstart[w] = (uint16_t)loff; S.y.w.wseg[w] = (uint16_t)tid; }
        // special-token literals in the cleaned tile (skipped when no byte can start one)
        if (S.has_sp) for (int32_t x = tid; x < tot_out; x += F3_T) {
          const uint8_t b = S.out[x];
          if ((T.sp_first[b >> 6] >> (b & 63)) & 1ULL) {
            int lo = 0, hi = ns;
            while (hi - lo > 1) { int mid = (lo + hi) >> 1; if (S.loff[mid] <= x) lo = mid; else hi = mid; }
            const int32_t end = S.loff[lo] + S.len[lo];
            for (uint32_t q = 0; q < T.n_sp; ++q) {
              const uint32_t o = T.sp_off[q], L = T.sp_off[q + 1] - o;
              if (L == 0 || T.sp_bytes[o] != b || x + (int32_t)L > end) continue;
              bool eq = true;
              for (uint32_t j = 1; j < L; ++j) if (S.out[x + j] != T.sp_bytes[o + j]) { eq = false; break; }
              if (eq) { atomicMin(&S.spmin[lo], (x << 4) | (int32_t)q); break; }
            }
          }
        }
      }
    } else {
      // ================= slow lane: sequential per-segment routines on global memory; giants come pre-cleaned =====
      const uint8_t* src = D.text + base;
      int32_t a = 0, b = 0;
      if (tid < ns) {
        a = S.rel[tid]; b = S.rel[tid + 1];
        gq = A.G.of_seg ? A.G.of_seg[s0 + tid] : -1;
        if (gq >= 0) { mylen = (int32_t)(A.G.toff[gq + 1] - A.G.toff[gq]); mychars = A.G.chars[gq]; mykeep = mylen > 0; }
        else {
          TrcCleanOut r = trc_clean_segment<false>(T, src + a, b - a, nullptr);
          mylen = (int32_t)r.bytes; mychars = (int32_t)r.chars; mykeep = r.bytes > 0;
        }
      }
      int32_t d1;
      f3_scan3(S, lane, wid, mylen, mykeep, 0, &loff, &koff, &d1);
      tot_out = S.tot[0]; tot_kept = S.tot[1];
      if (tid < F3_SEGS) { S.loff[tid] = loff; S.spmin[tid] = tid < ns ? gq : -1; S.len[tid] = mylen; }
      if (wid == 0) {
        long long gb = fuse_lookback(A.desc_bytes, tile, tot_out, lane);
        long long kb = fuse_lookback(A.desc_kept, tile, tot_kept, lane);
        if (lane == 0) { S.gbase = gb; S.kbase = kb; }
      }
      __syncthreads();
      gbase = S.gbase; kbase = S.kbase;
      if (tid < ns && mylen > 0 && gq < 0) trc_clean_segment<true>(T, src + a, b - a, D.C + gbase + loff);
      for (int i = 0; i < ns; ++i) {          // giants: the whole CTA copies their cleaned text
        const int32_t q = S.spmin[i];
        if (q < 0) continue;
        const uint8_t* gsrc = A.G.text + A.G.toff[q];
        uint8_t* gdst = D.C + gbase + S.loff[i];
        const int32_t gl = S.len[i];
        for (int32_t k = tid; k < gl; k += F3_T) gdst[k] = gsrc[k];
      }
    }
    if (tid < F3_SEGS) S.tok[tid] = 0;
    __syncthreads();
    // ---- tokens of " " + cleaned text
    int32_t mysp = -1;
    if (fast_out) {
      FUSE_PH(6);
      if (tid == 0) S.nslow = 0;
      __syncthreads();
      uint16_t* slow = S.z.slow;                 // the masks are dead: two word lists live there
      uint16_t* miss = S.z.slow + F3_WORD_CAP;
      if (tid == 0) S.nmiss = 0;
      __syncthreads();
      // pass 1 (uniform): probe the word cache, then the launch memo (L2); what neither knows is queued.  Two words
      // per thread and round, so that the two memo loads are in flight together.
      auto p1_key = [&](int32_t w, uint64_t& lo, uint64_t& hi, int& seg, int& mc, uint32_t& slot8) -> bool {
        mc = -1; seg = -1; lo = hi = 0; slot8 = 0xffffffffu;
        if (w >= tot_words) return false;
        seg = S.y.w.wseg[w];
        const int32_t st = S.y.w.wstart[w];
        const int32_t sl = S.loff[seg];
        const int32_t en = (w + 1 < tot_words && S.y.w.wseg[w + 1] == seg) ? (int32_t)S.y.w.wstart[w + 1] : sl + S.len[seg];
        const int vs = st == sl ? 1 : 0, Lw = en - st + vs;
        if (Lw > 15) return false;
        const uint32_t* p32 = reinterpret_cast<const uint32_t*>(S.out + (st & ~3));
        const unsigned sh = (unsigned)(st & 3) * 8u;
        const uint32_t w0 = p32[0], w1 = p32[1], w2 = p32[2], w3 = p32[3], w4 = p32[4];
        lo = (uint64_t)__funnelshift_r(w0, w1, sh) | ((uint64_t)__funnelshift_r(w1, w2, sh) << 32);
        hi = (uint64_t)__funnelshift_r(w2, w3, sh) | ((uint64_t)__funnelshift_r(w3, w4, sh) << 32);
        if (vs) { hi = (hi << 8) | (lo >> 56); lo = (lo << 8) | (uint64_t)' '; }
        if (Lw < 8) { lo &= (1ULL << (8 * Lw)) - 1; hi = 0; } else if (Lw == 8) hi = 0; else hi &= (1ULL << (8 * (Lw - 8))) - 1;
        if (Lw <= 7) {
          slot8 = trc_key8_slot(lo, TRC_HOT8 - 1);
          const uint64_t ce = S.hot8[slot8];
          if ((ce & 0x00ffffffffffffffULL) == lo) { mc = (int)(ce >> 56); return false; }
        }
        return A.memo != nullptr && lo != 0;
      };
      auto p1_fin = [&](int32_t w, uint64_t lo, uint64_t hi, int seg, int mc, uint32_t slot8, bool asked, const ulonglong2& e) {
        if (seg < 0) return;
        if (asked && e.x == lo && (e.y & 0x00ffffffffffffffULL) == hi && (e.y >> 56)) {
          mc = (int)(e.y >> 56) - 1;
          if (slot8 != 0xffffffffu) S.hot8[slot8] = lo | ((uint64_t)mc << 56);
        }
        if (mc < 0) miss[atomicAdd(&S.nmiss, 1)] = (uint16_t)w;
        else if (mc == F3_MEMO_SLOW) slow[atomicAdd(&S.nslow, 1)] = (uint16_t)w;
        else atomicAdd(&S.tok[seg], mc);
      };
      for (int32_t wb = tid; wb < tot_words; wb += 2 * F3_T) {
        uint64_t lo0, hi0, lo1, hi1; int sg0, sg1, mc0, mc1; uint32_t s80, s81;
        const bool ask0 = p1_key(wb, lo0, hi0, sg0, mc0, s80);
        const bool ask1 = p1_key(wb + F3_T, lo1, hi1, sg1, mc1, s81);
        ulonglong2 e0 = make_ulonglong2(0, 0), e1 = make_ulonglong2(0, 0);
        if (ask0) e0 = __ldcg(reinterpret_cast<const ulonglong2*>(A.memo) + f3_memo_slot(lo0, hi0, A.memo_mask));
        if (ask1) e1 = __ldcg(reinterpret_cast<const ulonglong2*>(A.memo) + f3_memo_slot(lo1, hi1, A.memo_mask));
        p1_fin(wb, lo0, hi0, sg0, mc0, s80, ask0, e0);
        p1_fin(wb + F3_T, lo1, hi1, sg1, mc1, s81, ask1, e1);
      }
      __syncthreads();
      FUSE_PH(11);
      // pass 2 (dense): packed-word matcher for the queued words; what it cannot decide goes to the slow list
      const int32_t nmiss = S.nmiss;
      auto match_words = [&](int32_t first, int32_t stride) {
        for (int32_t q = first; q < nmiss; q += stride) {
          const int32_t w = miss[q];
          const int seg = S.y.w.wseg[w];
          const int32_t st = S.y.w.wstart[w];
          const int32_t sl = S.loff[seg];
          const int32_t en = (w + 1 < tot_words && S.y.w.wseg[w + 1] == seg) ? (int32_t)S.y.w.wstart[w + 1] : sl + S.len[seg];
          uint32_t c = 0;
          const int vs = st == sl ? 1 : 0, n = en - st, Lw = n + vs;
          bool done = false;
          if (Lw <= 16) {
            const uint32_t* p32 = reinterpret_cast<const uint32_t*>(S.out + (st & ~3));
            const unsigned sh = (unsigned)(st & 3) * 8u;
            const uint32_t w0 = p32[0], w1 = p32[1], w2 = p32[2], w3 = p32[3], w4 = p32[4];
            uint64_t lo = (uint64_t)__funnelshift_r(w0, w1, sh) | ((uint64_t)__funnelshift_r(w1, w2, sh) << 32);
            uint64_t hi = (uint64_t)__funnelshift_r(w2, w3, sh) | ((uint64_t)__funnelshift_r(w3, w4, sh) << 32);
            if (vs) { hi = (hi << 8) | (lo >> 56); lo = (lo << 8) | (uint64_t)' '; }
            if (Lw < 8) { lo &= (1ULL << (8 * Lw)) - 1; hi = 0; } else if (Lw == 8) hi = 0; else if (Lw < 16) hi &= (1ULL << (8 * (Lw - 8))) - 1;
            done = trc_count_word16(T, F, lo, hi, Lw, &c);       // what it cannot decide needs the generic matcher: remembered too
            if (Lw <= 7) S.hot8[trc_key8_slot(lo, TRC_HOT8 - 1)] = lo | ((uint64_t)(done ? c : (uint32_t)F3_MEMO_SLOW) << 56);
            if (Lw <= 15) f3_memo_put(A, lo, hi, done ? c : (uint32_t)F3_MEMO_SLOW);
          } else done = trc_count_word_fast(T, F, S.out + st, n, vs, &c);
          if (A.phase && Lw > 15) atomicAdd(&A.phase[15], 1ULL);
          if (done) atomicAdd(&S.tok[seg], (int32_t)c);
          else slow[atomicAdd(&S.nslow, 1)] = (uint16_t)w;
        }
      };
      // Few queued words (the usual case once the memo is warm): the matcher is a latency chain on a handful of
      // lanes, so four warps run it while the other twelve resolve the look-back and copy the cleaned tile out.
      const bool split = nmiss <= 2 * F3_MATCH_T;
      if (split) {
        if (wid < F3_MATCH_T / 32) match_words(tid, F3_MATCH_T);
        else {
          if (wid == F3_MATCH_T / 32) {
            long long gb = fuse_resolve(A.desc_bytes, tile, tot_out, lane);
            long long kb = fuse_resolve(A.desc_kept, tile, tot_kept, lane);
            if (lane == 0) { S.gbase = gb; S.kbase = kb; }
          }
          asm volatile("bar.sync 1, %0;" ::"n"(F3_T - F3_MATCH_T) : "memory");
          f3_copy_out(D.C + S.gbase, S.out, tot_out, tid - F3_MATCH_T, F3_T - F3_MATCH_T);
        }
        __syncthreads();
      } else {
        match_words(tid, F3_T);
        __syncthreads();
      }
      if (A.phase && tid == 0) {               // diagnostics: how the words split over the passes
        atomicAdd(&A.phase[12], (unsigned long long)tot_words); atomicAdd(&A.phase[13], (unsigned long long)nmiss);
        atomicAdd(&A.phase[14], (unsigned long long)S.nslow);
      }
      FUSE_PH(9);
      if (!split) {
        // the global offsets are needed from here on (long-span work list, outputs): predecessors published long ago
        if (wid == 0) {
          long long gb = fuse_resolve(A.desc_bytes, tile, tot_out, lane);
          long long kb = fuse_resolve(A.desc_kept, tile, tot_kept, lane);
          if (lane == 0) { S.gbase = gb; S.kbase = kb; }
        }
        __syncthreads();
        f3_copy_out(D.C + S.gbase, S.out, tot_out, tid, F3_T);   // the cleaned tile leaves shared memory once, coalesced
      }
      gbase = S.gbase; kbase = S.kbase;
      FUSE_PH(10);
      const int32_t nslow = S.nslow;
      // words for the generic matcher: handed to k_count_slow_words (full occupancy) unless its list is full
      if (tid == 0) {
        unsigned long long b0 = nslow ? atomicAdd(A.sl_cnt, (unsigned long long)nslow) : 0ULL;
        S.slbase = ((int64_t)b0 + nslow <= A.sl_cap) ? (long long)b0 : ((int64_t)b0 < A.sl_cap ? -2LL - (long long)b0 : -1LL);
      }
      __syncthreads();
      const long long slbase = S.slbase;
      if (slbase < -1) {                       // the reservation ran over the end of the list: its part inside the list
        SlowWord e; e.pos = 0; e.len_vs = 0; e.unit = 0;   // becomes empty words, the tile counts its words inline
        for (int64_t q = -2 - slbase + tid; q < A.sl_cap; q += F3_T) A.sl[q] = e;
      }
      for (int32_t q = tid; q < nslow; q += F3_T) {
        const int32_t w = slow[q];
        const int seg = S.y.w.wseg[w];
        const int32_t st = S.y.w.wstart[w];
        const int32_t sl = S.loff[seg];
        const int32_t en = (w + 1 < tot_words && S.y.w.wseg[w + 1] == seg) ? (int32_t)S.y.w.wstart[w + 1] : sl + S.len[seg];
        const int vs = st == sl ? 1 : 0;
        const int64_t j = kbase + S.koff[seg];
        if (slbase >= 0) {
          SlowWord e; e.pos = gbase + st; e.len_vs = (en - st) | (vs << 31); e.unit = (int32_t)j;
          A.sl[slbase + q] = e;
          continue;
        }
        uint32_t c = trc_count_unit(T, F, byte_id, S.out + st, en - st, vs, [&](int64_t ls, int64_t len) -> uint32_t {
          unsigned long long slot = atomicAdd(&A.wl_cnt[0], 1ULL);
          unsigned long long off = atomicAdd(&A.wl_cnt[1], (unsigned long long)len);
          if ((int64_t)slot < A.wl_cap) { LongEntry e; e.unit = j; e.lstart = ls + (vs ? 0 : (st - sl + 1)); e.len = len; e.scratch_off = (int64_t)off; A.wl[slot] = e; }
          return 0u;
        });
        atomicAdd(&S.tok[seg], (int32_t)c);
      }
    } else if (tid < ns && mylen > 0 && gq >= 0) {
      S.tok[tid] = (int32_t)A.G.tok[gq];
      const long long v = A.G.spw[gq];
      mysp = v == 0x7fffffffffffffffLL ? -1 : (int32_t)(v & 15);
    } else if (tid < ns && mylen > 0) {
      const int64_t j = kbase + koff;
      const uint8_t* p = D.C + gbase + loff;
      uint32_t c = trc_count_unit(T, F, byte_id, p, mylen, 1, [&](int64_t ls, int64_t len) -> uint32_t {
        unsigned long long slot = atomicAdd(&A.wl_cnt[0], 1ULL);
        unsigned long long off = atomicAdd(&A.wl_cnt[1], (unsigned long long)len);
        if ((int64_t)slot < A.wl_cap) { LongEntry e; e.unit = j; e.lstart = ls; e.len = len; e.scratch_off = (int64_t)off; A.wl[slot] = e; }
        return 0u;
      });
      S.tok[tid] = (int32_t)c;
      mysp = trc_find_special(T, p, mylen);
    }
    __syncthreads();
    if (fast_out && tid < ns) { const int32_t v = S.spmin[tid]; mysp = v == 0x7fffffff ? -1 : (v & 15); }
    FUSE_PH(7);
    // ---- outputs
    if (tid < ns) {
      D.cchars[s0 + tid] = mychars;
      if (mykeep) {
        const int64_t j = kbase + koff;
        const int64_t i = s0 + tid;
        D.k2r[j] = (int32_t)i; D.coff[j] = gbase + loff; D.ctok[j] = (uint32_t)S.tok[tid]; A.seg_special[j] = mysp;
        int64_t lo = S.doc0;                  // documents are rarely shorter than a few segments: walk, then search
        for (int q = 0; q < 4 && D.doc_off[lo + 1] <= i; ++q) ++lo;
        if (D.doc_off[lo + 1] <= i) {
          int64_t hi = D.n_doc;
          while (hi - lo > 1) { int64_t mid = (lo + hi) >> 1; if (D.doc_off[mid] <= i) lo = mid; else hi = mid; }
        }
        D.sdoc[j] = (int32_t)lo;
      }
    }
    if (tid == 0 && s0 + ns == D.n) {
      D.coff[kbase + tot_kept] = gbase + tot_out;
      A.totals[0] = kbase + tot_kept; A.totals[1] = gbase + tot_out;
    }
    __syncthreads();
    FUSE_PH(8);
  }
}
