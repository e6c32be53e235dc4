// K1a + K2 fused: one pass over the raw text.  Persistent CTAs take tiles of FUSE_SEGS raw segments from
// an atomic counter.  A tile is staged in shared memory with 16-byte loads and cleaned there with the
// byte-parallel local rules of trc_clean_par.cuh (preprocessor.py:69-89): all lanes of a warp work on 32
// consecutive bytes, output positions come from warp ballots + a scan of per-strip totals.  The cleaned
// tile is tokenised in shared memory word by word (big_chunkeroosky.py:83 via the decomposition of
// thread per whitespace-delimited word, uniform work per thread) and written out once, at
// the offset obtained by a decoupled look-back over the tiles' (cleaned bytes, kept segments) aggregates.
// Replaces the count -> scan -> compact -> write -> tokenise sequence of the v1 kernels.
// Tiles that do not fit the staging buffers take the slow lane (sequential per-segment routines on global
// memory) inside the same kernel.  CUDA only; the per-unit routines it calls are host-testable.
#pragma once

#define FUSE_T 512
#define FUSE_NW (FUSE_T / 32)
#define FUSE_SEGS 256
#define FUSE_IN_CAP 18432
#define FUSE_OUT_CAP 26624
#define FUSE_WORD_CAP 4608
#define FUSE_CAND_CAP 3072
#define FUSE_HOT TRC_HOT8
#define FUSE_STRIPS (FUSE_IN_CAP / 32)

struct FuseSmem {
  union {                                     // raw tile; after the scatter: sorted word list
    uint8_t in[FUSE_IN_CAP + 32];
    struct { uint16_t wstart[FUSE_WORD_CAP]; uint16_t wseg[FUSE_WORD_CAP]; } w;
  } y;
  union { uint8_t cls[FUSE_IN_CAP]; uint8_t out[FUSE_OUT_CAP]; } x;   // classes (+segment delta); later the cleaned tile
  union { uint8_t code[FUSE_IN_CAP]; uint16_t slow[FUSE_WORD_CAP]; } z; // output codes (+segment delta); later the slow-word list
  uint64_t hot8[FUSE_HOT];
  uint32_t cand[FUSE_CAND_CAP];               // duplicate-word candidates: byte index | segment << 16
  uint16_t st_out[FUSE_STRIPS + 1], st_sp[FUSE_STRIPS + 1], st_ct[FUSE_STRIPS + 1];   // per strip: bytes, spaces, continuation bytes
  uint16_t strip_seg[FUSE_STRIPS];
  uint8_t pre_out[FUSE_SEGS + 1], pre_sp[FUSE_SEGS + 1], pre_ct[FUSE_SEGS + 1];       // strip-relative prefixes at segment starts
  uint16_t spb[FUSE_SEGS + 1], ctb[FUSE_SEGS + 1];
  int32_t rel[FUSE_SEGS + 1];
  int32_t loff[FUSE_SEGS + 1];
  int32_t fns[FUSE_SEGS], len[FUSE_SEGS], koff[FUSE_SEGS], tok[FUSE_SEGS], spmin[FUSE_SEGS];
  int32_t wsum[3][FUSE_NW];
  int32_t tot[3];
  int32_t ncand, nslow;
  unsigned int tile;
  long long gbase, kbase;
};
static_assert(sizeof(FuseSmem) <= 115200, "two CTAs per SM");

struct GiantView {                  // what the fused kernel needs about pre-processed giant segments (trc_giant.cuh)
  const int32_t* of_seg;            // per raw segment: giant index or -1; nullptr when the batch has no giants
  const int64_t* toff; const uint8_t* text; const int32_t* chars; const uint32_t* tok; const long long* spw;
};

struct SlowWord { int64_t pos; int32_t len_vs; int32_t unit; };   // byte offset in the cleaned text, length | vs << 31, kept segment

struct FuseArgs {
  unsigned long long* tile_counter;
  unsigned long long* desc_bytes;   // look-back descriptors, one per tile
  unsigned long long* desc_kept;
  int64_t* totals;                  // [0] kept segments, [1] cleaned bytes
  int64_t text_bytes;
  int32_t* seg_special;
  LongEntry* wl; unsigned long long* wl_cnt; int64_t wl_cap;
  unsigned long long* phase;        // optional: cycles per phase summed over CTAs (diagnostics, TRCHUNK_PHASES=1)
  SlowWord* sl; unsigned long long* sl_cnt; int64_t sl_cap;   // words left to the generic matcher (k_count_slow_words)
  GiantView G;
  // word memo of the launch (k_clean_tok_fused3): packed word of <= 15 bytes -> token count, write-once entries of
  // two 64-bit words {bytes 0..7, bytes 8..14 | (count+1) << 56}, claimed through memo_lock; cleared before every launch
  unsigned long long* memo; uint32_t* memo_lock; uint32_t memo_mask;
  // second level of the look-back of k_clean_tok_w32: per group of 32 tiles an arrival counter and (status, value) words
  unsigned long long* grp_cnt; unsigned long long* grp_bytes; unsigned long long* grp_kept;
  unsigned long long* wstat;        // [0] words through the word-level counter, [1] first sightings, [2] words left to the generic matcher
  const int32_t* tile_start;        // k_clean_tok_c32: first segment of every tile (+ n at the end), or null: tile t = segments [t * tsegs, ...)
  // cut mode of k_clean_tok_c32 (oversized segments cut into parts, trc_giant.cuh): per segment of the launch the caller's
  // segment it belongs to, and whether it continues one; per kept record the first special-token hit (position << 4 | k)
  const int32_t* vorig; const uint8_t* vcont; unsigned long long* special_pos;
  uint8_t* wtok;                    // cut mode: per byte of the cleaned text, tokens of the word that starts there (TrcWordSums::wtok)
};
#define FUSE_PH(i) do { if (A.phase && tid == 0) { long long now_ = clock64(); atomicAdd(&A.phase[i], (unsigned long long)(now_ - ph_t)); ph_t = now_; } } while (0)

__device__ __forceinline__ long long fuse_lookback(unsigned long long* d, unsigned tile, long long aggregate, int lane) {
  long long excl = 0;
  if (tile == 0) { if (lane == 0) st_relaxed(&d[0], (2ULL << 62) | (unsigned long long)aggregate); return 0; }
  if (lane == 0) st_relaxed(&d[tile], (1ULL << 62) | (unsigned long long)aggregate);
  long long idx = (long long)tile - 1;
  for (;;) {
    long long my = idx - lane;
    unsigned long long dv = (2ULL << 62);
    if (my >= 0) { dv = ld_relaxed(&d[my]); while ((dv >> 62) == 0) dv = ld_relaxed(&d[my]); }
    unsigned st = (unsigned)(dv >> 62);
    long long val = (long long)(dv & ((1ULL << 62) - 1));
    unsigned ball = __ballot_sync(0xffffffffu, st == 2);
    int first = ball ? __ffs(ball) - 1 : 31;
    long long c = lane <= first ? val : 0;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
    excl += c;
    if (ball) break;
    idx -= 32;
  }
  if (lane == 0) st_relaxed(&d[tile], (2ULL << 62) | (unsigned long long)(excl + aggregate));
  return excl;
}

// The same protocol in two halves: publish the aggregate as soon as it is known, resolve the exclusive
// prefix only when it is needed (by then the predecessors have long published theirs).
__device__ __forceinline__ void fuse_publish(unsigned long long* d, unsigned tile, long long aggregate) {
  st_relaxed(&d[tile], ((tile == 0 ? 2ULL : 1ULL) << 62) | (unsigned long long)aggregate);
}
__device__ __forceinline__ long long fuse_resolve(unsigned long long* d, unsigned tile, long long aggregate, int lane) {
  if (tile == 0) return 0;
  long long excl = 0;
  long long idx = (long long)tile - 1;
  for (;;) {
    long long my = idx - lane;
    unsigned long long dv = (2ULL << 62);
    if (my >= 0) { dv = ld_relaxed(&d[my]); while ((dv >> 62) == 0) dv = ld_relaxed(&d[my]); }
    unsigned st = (unsigned)(dv >> 62);
    long long val = (long long)(dv & ((1ULL << 62) - 1));
    unsigned ball = __ballot_sync(0xffffffffu, st == 2);
    int first = ball ? __ffs(ball) - 1 : 31;
    long long c = lane <= first ? val : 0;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(0xffffffffu, c, o);
    excl += c;
    if (ball) break;
    idx -= 32;
  }
  if (lane == 0) st_relaxed(&d[tile], (2ULL << 62) | (unsigned long long)(excl + aggregate));
  return excl;
}

// exclusive block scan of three int32 values; totals land in S.tot[]
__device__ __forceinline__ void fuse_scan3(FuseSmem& S, int lane, int wid, int32_t v0, int32_t v1, int32_t v2, int32_t* e0, int32_t* e1,
                                           int32_t* e2) {
  int32_t x0 = v0, x1 = v1, x2 = v2;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    int32_t y0 = __shfl_up_sync(0xffffffffu, x0, o), y1 = __shfl_up_sync(0xffffffffu, x1, o), y2 = __shfl_up_sync(0xffffffffu, x2, o);
    if (lane >= o) { x0 += y0; x1 += y1; x2 += y2; }
  }
  __syncthreads();                      // protects wsum/tot of a previous use
  if (lane == 31) { S.wsum[0][wid] = x0; S.wsum[1][wid] = x1; S.wsum[2][wid] = x2; }
  __syncthreads();
  if (wid == 0) {
    int32_t w0 = lane < FUSE_NW ? S.wsum[0][lane] : 0, w1 = lane < FUSE_NW ? S.wsum[1][lane] : 0, w2 = lane < FUSE_NW ? S.wsum[2][lane] : 0;
#pragma unroll
    for (int o = 1; o < FUSE_NW; o <<= 1) {
      int32_t y0 = __shfl_up_sync(0xffffffffu, w0, o), y1 = __shfl_up_sync(0xffffffffu, w1, o), y2 = __shfl_up_sync(0xffffffffu, w2, o);
      if (lane >= o) { w0 += y0; w1 += y1; w2 += y2; }
    }
    if (lane < FUSE_NW) { S.wsum[0][lane] = w0; S.wsum[1][lane] = w1; S.wsum[2][lane] = w2; }
    if (lane == FUSE_NW - 1) { S.tot[0] = w0; S.tot[1] = w1; S.tot[2] = w2; }
  }
  __syncthreads();
  *e0 = (wid ? S.wsum[0][wid - 1] : 0) + x0 - v0;
  *e1 = (wid ? S.wsum[1][wid - 1] : 0) + x1 - v1;
  *e2 = (wid ? S.wsum[2][wid - 1] : 0) + x2 - v2;
}

__global__ void __launch_bounds__(FUSE_T, 2) k_clean_tok_fused(TrcDev D, TrcTables T, const uint32_t* __restrict__ byte_id, FuseArgs A,
                                                               int64_t ntiles) {
  extern __shared__ __align__(16) unsigned char fuse_raw[];
  FuseSmem& S = *reinterpret_cast<FuseSmem*>(fuse_raw);
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  const unsigned lt = (1u << lane) - 1u;
  for (int k = tid; k < FUSE_HOT; k += FUSE_T) S.hot8[k] = T.hot8_src[k];
  TrcFast F; F.acls = T.acls; F.hot8 = S.hot8; F.hot8_mask = FUSE_HOT - 1;
  __syncthreads();
  long long ph_t = clock64();
  for (;;) {
    if (tid == 0) S.tile = (unsigned int)atomicAdd(A.tile_counter, 1ULL);
    __syncthreads();
    const unsigned tile = S.tile;
    if ((int64_t)tile >= ntiles) break;
    const int64_t s0 = (int64_t)tile * FUSE_SEGS;
    const int ns = (int)((D.n - s0) < FUSE_SEGS ? (D.n - s0) : FUSE_SEGS);
    const int64_t raw0 = D.seg_off[s0], raw1 = D.seg_off[s0 + ns];
    const int al = (int)(raw0 & 15);
    const int64_t base = raw0 - al;
    const int32_t span = (int32_t)((raw1 - base) < (1 << 30) ? (raw1 - base) : (1 << 30));
    const bool fast_in = (raw1 - base) <= FUSE_IN_CAP;
    for (int k = tid; k <= ns; k += FUSE_T) S.rel[k] = (int32_t)(D.seg_off[s0 + k] - base);
    __syncthreads();                          // rel[] is read across threads by both lanes
    int32_t mylen = 0, mynw = 0, mykeep = 0, mychars = 0, loff = 0, woff = 0, koff = 0;
    int32_t tot_out = 0, tot_words = 0, tot_kept = 0;
    long long gbase = 0, kbase = 0;
    bool fast_out = false;
    if (fast_in) {
      // ================= fast lane: byte-parallel clean in shared memory =================
      for (int32_t k = tid * 16; k < span; k += FUSE_T * 16) {
        if (base + k + 16 <= A.text_bytes) *reinterpret_cast<int4*>(&S.y.in[k]) = *reinterpret_cast<const int4*>(D.text + base + k);
        else for (int q = 0; q < 16 && base + k + q < A.text_bytes; ++q) S.y.in[k + q] = D.text[base + k + q];
      }
      const int nstrips = (span + 31) >> 5;
      __syncthreads();
      FUSE_PH(0);
      // segment of the first byte of every strip
      if (tid < ns) {
        int32_t a = S.rel[tid], b = S.rel[tid + 1];
        if (b > a) for (int32_t s = (a + 31) >> 5; (s << 5) < b; ++s) S.strip_seg[s] = (uint16_t)tid;
      }
      if (tid == 0) S.strip_seg[0] = 0;
      __syncthreads();
      // P1 classify; bits 3..7 of cls keep the byte's segment relative to its strip's first segment
      for (int s = wid; s < nstrips; s += FUSE_NW) {
        int32_t k = (s << 5) + lane;
        if (k < al) { S.x.cls[k] = 0; S.z.code[k] = 0; }      // bytes of the previous tile in the aligned prefix
        else if (k < span) {
          const int i0 = S.strip_seg[s];
          int i = i0;
          while (i + 1 < ns && S.rel[i + 1] <= k) ++i;
          cp_classify(T, S.y.in, S.x.cls, S.z.code, k, S.rel[i + 1]);
          const uint8_t c = S.y.in[k];
          if (c < 0x80 || c >= 0xC0) {
            int d = i - i0; if (d > 31) d = 31;
            S.x.cls[k] |= (uint8_t)(d << 3);
            if (c >= 0xC0) {
              const int len = trc_u8len(c);
              for (int q = 1; q < len && k + q < span; ++q) { int dd = i - (int)S.strip_seg[(k + q) >> 5]; if (dd > 31) dd = 31; S.x.cls[k + q] |= (uint8_t)(dd << 3); }
            }
          }
        }
      }
      __syncthreads();
      FUSE_PH(1);
      // P2 duplicate-word marks: cheap per-byte filter, candidates compacted, then handled densely
      if (tid == 0) S.ncand = 0;
      if (tid < ns) S.fns[tid] = cp_first_nonspace(S.x.cls, S.rel[tid], S.rel[tid + 1]);
      __syncthreads();
      for (int s = wid; s < nstrips; s += FUSE_NW) {
        int32_t k = (s << 5) + lane;
        bool cnd = false; int i = 0;
        if (k >= al && k < span) {
          const int d = S.x.cls[k] >> 3;
          i = S.strip_seg[s] + d;
          if (d == 31) while (i + 1 < ns && S.rel[i + 1] <= k) ++i;
          cnd = cp_dup_candidate(S.x.cls, k, S.rel[i]);
        }
        unsigned m = __ballot_sync(0xffffffffu, cnd);
        if (m) {
          int32_t b0 = 0;
          if (lane == 0) b0 = atomicAdd(&S.ncand, __popc(m));
          b0 = __shfl_sync(0xffffffffu, b0, 0);
          int32_t slot = b0 + __popc(m & lt);
          if (cnd && slot < FUSE_CAND_CAP) S.cand[slot] = (uint32_t)k | ((uint32_t)i << 16);
        }
      }
      __syncthreads();
      {
        const int32_t nc = S.ncand;
        if (nc <= FUSE_CAND_CAP) {
          for (int32_t c = tid; c < nc; c += FUSE_T) {
            uint32_t e = S.cand[c];
            int i = (int)(e >> 16);
            cp_mark_dup(S.y.in, S.x.cls, S.z.code, (int32_t)(e & 0xffffu), S.rel[i], S.rel[i + 1]);
          }
        } else {                                 // list overflow: undecimated pass (same marks)
          for (int s = wid; s < nstrips; s += FUSE_NW) {
            int32_t k = (s << 5) + lane;
            if (k >= al && k < span) {
              int i = S.strip_seg[s];
              while (i + 1 < ns && S.rel[i + 1] <= k) ++i;
              cp_mark_dup(S.y.in, S.x.cls, S.z.code, k, S.rel[i], S.rel[i + 1]);
            }
          }
        }
      }
      __syncthreads();
      FUSE_PH(2);
      // P3 output codes; per strip: output bytes, emitted spaces, emitted continuation bytes; the lane that
      // owns a segment's first byte records the strip-relative prefixes for that segment
      for (int s = wid; s < nstrips; s += FUSE_NW) {
        int32_t k = (s << 5) + lane;
        uint32_t r = 0; bool cont = false; int i = 0;
        const bool live = k >= al && k < span;
        if (live) {
          const uint8_t cl = S.x.cls[k];
          const int d = cl >> 3;
          i = S.strip_seg[s] + d;
          if (d == 31) while (i + 1 < ns && S.rel[i + 1] <= k) ++i;
          r = cp_code(S.y.in, S.x.cls, S.z.code, k, S.rel[i], S.rel[i + 1], S.fns[i]);
          cont = r == 1 && (cl & CP_CONT);
          S.z.code[k] = (uint8_t)(r | (d << 2));
        }
        const unsigned m1 = __ballot_sync(0xffffffffu, r != 0), m2 = __ballot_sync(0xffffffffu, r == 3);
        const unsigned ms = __ballot_sync(0xffffffffu, r >= 2), mc = __ballot_sync(0xffffffffu, cont);
        if (lane == 0) { S.st_out[s] = (uint16_t)(__popc(m1) + __popc(m2)); S.st_sp[s] = (uint16_t)__popc(ms); S.st_ct[s] = (uint16_t)__popc(mc); }
        if (live && S.rel[i] == k) {
          const uint8_t po = (uint8_t)(__popc(m1 & lt) + __popc(m2 & lt)), ps = (uint8_t)__popc(ms & lt), pc = (uint8_t)__popc(mc & lt);
          for (int ii = i; ii >= 0 && S.rel[ii] == k; --ii) { S.pre_out[ii] = po; S.pre_sp[ii] = ps; S.pre_ct[ii] = pc; }
        }
      }
      __syncthreads();
      if (wid == 0) {                       // exclusive scans of the three strip totals
        int32_t c0 = 0, c1 = 0, c2 = 0;
        for (int b0 = 0; b0 < nstrips; b0 += 32) {
          const bool in_r = b0 + lane < nstrips;
          int32_t v0 = in_r ? S.st_out[b0 + lane] : 0, v1 = in_r ? S.st_sp[b0 + lane] : 0, v2 = in_r ? S.st_ct[b0 + lane] : 0;
          int32_t x0 = v0, x1 = v1, x2 = v2;
#pragma unroll
          for (int o = 1; o < 32; o <<= 1) {
            int32_t y0 = __shfl_up_sync(0xffffffffu, x0, o), y1 = __shfl_up_sync(0xffffffffu, x1, o), y2 = __shfl_up_sync(0xffffffffu, x2, o);
            if (lane >= o) { x0 += y0; x1 += y1; x2 += y2; }
          }
          if (in_r) { S.st_out[b0 + lane] = (uint16_t)(c0 + x0 - v0); S.st_sp[b0 + lane] = (uint16_t)(c1 + x1 - v1); S.st_ct[b0 + lane] = (uint16_t)(c2 + x2 - v2); }
          c0 += __shfl_sync(0xffffffffu, x0, 31); c1 += __shfl_sync(0xffffffffu, x1, 31); c2 += __shfl_sync(0xffffffffu, x2, 31);
        }
        if (lane == 0) { S.st_out[nstrips] = (uint16_t)c0; S.st_sp[nstrips] = (uint16_t)c1; S.st_ct[nstrips] = (uint16_t)c2; }
      }
      __syncthreads();
      tot_out = S.st_out[nstrips];
      const int32_t tot_sp = S.st_sp[nstrips], tot_ct = S.st_ct[nstrips];
      for (int i = tid; i <= ns; i += FUSE_T) {           // prefixes at every segment start
        const int32_t k = S.rel[i];
        if (k >= span) { S.loff[i] = tot_out; S.spb[i] = (uint16_t)tot_sp; S.ctb[i] = (uint16_t)tot_ct; }
        else { const int s = k >> 5; S.loff[i] = S.st_out[s] + S.pre_out[i]; S.spb[i] = (uint16_t)(S.st_sp[s] + S.pre_sp[i]); S.ctb[i] = (uint16_t)(S.st_ct[s] + S.pre_ct[i]); }
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
      fuse_scan3(S, lane, wid, mykeep, mynw, 0, &koff, &woff, &d1);
      tot_kept = S.tot[0]; tot_words = S.tot[1];
      if (tid < ns) { S.len[tid] = mylen; S.koff[tid] = koff; S.spmin[tid] = 0x7fffffff; }
      if (wid == 0) {
        long long gb = fuse_lookback(A.desc_bytes, tile, tot_out, lane);
        long long kb = fuse_lookback(A.desc_kept, tile, tot_kept, lane);
        if (lane == 0) { S.gbase = gb; S.kbase = kb; }
      }
      __syncthreads();
      gbase = S.gbase; kbase = S.kbase;
      FUSE_PH(4);
      fast_out = tot_out <= FUSE_OUT_CAP;
      // scatter (cls is dead: out may overwrite it)
      {
        uint8_t* dst = fast_out ? S.x.out : D.C + gbase;
        for (int s = wid; s < nstrips; s += FUSE_NW) {
          int32_t k = (s << 5) + lane;
          uint32_t r = (k >= al && k < span) ? (S.z.code[k] & 3u) : 0u;
          unsigned m1 = __ballot_sync(0xffffffffu, r != 0), m2 = __ballot_sync(0xffffffffu, r == 3);
          int32_t pos = S.st_out[s] + __popc(m1 & lt) + __popc(m2 & lt);
          if (r == 1) dst[pos] = S.y.in[k];
          else if (r == 2) dst[pos] = ' ';
          else if (r == 3) { dst[pos] = ' '; dst[pos + 1] = S.y.in[k]; }
        }
      }
      __syncthreads();
      if (fast_out) {                        // the cleaned tile leaves shared memory once, coalesced
        uint8_t* dstg = D.C + gbase;
        for (int32_t k = tid; k < tot_out; k += FUSE_T) dstg[k] = S.x.out[k];
      }
      FUSE_PH(5);
      fast_out = fast_out && tot_words <= FUSE_WORD_CAP;
      if (fast_out) {
        // sorted word list (the raw tile is dead): a word starts at every emitted space and at every segment start
        for (int s = wid; s < nstrips; s += FUSE_NW) {
          int32_t k = (s << 5) + lane;
          uint32_t cd = (k >= al && k < span) ? S.z.code[k] : 0u;
          const uint32_t r = cd & 3u;
          unsigned m1 = __ballot_sync(0xffffffffu, r != 0), m2 = __ballot_sync(0xffffffffu, r == 3), ms = __ballot_sync(0xffffffffu, r >= 2);
          if (r >= 2) {
            const int d = (int)(cd >> 2);
            int i = S.strip_seg[s] + d;
            if (d == 31) while (i + 1 < ns && S.rel[i + 1] <= k) ++i;
            const int32_t pos = S.st_out[s] + __popc(m1 & lt) + __popc(m2 & lt);
            const int32_t w = S.st_sp[s] + __popc(ms & lt) + S.koff[i] + 1;
            S.y.w.wstart[w] = (uint16_t)pos; S.y.w.wseg[w] = (uint16_t)i;
          }
        }
        if (tid < ns && mylen > 0) { const int32_t w = myspb + koff; S.y.w.wstart[w] = (uint16_t)loff; S.y.w.wseg[w] = (uint16_t)tid; }
        // special-token literals in the cleaned tile
        for (int32_t x = tid; x < tot_out; x += FUSE_T) {
          const uint8_t b = S.x.out[x];
          if ((T.sp_first[b >> 6] >> (b & 63)) & 1ULL) {
            int lo = 0, hi = ns;                   // segment of output byte x: last i with loff[i] <= x
            while (hi - lo > 1) { int mid = (lo + hi) >> 1; if (S.loff[mid] <= x) lo = mid; else hi = mid; }
            const int32_t end = S.loff[lo] + S.len[lo];
            for (uint32_t q = 0; q < T.n_sp; ++q) {
              const uint32_t o = T.sp_off[q], L = T.sp_off[q + 1] - o;
              if (L == 0 || T.sp_bytes[o] != b || x + (int32_t)L > end) continue;
              bool eq = true;
              for (uint32_t j = 1; j < L; ++j) if (S.x.out[x + j] != T.sp_bytes[o + j]) { eq = false; break; }
              if (eq) { atomicMin(&S.spmin[lo], (x << 4) | (int32_t)q); break; }
            }
          }
        }
      }
    } else {
      // ================= slow lane: sequential per-segment routines on global memory =================
      const uint8_t* src = D.text + base;
      int32_t a = 0, b = 0;
      if (tid < ns) {
        a = S.rel[tid]; b = S.rel[tid + 1];
        TrcCleanOut r = trc_clean_segment<false>(T, src + a, b - a, nullptr);
        mylen = (int32_t)r.bytes; mychars = (int32_t)r.chars; mykeep = r.bytes > 0;
      }
      int32_t d1;
      fuse_scan3(S, lane, wid, mylen, mykeep, 0, &loff, &koff, &d1);
      tot_out = S.tot[0]; tot_kept = S.tot[1];
      if (wid == 0) {
        long long gb = fuse_lookback(A.desc_bytes, tile, tot_out, lane);
        long long kb = fuse_lookback(A.desc_kept, tile, tot_kept, lane);
        if (lane == 0) { S.gbase = gb; S.kbase = kb; }
      }
      __syncthreads();
      gbase = S.gbase; kbase = S.kbase;
      if (tid < ns && mylen > 0) trc_clean_segment<true>(T, src + a, b - a, D.C + gbase + loff);
    }
    if (tid < FUSE_SEGS) S.tok[tid] = 0;
    __syncthreads();
    // ---- tokens of " " + cleaned text, special-token scan
    int32_t mysp = -1;
    if (fast_out) {
      FUSE_PH(6);
      // fast pass: words made only of the common ASCII span shapes; the rest is queued and handled densely
      if (tid == 0) S.nslow = 0;
      __syncthreads();
      uint16_t* slow = S.z.slow;
      for (int32_t w = tid; w < tot_words; w += FUSE_T) {
        const int seg = S.y.w.wseg[w];
        const int32_t st = S.y.w.wstart[w];
        const int32_t sl = S.loff[seg];
        const int32_t en = (w + 1 < tot_words && S.y.w.wseg[w + 1] == seg) ? (int32_t)S.y.w.wstart[w + 1] : sl + S.len[seg];
        uint32_t c = 0;
        if (trc_count_word_fast(T, F, S.x.out + st, en - st, st == sl ? 1 : 0, &c)) atomicAdd(&S.tok[seg], (int32_t)c);
        else slow[atomicAdd(&S.nslow, 1)] = (uint16_t)w;
      }
      __syncthreads();
      const int32_t nslow = S.nslow;
      for (int32_t q = tid; q < nslow; q += FUSE_T) {
        const int32_t w = slow[q];
        const int seg = S.y.w.wseg[w];
        const int32_t st = S.y.w.wstart[w];
        const int32_t sl = S.loff[seg];
        const int32_t en = (w + 1 < tot_words && S.y.w.wseg[w + 1] == seg) ? (int32_t)S.y.w.wstart[w + 1] : sl + S.len[seg];
        const int vs = st == sl ? 1 : 0;
        const int64_t j = kbase + S.koff[seg];
        uint32_t c = trc_count_unit(T, F, byte_id, S.x.out + st, en - st, vs, [&](int64_t ls, int64_t len) -> uint32_t {
          unsigned long long slot = atomicAdd(&A.wl_cnt[0], 1ULL);
          unsigned long long off = atomicAdd(&A.wl_cnt[1], (unsigned long long)len);
          if ((int64_t)slot < A.wl_cap) { LongEntry e; e.unit = j; e.lstart = ls + (vs ? 0 : (st - sl + 1)); e.len = len; e.scratch_off = (int64_t)off; A.wl[slot] = e; }
          return 0u;
        });
        atomicAdd(&S.tok[seg], (int32_t)c);
      }
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
        int64_t lo = 0, hi = D.n_doc;
        while (hi - lo > 1) { int64_t mid = (lo + hi) >> 1; if (D.doc_off[mid] <= i) lo = mid; else hi = mid; }
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

// kdoc_off[x] = first kept segment whose document is >= x   (sdoc is non-decreasing).  One thread per kept segment:
// where the document changes between j-1 and j, every document in (sdoc[j-1], sdoc[j]] starts at j.  Streams sdoc once
// instead of a 21-step binary search per document.
__global__ void k_doc_ranges_stream(const int32_t* __restrict__ sdoc, int64_t n_kept, int64_t* kdoc_off, int64_t n_doc) {
  GSTRIDE(j, n_kept + 1) {
    const int64_t a = j == 0 ? -1 : (int64_t)sdoc[j - 1];
    const int64_t b = j == n_kept ? n_doc : (int64_t)sdoc[j];
    for (int64_t x = a + 1; x <= b; ++x) kdoc_off[x] = j;
  }
}
// (binary-search form, kept for the older fused kernels)
__global__ void k_doc_ranges(const int32_t* __restrict__ sdoc, int64_t n_kept, int64_t* kdoc_off, int64_t n_doc) {
  GSTRIDE(x, n_doc + 1) {
    int64_t lo = 0, hi = n_kept;
    while (lo < hi) { int64_t mid = (lo + hi) >> 1; if (sdoc[mid] < x) lo = mid + 1; else hi = mid; }
    kdoc_off[x] = lo;
  }
}


// Words the fused kernel could not decide with its fast paths (non-ASCII, spans that need byte-pair merging):
// one thread per word with the generic matcher, at full occupancy instead of a few lanes of a busy CTA.
__global__ void k_count_slow_words(TrcDev D, TrcTables T, const uint32_t* __restrict__ byte_id, const SlowWord* __restrict__ sl, int64_t n,
                                   LongEntry* wl, unsigned long long* wl_cnt, int64_t wl_cap) {
  const TrcFast F = trc_fast_default(T);
  GSTRIDE(q, n) {
    const SlowWord w = sl[q];
    const int vs = (int)((uint32_t)w.len_vs >> 31);
    const int64_t len = w.len_vs & 0x7fffffff;
    const int64_t j = w.unit;
    const int64_t rel = w.pos - D.coff[j];
    uint32_t c = trc_count_unit(T, F, byte_id, D.C + w.pos, len, vs, [&](int64_t ls, int64_t slen) -> uint32_t {
      unsigned long long slot = atomicAdd(&wl_cnt[0], 1ULL);
      unsigned long long off = atomicAdd(&wl_cnt[1], (unsigned long long)slen);
      if ((int64_t)slot < wl_cap) { LongEntry e; e.unit = j; e.lstart = ls + (vs ? 0 : rel + 1); e.len = slen; e.scratch_off = (int64_t)off; wl[slot] = e; }
      return 0u;
    });
    atomicAdd(&D.ctok[j], c);
  }
}
