// K1a + K2 fused, fifth generation: the CTA tile of the third generation with the per-byte work of the fourth.
// (preprocessor.py:38,69-89 + big_chunkeroosky.py:83 via DESIGN.md §3; results identical to k_clean_tok_fused3.)
//
//  * tile = up to 256 raw segments, at most 16 KB, one CTA of 512 threads, two CTAs per SM; tiles come from an atomic
//    counter; a flat decoupled look-back over ~300 tiles in flight gives the global offsets (profiles/r2g: a tile per
//    WARP put 2 400 tiles in flight and lost a third of the kernel in look-back spins and exposed L2 latency - the
//    lock-stepped warps of a CTA tile hide both);
//  * a THREAD owns one 32-byte chunk of the tile: every phase is a single pass of the CTA, every mask operation decides
//    32 bytes (trc_w32.cuh; the third generation spent the same instruction on an 8-byte cell and needed five passes);
//  * bytes are classified through a 256-entry table in shared memory (two multiply-adds per byte, eight class planes);
//  * duplicate-word candidates are decided from mask windows held in registers (one byte compare rejects most);
//  * the raw tile arrives by ONE bulk asynchronous copy (cp.async.bulk global -> shared, completion on an mbarrier)
//    issued by thread 0: no thread spends issue slots on the load.  Source and destination are this GPU's own memory;
//  * token counting: one pass probes the word cache (shared) and the launch memo (L2) with ballot-compacted miss
//    lists instead of per-word atomics; counts are bytes in shared memory and a thread per segment sums its words;
//  * buffers hold the worst case of a tile that fits (cleaned <= 1.5 x raw), so besides the word-list overflow there is
//    one other path only: the slow lane for tiles that do not fit (sequential per-segment routines; segments of more
//    than 4 KB arrive pre-cleaned from trc_giant.cuh).
#pragma once

#ifndef F5_T
#define F5_T 256                                // threads per CTA = chunks per tile
#endif
#define F5_CTAS (1024 / F5_T)                   // CTAs per SM
#define F5_NW (F5_T / 32)
#define F5_SEGS F5_T                            // a thread per segment in the per-segment phases
#define F5_IN_CAP (F5_T * 32)
#define F5_NCH (F5_IN_CAP / 32)                 // one chunk per thread
#define F5_OUT_CAP (F5_IN_CAP + F5_IN_CAP / 2)
#define F5_WORD_CAP (F5_T * 11)
#define F5_SLOW_CAP (F5_T * 2)
#define F5_NM 15
#define F5_NOLEAD 14                            // mask row: starts of segments that continue a cut oversized segment
#define F5_MATCH_T (F5_T / 4)                   // threads that run the packed-word matcher when only a few words are queued
#define F5_HOT (F5_T * 4)                       // entries of the CTA's word cache

struct F5Smem {
  union alignas(16) {                           // raw tile; after the scatter: word starts (position | first word of its segment << 15)
    uint8_t in[F5_IN_CAP + 32];
    uint16_t wstart[F5_WORD_CAP + 8];
  } y;
  alignas(16) uint8_t out[F5_OUT_CAP + 32];
  union alignas(16) {
    uint32_t m[F5_NM][F5_NCH + 2];              // mask rows (F4_* indices), chunk c at [c + 1]
    struct { uint8_t cnt[F5_WORD_CAP + 8]; uint16_t miss[F5_WORD_CAP + 8]; uint16_t slow[F5_SLOW_CAP]; } t;   // after the word starts
  } u;
  uint64_t lut[256];
  uint64_t hot8[F5_HOT];
  alignas(16) uint4 kmask[17];                  // byte masks of the first 0..16 bytes of a packed word
  int32_t rel[F5_SEGS + 1];
  int32_t xtok[F5_SEGS];
  int32_t spmin[F5_SEGS];
  uint16_t loff[F5_SEGS + 1];
  uint16_t wfirst[F5_SEGS + 1];
  uint8_t cont[F5_SEGS];                        // cut mode: the segment continues an oversized segment
  int16_t koff[F5_SEGS];                        // kept record of the segment relative to the tile's first (-1: a record of an earlier tile)
  uint16_t kr[F5_SEGS + 4];                     // by rank among the non-empty raw segments: kept segments up to it
  uint32_t wsa[F5_NW], wsb[F5_NW];              // warp totals of the block scans
  int32_t wsum[3][F5_NW];
  int32_t tot[3];
  uint32_t tot_os, tot_tb;
  int32_t nmiss, nslow, flag, sfflag;
  unsigned int tile;
  long long next_raw[2];                        // the tile's seg_off[s0], seg_off[s0 + ns], loaded once by thread 0
  long long tile_s0; int32_t tile_ns;           // its first segment and segment count
  int32_t doc0;
  long long gbase, kbase, slbase;
  alignas(8) unsigned long long mbar;
};
static_assert(sizeof(((F5Smem*)0)->u.t) <= sizeof(((F5Smem*)0)->u.m), "token lists fit the mask rows");
static_assert(sizeof(F5Smem) * F5_CTAS <= 227 * 1024 - F5_CTAS * 1024, "CTAs per SM");

__device__ __forceinline__ void f5_scan3(F5Smem& S, int lane, int wid, int32_t v0, int32_t v1, int32_t v2, int32_t* e0, int32_t* e1, int32_t* e2) {
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
    int32_t w0 = lane < F5_NW ? S.wsum[0][lane] : 0, w1 = lane < F5_NW ? S.wsum[1][lane] : 0, w2 = lane < F5_NW ? S.wsum[2][lane] : 0;
#pragma unroll
    for (int o = 1; o < F5_NW; o <<= 1) {
      int32_t y0 = __shfl_up_sync(0xffffffffu, w0, o), y1 = __shfl_up_sync(0xffffffffu, w1, o), y2 = __shfl_up_sync(0xffffffffu, w2, o);
      if (lane >= o) { w0 += y0; w1 += y1; w2 += y2; }
    }
    if (lane < F5_NW) { S.wsum[0][lane] = w0; S.wsum[1][lane] = w1; S.wsum[2][lane] = w2; }
    if (lane == F5_NW - 1) { S.tot[0] = w0; S.tot[1] = w1; S.tot[2] = w2; }
  }
  __syncthreads();
  *e0 = (wid ? S.wsum[0][wid - 1] : 0) + x0 - v0;
  *e1 = (wid ? S.wsum[1][wid - 1] : 0) + x1 - v1;
  *e2 = (wid ? S.wsum[2][wid - 1] : 0) + x2 - v2;
}

// packed key of word w of the cleaned tile: up to 16 bytes, the segment's first word with its virtual space
__device__ __forceinline__ void f5_word_key(const F5Smem& S, int32_t w, uint64_t& lo, uint64_t& hi, int& Lw, int& vs, int32_t& st, int32_t& n) {
  const uint32_t e = S.y.wstart[w];
  st = (int32_t)(e & 0x7fffu); vs = (int)(e >> 15);
  n = (int32_t)(S.y.wstart[w + 1] & 0x7fffu) - st;
  Lw = n + vs;
  lo = hi = 0;
  if (Lw > 16) return;
  const uint32_t* p32 = reinterpret_cast<const uint32_t*>(S.out + (st & ~3));
  const unsigned sh = (unsigned)(st & 3) * 8u;
  const uint32_t w0 = p32[0], w1 = p32[1], w2 = p32[2], w3 = p32[3], w4 = p32[4];
  uint32_t k0 = __funnelshift_r(w0, w1, sh), k1 = __funnelshift_r(w1, w2, sh), k2 = __funnelshift_r(w2, w3, sh), k3 = __funnelshift_r(w3, w4, sh);
  if (vs) { k3 = (k3 << 8) | (k2 >> 24); k2 = (k2 << 8) | (k1 >> 24); k1 = (k1 << 8) | (k0 >> 24); k0 = (k0 << 8) | (uint32_t)' '; }
  const uint4 km = S.kmask[Lw];                 // the first Lw bytes
  k0 &= km.x; k1 &= km.y; k2 &= km.z; k3 &= km.w;
  lo = (uint64_t)k0 | ((uint64_t)k1 << 32);
  hi = (uint64_t)k2 | ((uint64_t)k3 << 32);
}

// Launch memo, two ways: a word may sit in slot s or s ^ 1 (one 32-byte sector, both loads in flight together).  Entries
// are write-once as in f3_memo_put (claimed through memo_lock, two aligned 64-bit stores, count + 1 in the top byte).
__device__ __forceinline__ void f5_memo_put(const FuseArgs& A, uint64_t lo, uint64_t hi, uint32_t c) {
  if (!A.memo || lo == 0 || c > (uint32_t)F3_MEMO_SLOW) return;
  uint32_t slot = f3_memo_slot(lo, hi, A.memo_mask);
  if (atomicCAS(&A.memo_lock[slot], 0u, 1u) != 0u) { slot ^= 1u; if (atomicCAS(&A.memo_lock[slot], 0u, 1u) != 0u) return; }
  st_relaxed(&A.memo[2 * (size_t)slot], lo);
  st_relaxed(&A.memo[2 * (size_t)slot + 1], hi | ((uint64_t)(c + 1u) << 56));
}
__device__ __forceinline__ int f5_memo_hit(const ulonglong2& e, uint64_t lo, uint64_t hi) {
  return (e.x == lo && (e.y & 0x00ffffffffffffffULL) == hi && (e.y >> 56)) ? (int)(e.y >> 56) - 1 : -1;
}

// Decoupled look-back over BOTH descriptor rows at once (cleaned bytes, kept segments): the two loads of a round are in
// flight together, so the resolve costs the round trips of one row.  Each row keeps its own "inclusive prefix found" state.
__device__ __forceinline__ void f5_resolve2(unsigned long long* db, unsigned long long* dk, unsigned tile, long long agg_b, long long agg_k,
                                            int lane, long long* out_b, long long* out_k) {
  long long eb = 0, ek = 0;
  if (tile > 0) {
    bool done_b = false, done_k = false;
    long long idx = (long long)tile - 1;
    while (!(done_b && done_k)) {
      const long long my = idx - lane;
      unsigned long long vb = (2ULL << 62), vk = (2ULL << 62);
      if (my >= 0) {
        vb = ld_relaxed(&db[my]); vk = ld_relaxed(&dk[my]);
        while ((vb >> 62) == 0) vb = ld_relaxed(&db[my]);
        while ((vk >> 62) == 0) vk = ld_relaxed(&dk[my]);
      }
      const unsigned b2 = __ballot_sync(0xffffffffu, (vb >> 62) == 2), k2 = __ballot_sync(0xffffffffu, (vk >> 62) == 2);
      const int fb = b2 ? __ffs(b2) - 1 : 31, fk = k2 ? __ffs(k2) - 1 : 31;
      long long cb = (!done_b && lane <= fb) ? (long long)(vb & ((1ULL << 62) - 1)) : 0;
      long long ck = (!done_k && lane <= fk) ? (long long)(vk & ((1ULL << 62) - 1)) : 0;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) { cb += __shfl_xor_sync(0xffffffffu, cb, o); ck += __shfl_xor_sync(0xffffffffu, ck, o); }
      eb += cb; ek += ck;
      done_b = done_b || b2 != 0; done_k = done_k || k2 != 0;
      idx -= 32;
    }
    if (lane == 0) {
      st_relaxed(&db[tile], (2ULL << 62) | (unsigned long long)(eb + agg_b));
      st_relaxed(&dk[tile], (2ULL << 62) | (unsigned long long)(ek + agg_k));
    }
  }
  *out_b = eb; *out_k = ek;
}

__device__ __forceinline__ int f5_seg_of_word(const F5Smem& S, int ns, int32_t w) {
  int lo = 0, hi = ns;
  while (hi - lo > 1) { const int mid = (lo + hi) >> 1; if ((int32_t)S.wfirst[mid] <= w) lo = mid; else hi = mid; }
  return lo;
}

template <bool CUT>
__global__ void __launch_bounds__(F5_T, F5_CTAS) k_clean_tok_c32(TrcDev D, TrcTables T, const uint32_t* __restrict__ byte_id, FuseArgs A,
                                                           int64_t ntiles, int tsegs) {
  extern __shared__ __align__(16) unsigned char fuse_raw[];
  F5Smem& S = *reinterpret_cast<F5Smem*>(fuse_raw);
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  for (int k = tid; k < 256; k += F5_T) S.lut[k] = w32_lut_entry(T, (uint32_t)k);
  for (int k = tid; k < F5_HOT; k += F5_T) S.hot8[k] = 0ULL;
  if (tid < 17) {
    uint32_t q[4];
    for (int j = 0; j < 4; ++j) { const int r = tid - 4 * j; q[j] = r >= 4 ? 0xffffffffu : (r <= 0 ? 0u : ((1u << (8 * r)) - 1u)); }
    S.kmask[tid] = make_uint4(q[0], q[1], q[2], q[3]);
  }
  if (tid == 0) { f4_mbar_init(f4_saddr(&S.mbar), 1u); f4_fence_async(); }
  TrcFast F; F.acls = T.acls; F.hot8 = nullptr; F.hot8_mask = 0;
  W32 M;
  M.sp = S.u.m[F4_SP] + 1; M.w = S.u.m[F4_W] + 1; M.ct = S.u.m[F4_CT] + 1; M.bd = S.u.m[F4_BD] + 1; M.ep = S.u.m[F4_EP] + 1;
  M.lt = S.u.m[F4_LT] + 1; M.drop = S.u.m[F4_DROP] + 1; M.ds = S.u.m[F4_DS] + 1;
  // Cut mode (A.vcont): oversized segments arrive cut into parts in front of whitespace runs (k_cut_find); a part that
  // continues a segment keeps its leading whitespace (one space), opens no kept record and adds its bytes, tokens and
  // code points to the record of the part that started the segment - the most recent record before it.
  constexpr bool cutmode = CUT;                  // (two instantiations: the bench workloads pay nothing for the cut-mode branches)
  if (cutmode) M.nolead = S.u.m[F5_NOLEAD] + 1;
  uint32_t* const c_cp = S.u.m[F4_COPY] + 1; uint32_t* const c_se = S.u.m[F4_SPACE] + 1; uint32_t* const c_in = S.u.m[F4_INS] + 1;
  uint32_t* const p_os = S.u.m[F4_PRE_OS] + 1; uint32_t* const p_tb = S.u.m[F4_PRE_TB] + 1; uint32_t* const m_ns = S.u.m[F4_NS] + 1;
  uint32_t parity = 0;
  const uint32_t lt_mask = (1u << lane) - 1u;
  __syncthreads();
  long long ph_t = clock64();
  for (;;) {
    // (Claiming the next tile ahead - during the matcher phase, with its byte range - was measured twice: no gain on uniform
    // tiles, the other CTAs of the SM already cover the claim's round trips, and 1.3 -> 4.2 ms on config 4, where a tile id
    // held back behind slow-lane tiles stalls the look-back of every later tile.  Tiles are claimed when they are started.)
    if (tid == 0) {
      const unsigned t = (unsigned int)atomicAdd(A.tile_counter, 1ULL);
      S.tile = t;
      if ((int64_t)t < ntiles) {
        int64_t q0, qn;
        if (A.tile_start) { q0 = A.tile_start[t]; qn = A.tile_start[t + 1] - q0; }
        else { q0 = (int64_t)t * tsegs; qn = (D.n - q0) < tsegs ? (D.n - q0) : tsegs; }
        S.next_raw[0] = D.seg_off[q0]; S.next_raw[1] = D.seg_off[q0 + qn];
        S.tile_s0 = q0; S.tile_ns = (int32_t)qn;
      }
    }
    __syncthreads();
    const unsigned tile = S.tile;
    if ((int64_t)tile >= ntiles) break;
    const int64_t s0 = S.tile_s0;
    const int ns = S.tile_ns;
    const int64_t raw0 = S.next_raw[0], raw1 = S.next_raw[1];
    const int al = (int)(raw0 & 15);
    const int64_t base = raw0 - al;
    const bool fast_in = (raw1 - base) <= F5_IN_CAP;
    const int32_t span = fast_in ? (int32_t)(raw1 - base) : 0;
    // the raw tile: one bulk copy of whole 16-byte units inside the text buffer; the buffer's last bytes by hand
    uint32_t n16 = 0;
    if (fast_in) {
      int64_t want = ((int64_t)span + 15) & ~(int64_t)15;
      const int64_t room = (A.text_bytes - base) & ~(int64_t)15;
      if (want > room) want = room;
      n16 = want > 0 ? (uint32_t)want : 0u;
      if (tid == 0 && n16) {
        f4_fence_async();
        f4_mbar_expect(f4_saddr(&S.mbar), n16);
        f4_bulk_g2s(f4_saddr(S.y.in), D.text + base, n16, f4_saddr(&S.mbar));
      }
    }
    if (wid == F5_NW - 1) {                   // document of the first segment: one 32-ary search per tile, under the tile load
      const int64_t s0o = cutmode ? (int64_t)A.vorig[s0] : s0;       // (doc_off counts the caller's segments)
      int64_t lo = 0, hi = D.n_doc;
      while (hi - lo > 1) {
        const int64_t step = (hi - lo + 31) >> 5;
        const int64_t pv = lo + step * (lane + 1);
        const bool le = pv < hi && D.doc_off[pv] <= s0o;
        const int cnt = __popc(__ballot_sync(0xffffffffu, le));
        const int64_t nhi = lo + step * (cnt + 1);
        lo += step * cnt;
        if (nhi < hi) hi = nhi;
      }
      if (lane == 0) S.doc0 = (int32_t)lo;
    }
    for (int k = tid; k <= ns; k += F5_T) S.rel[k] = (int32_t)((D.seg_off[s0 + k] - base) < (1 << 30) ? (D.seg_off[s0 + k] - base) : (1 << 30));
    int32_t mylen = 0, mynw = 0, mykeep = 0, mychars = 0, loff = 0, woff = 0, koff = 0, mytok = 0, mysp = -1;
    const int mycont = (cutmode && tid < ns) ? (int)A.vcont[s0 + tid] : 0;
    int32_t tot_out = 0, tot_words = 0, tot_kept = 0;
    long long gbase = 0, kbase = 0;
    bool fast_out = false;
    int32_t gq = -1;
    if (fast_in) {
      const int nch = (span + 31) >> 5;
      const int c = tid;                                               // this thread's chunk
      // rows built with atomic ORs, "nothing there" on both sides of the tile
      S.u.m[F4_BD][tid] = 0u; S.u.m[F4_DROP][tid] = 0u; S.u.m[F4_DS][tid] = 0u;
      if (tid < 2) { S.u.m[F4_BD][F5_NCH + tid] = 0u; S.u.m[F4_DROP][F5_NCH + tid] = 0u; S.u.m[F4_DS][F5_NCH + tid] = 0u; }
      if (cutmode) { S.u.m[F5_NOLEAD][tid] = 0u; if (tid < 2) S.u.m[F5_NOLEAD][F5_NCH + tid] = 0u; }
      if (tid < F5_NM) { S.u.m[tid][0] = 0u; S.u.m[tid][nch + 1] = 0u; }
      if (tid < F5_SEGS) { S.xtok[tid] = 0; S.spmin[tid] = 0x7fffffff; }
      if (tid == 0) S.sfflag = 0;
      if (n16) { f4_mbar_wait(f4_saddr(&S.mbar), parity); parity ^= 1u; }
      for (int32_t k = (int32_t)n16 + tid; k < span; k += F5_T) S.y.in[k] = D.text[base + k];
      __syncthreads();
      FUSE_PH(0);
      if (tid < ns) { const int32_t a = S.rel[tid], b = S.rel[tid + 1]; if (b > a) { w32_or_bit(M.bd, a); if (mycont) w32_or_bit(S.u.m[F5_NOLEAD] + 1, a); } }
      // ---- P1: classes of this thread's chunk (lead bytes of multi-byte characters are parked in the COPY row)
      uint32_t lead = 0, sf = 0;
      const uint32_t valid = w32_valid(c, al, span);
      if (c < nch) {
        const uint4 q0 = *reinterpret_cast<const uint4*>(S.y.in + 32 * c), q1 = *reinterpret_cast<const uint4*>(S.y.in + 32 * c + 16);
        const uint32_t x[8] = {q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, q1.z, q1.w};
        const W32Cls cl = w32_classify(S.lut, x);
        lead = cl.na & ~cl.ct & valid;
        M.sp[c] = cl.sp & valid; M.w[c] = cl.w & valid; M.ct[c] = cl.ct & valid; M.ep[c] = cl.ep & valid; M.lt[c] = cl.lt & valid;
        m_ns[c] = (cl.ns | lead) & valid;
        sf = cl.sf & valid;
      }
      if (sf) S.sfflag = 1;
      const int any_lead = __syncthreads_or(lead != 0);
      if (any_lead) {
        if (lead) w32_fix_multibyte(T, S.y.in, M, c, lead, span);
        __syncthreads();
      }
      const bool has_sp = S.sfflag != 0 && T.n_sp > 0;
      FUSE_PH(1);
      // ---- P2: duplicate words
      if (c < nch) {
        uint32_t generic = w32_dup_chunk_simple(S.y.in, M, c, valid, span);
        while (generic) {
          const int bit = __ffs(generic) - 1;
          generic &= generic - 1;
          const int32_t k = (c << 5) + bit;
          if (w32_dup_quick_reject(S.y.in, M, k, span)) continue;
          int lo = 0, hi = ns;
          while (hi - lo > 1) { const int mid = (lo + hi) >> 1; if (S.rel[mid] <= k) lo = mid; else hi = mid; }
          w32_mark_dup(S.y.in, M, k, S.rel[lo], S.rel[lo + 1]);
        }
      }
      __syncthreads();
      FUSE_PH(2);
      // ---- P3: output codes and their prefixes (block scan of two packed words)
      W32Codes r; r.copy = r.space = r.ins = 0;
      {
        uint32_t ctm = 0, bd = 0;
        if (c < nch) { r = w32_chunk_codes(S.y.in, M, c, al, span); ctm = M.ct[c]; bd = M.bd[c]; }
        const uint32_t s = __popc(r.space) + __popc(r.ins), o = __popc(r.copy) + s;
        const uint32_t va = o | (s << 16), vb = (uint32_t)__popc(r.copy & ctm) | ((uint32_t)__popc(bd) << 16);
        uint32_t xa = va, xb = vb;
#pragma unroll
        for (int q = 1; q < 32; q <<= 1) {
          const uint32_t ya = __shfl_up_sync(0xffffffffu, xa, q), yb = __shfl_up_sync(0xffffffffu, xb, q);
          if (lane >= q) { xa += ya; xb += yb; }
        }
        if (lane == 31) { S.wsa[wid] = xa; S.wsb[wid] = xb; }
        __syncthreads();
        if (wid == 0) {
          uint32_t ta = lane < F5_NW ? S.wsa[lane] : 0u, tb = lane < F5_NW ? S.wsb[lane] : 0u;
          const uint32_t ta0 = ta, tb0 = tb;
#pragma unroll
          for (int q = 1; q < F5_NW; q <<= 1) {
            const uint32_t ya = __shfl_up_sync(0xffffffffu, ta, q), yb = __shfl_up_sync(0xffffffffu, tb, q);
            if (lane >= q) { ta += ya; tb += yb; }
          }
          if (lane < F5_NW) { S.wsa[lane] = ta - ta0; S.wsb[lane] = tb - tb0; }
          if (lane == F5_NW - 1) { S.tot_os = ta; S.tot_tb = tb; }
        }
        __syncthreads();
        if (c < nch) { c_cp[c] = r.copy; c_se[c] = r.space; c_in[c] = r.ins; p_os[c] = S.wsa[wid] + xa - va; p_tb[c] = S.wsb[wid] + xb - vb; }
      }
      __syncthreads();
      FUSE_PH(3);
      // ---- P4: per segment (thread i <-> segment i): offsets, lengths, words, code points
      tot_out = (int32_t)(S.tot_os & 0xffffu);
      const int32_t tot_sp = (int32_t)(S.tot_os >> 16), tot_ct = (int32_t)(S.tot_tb & 0xffffu);
      bool nonempty = false;
      int32_t br = 0;
      if (tid < ns) {
        int32_t pv[2][3];
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          const int32_t k = S.rel[tid + e];
          pv[e][0] = tot_out; pv[e][1] = tot_sp; pv[e][2] = tot_ct;
          if (k < span) {
            const int cc = k >> 5;
            const uint32_t below = (1u << (k & 31)) - 1u;
            const uint32_t cp = c_cp[cc], se = c_se[cc], in = c_in[cc];
            const uint32_t os = p_os[cc], tb = p_tb[cc];
            pv[e][0] = (int32_t)(os & 0xffffu) + __popc((cp | se) & below) + __popc(in & below);
            pv[e][1] = (int32_t)(os >> 16) + __popc((se | in) & below);
            pv[e][2] = (int32_t)(tb & 0xffffu) + __popc(cp & M.ct[cc] & below);
            if (e == 0) br = (int32_t)(tb >> 16) + __popc(M.bd[cc] & below);
          }
        }
        loff = pv[0][0]; mylen = pv[1][0] - pv[0][0]; mykeep = mylen > 0 && !mycont;       // a record, and a first word with a virtual space
        mynw = mylen > 0 ? pv[1][1] - pv[0][1] + (mycont ? 0 : 1) : 0;
        mychars = mylen - (pv[1][2] - pv[0][2]);
        nonempty = S.rel[tid + 1] > S.rel[tid];
      }
      int32_t d1;
      f5_scan3(S, lane, wid, mykeep, mynw, 0, &koff, &woff, &d1);
      tot_kept = S.tot[0]; tot_words = S.tot[1];
      if (tid == 0) { fuse_publish(A.desc_bytes, tile, tot_out); fuse_publish(A.desc_kept, tile, tot_kept); }
      fast_out = tot_words <= F5_WORD_CAP;
      if (tid < ns) { S.loff[tid] = (uint16_t)loff; S.wfirst[tid] = (uint16_t)woff; S.koff[tid] = (int16_t)(koff - mycont); if (cutmode) S.cont[tid] = (uint8_t)mycont; }
      if (tid == 0) { S.loff[ns] = (uint16_t)tot_out; S.wfirst[ns] = (uint16_t)tot_words; }
      if (nonempty) S.kr[br] = (uint16_t)(koff + (mykeep ? 1 : 0));
      FUSE_PH(4);
      // ---- P5a: scatter into the cleaned tile
      if (c < nch && (r.copy | r.space)) {
        const uint32_t cp = r.copy, se = r.space, in = r.ins;
        int32_t pos = (int32_t)(p_os[c] & 0xffffu);
        const uint4 q0 = *reinterpret_cast<const uint4*>(S.y.in + 32 * c), q1 = *reinterpret_cast<const uint4*>(S.y.in + 32 * c + 16);
        const uint32_t x[8] = {q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, q1.z, q1.w};
        const uint32_t odd = se & m_ns[c];            // emitted spaces whose source byte is not U+0020
        if (in == 0u && (cp | se) == 0xffffffffu && !odd) f4_store32(S.out + pos, x);
        else {
          // Four bytes at a time.  Groups that are copied as they are go first, in chunk order (the lanes of a warp agree
          // on most of them); the others - a dropped byte, an inserted or replaced space - are left for a second loop in
          // which every lane walks ITS OWN list of such groups, so the lanes run the byte-wise code together instead of
          // once per group index (profiles/r2s: that code ran with 2.4 of 32 lanes).
          uint8_t* dst = S.out;
          const uint32_t any = cp | se, bytewise = in | odd;
          const int32_t pos0 = pos;
          uint32_t nt = 0;
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const uint32_t a4 = (any >> (4 * j)) & 15u;
            if (!a4) continue;
            if (a4 == 15u && !((bytewise >> (4 * j)) & 15u)) {
              const uint32_t xv = x[j];
              dst[pos] = (uint8_t)xv; dst[pos + 1] = (uint8_t)(xv >> 8); dst[pos + 2] = (uint8_t)(xv >> 16); dst[pos + 3] = (uint8_t)(xv >> 24);
              pos += 4;
            } else { nt |= 1u << j; pos += __popc(a4) + __popc((in >> (4 * j)) & 15u); }
          }
          while (nt) {
            const int j = (int)trc_ctz32(nt);
            nt &= nt - 1u;
            const uint32_t below = (1u << (4 * j)) - 1u;
            int32_t p = pos0 + __popc(any & below) + __popc(in & below);
            const uint32_t xv = *reinterpret_cast<const uint32_t*>(S.y.in + 32 * c + 4 * j);
            const uint32_t c4 = (cp >> (4 * j)) & 15u, s4 = (se >> (4 * j)) & 15u, i4 = (in >> (4 * j)) & 15u;
#pragma unroll
            for (int b = 0; b < 4; ++b) {
              if ((i4 >> b) & 1u) dst[p++] = ' ';
              if (((c4 | s4) >> b) & 1u) dst[p++] = ((c4 >> b) & 1u) ? (uint8_t)(xv >> (8 * b)) : (uint8_t)' ';
            }
          }
        }
      }
      __syncthreads();                                // the raw tile is dead: word starts take its place
      FUSE_PH(5);
      if (fast_out) {
        // ---- P5b: a word starts at every emitted space and at every kept segment's first byte
        if (c < nch) {
          const uint32_t cp = r.copy, se = r.space, in = r.ins;
          uint32_t sm = se | in;
          if (sm) {
            const uint32_t os = p_os[c];
            const int32_t pos = (int32_t)(os & 0xffffu), sbase = (int32_t)(os >> 16);
            const uint32_t bdm = M.bd[c];
            const int32_t brb = (int32_t)(p_tb[c] >> 16);
            while (sm) {
              const int bit = __ffs(sm) - 1;
              sm &= sm - 1;
              const uint32_t below = (1u << bit) - 1u;
              const int32_t wpos = pos + __popc((cp | se) & below) + __popc(in & below);
              const int32_t rank = brb + __popc(bdm & (below | (1u << bit))) - 1;     // segment whose start is at or before this byte
              const int32_t w = sbase + __popc((se | in) & below) + S.kr[rank];
              S.y.wstart[w] = (uint16_t)wpos;
            }
          }
        }
        if (tid < ns && mykeep) S.y.wstart[woff] = (uint16_t)(loff | 0x8000);
        if (tid == 0) { S.y.wstart[tot_words] = (uint16_t)tot_out; S.nmiss = 0; S.nslow = 0; }
        // special-token literals in the cleaned tile (skipped when no byte can start one)
        if (has_sp) for (int32_t x = tid; x < tot_out; x += F5_T) {
          const uint8_t b = S.out[x];
          if ((T.sp_first[b >> 6] >> (b & 63)) & 1ULL) {
            int lo = 0, hi = ns;
            while (hi - lo > 1) { int mid = (lo + hi) >> 1; if ((int32_t)S.loff[mid] <= x) lo = mid; else hi = mid; }
            const int32_t end = (int32_t)S.loff[lo + 1];
            for (uint32_t q = 0; q < T.n_sp; ++q) {
              const uint32_t o = T.sp_off[q], L = T.sp_off[q + 1] - o;
              if (L == 0 || T.sp_bytes[o] != b || x + (int32_t)L > end) continue;
              bool eq = true;
              for (uint32_t j = 1; j < L; ++j) if (S.out[x + j] != T.sp_bytes[o + j]) { eq = false; break; }
              if (eq) { atomicMin(&S.spmin[lo], (x << 4) | (int32_t)q); break; }
            }
          }
        }
        __syncthreads();                              // the mask rows are dead from here on: counts and lists live there
        FUSE_PH(6);
        // ---- P7: tokens, word by word
        uint8_t* const cnt = S.u.t.cnt; uint16_t* const miss = S.u.t.miss; uint16_t* const slow = S.u.t.slow;
        // one pass: the CTA's word cache (shared memory), then the launch memo (L2, two ways in one 32-byte sector).  (Queueing
        // the memo probes so that they go out in one round was slower: the second key build costs more than the latency,
        // which 32 resident warps already hide.)
        for (int32_t wb = 0; wb < tot_words; wb += F5_T) {
          const int32_t w = wb + tid;
          int mc = -2;                                 // -2 no word, -1 unknown, >= 0 count, F3_MEMO_SLOW: generic matcher
          if (w < tot_words) {
            uint64_t lo, hi; int Lw, vs; int32_t st, n;
            f5_word_key(S, w, lo, hi, Lw, vs, st, n);
            mc = -1;
            if (Lw <= 15) {
              uint32_t slot8 = 0xffffffffu;
              if (Lw <= 7) {
                slot8 = trc_key8_slot(lo, F5_HOT - 1);
                const uint64_t ce = S.hot8[slot8];
                if ((ce & 0x00ffffffffffffffULL) == lo) mc = (int)(ce >> 56);
              }
              if (mc < 0 && A.memo != nullptr && lo != 0) {
                const uint32_t ms = f3_memo_slot(lo, hi, A.memo_mask);
                const ulonglong2 ea = __ldcg(reinterpret_cast<const ulonglong2*>(A.memo) + ms);
                const ulonglong2 eb = __ldcg(reinterpret_cast<const ulonglong2*>(A.memo) + (ms ^ 1u));
                int h = f5_memo_hit(ea, lo, hi);
                if (h < 0) h = f5_memo_hit(eb, lo, hi);
                if (h >= 0) { mc = h; if (slot8 != 0xffffffffu) S.hot8[slot8] = lo | ((uint64_t)h << 56); }
              }
            }
            cnt[w] = (uint8_t)((mc >= 0 && mc != F3_MEMO_SLOW) ? mc : 0);
          }
          const uint32_t bm = __ballot_sync(0xffffffffu, mc == -1), bs = __ballot_sync(0xffffffffu, mc == F3_MEMO_SLOW);
          int mb = 0, sb = 0;
          if (lane == 0) { if (bm) mb = atomicAdd(&S.nmiss, __popc(bm)); if (bs) sb = atomicAdd(&S.nslow, __popc(bs)); }
          mb = __shfl_sync(0xffffffffu, mb, 0); sb = __shfl_sync(0xffffffffu, sb, 0);
          if (mc == -1) miss[mb + __popc(bm & lt_mask)] = (uint16_t)w;
          if (mc == F3_MEMO_SLOW) { const int q = sb + __popc(bs & lt_mask); if (q < F5_SLOW_CAP) slow[q] = (uint16_t)w; else cnt[w] = 255; }
        }
        __syncthreads();
        FUSE_PH(11);
        // first sightings: the packed-word matcher (regex alternatives 1-4 on bit masks, exact-key vocabulary sets)
        const int32_t nmiss = S.nmiss;
        auto match_words = [&](int32_t first, int32_t stride) {
          for (int32_t q = first; q < nmiss; q += stride) {
            const int32_t w = miss[q];
            uint64_t lo, hi; int Lw, vs; int32_t st, n;
            f5_word_key(S, w, lo, hi, Lw, vs, st, n);
            uint32_t cc = 0;
            bool done;
            if (Lw <= 16) {
              done = trc_count_word16(T, F, lo, hi, Lw, &cc);
              if (Lw <= 7) S.hot8[trc_key8_slot(lo, F5_HOT - 1)] = lo | ((uint64_t)(done ? cc : (uint32_t)F3_MEMO_SLOW) << 56);
              if (Lw <= 15) f5_memo_put(A, lo, hi, done ? cc : (uint32_t)F3_MEMO_SLOW);
            } else done = trc_count_word_fast(T, F, S.out + st, n, vs, &cc) && cc < 250u;
            if (done) cnt[w] = (uint8_t)cc;
            else { const int qq = atomicAdd(&S.nslow, 1); if (qq < F5_SLOW_CAP) slow[qq] = (uint16_t)w; else cnt[w] = 255; }
          }
        };
        // Few queued words (the usual case once the memo is warm): the matcher is a latency chain on a handful of lanes,
        // so four warps run it while the other twelve resolve the look-back and copy the cleaned tile out.
        const bool split = nmiss <= 2 * F5_MATCH_T;
        if (split) {
          if (wid < F5_MATCH_T / 32) match_words(tid, F5_MATCH_T);
          else {
            if (wid == F5_MATCH_T / 32) {
              long long gb, kb;
              f5_resolve2(A.desc_bytes, A.desc_kept, tile, tot_out, tot_kept, lane, &gb, &kb);
              if (lane == 0) { S.gbase = gb; S.kbase = kb; }
            }
            asm volatile("bar.sync 1, %0;" ::"n"(F5_T - F5_MATCH_T) : "memory");
            f3_copy_out(D.C + S.gbase, S.out, tot_out, tid - F5_MATCH_T, F5_T - F5_MATCH_T);
          }
          __syncthreads();
        } else {
          match_words(tid, F5_T);
          __syncthreads();
          if (wid == 0) {
            long long gb, kb;
            f5_resolve2(A.desc_bytes, A.desc_kept, tile, tot_out, tot_kept, lane, &gb, &kb);
            if (lane == 0) { S.gbase = gb; S.kbase = kb; }
          }
          __syncthreads();
          f3_copy_out(D.C + S.gbase, S.out, tot_out, tid, F5_T);
        }
        gbase = S.gbase; kbase = S.kbase;
        FUSE_PH(9);
        if (tid == 0) { atomicAdd(&A.wstat[0], (unsigned long long)tot_words); atomicAdd(&A.wstat[1], (unsigned long long)nmiss); atomicAdd(&A.wstat[2], (unsigned long long)S.nslow); }
        // words for the generic matcher: handed to k_count_slow_words (full occupancy) unless its list is full
        const int32_t nslow = S.nslow, nl = nslow < F5_SLOW_CAP ? nslow : F5_SLOW_CAP;
        if (tid == 0) {
          unsigned long long b0 = nl ? atomicAdd(A.sl_cnt, (unsigned long long)nl) : 0ULL;
          S.slbase = ((int64_t)b0 + nl <= A.sl_cap) ? (long long)b0 : ((int64_t)b0 < A.sl_cap ? -2LL - (long long)b0 : -1LL);
          S.flag = (nslow > nl || ((int64_t)b0 + nl > A.sl_cap)) ? 1 : 0;
        }
        __syncthreads();
        const long long slbase = S.slbase;
        if (slbase < -1) {
          SlowWord e; e.pos = 0; e.len_vs = 0; e.unit = 0;
          for (int64_t q = -2 - slbase + tid; q < A.sl_cap; q += F5_T) A.sl[q] = e;
        }
        for (int32_t q = tid; q < nl; q += F5_T) {
          const int32_t w = slow[q];
          if (slbase >= 0) {
            const uint32_t e = S.y.wstart[w];
            const int32_t st = (int32_t)(e & 0x7fffu); const int vs = (int)(e >> 15);
            const int32_t n = (int32_t)(S.y.wstart[w + 1] & 0x7fffu) - st;
            const int seg = f5_seg_of_word(S, ns, w);
            SlowWord sw; sw.pos = gbase + st; sw.len_vs = n | (vs << 31); sw.unit = (int32_t)(kbase + S.koff[seg]);
            A.sl[slbase + q] = sw;
          } else cnt[w] = 255;
        }
        if (cutmode && A.wtok) {                       // per-word counts for the sentence stage (words on the slow list: their kernels)
          __syncthreads();
          for (int32_t w = tid; w < tot_words; w += F5_T) if (cnt[w] != 255) A.wtok[gbase + (int32_t)(S.y.wstart[w] & 0x7fffu)] = cnt[w];
        }
        if (S.flag) {                                  // rare: words that fit neither list are counted here
          __syncthreads();
          for (int32_t w = tid; w < tot_words; w += F5_T) {
            if (cnt[w] != 255) continue;
            const uint32_t e = S.y.wstart[w];
            const int32_t st = (int32_t)(e & 0x7fffu); const int vs = (int)(e >> 15);
            const int32_t n = (int32_t)(S.y.wstart[w + 1] & 0x7fffu) - st;
            const int seg = f5_seg_of_word(S, ns, w);
            const int32_t sl = (int32_t)S.loff[seg];
            const int64_t j = kbase + S.koff[seg];
            bool deferred = false;
            const uint32_t cc = trc_count_unit(T, F, byte_id, S.out + st, n, vs, [&](int64_t ls, int64_t len) -> uint32_t {
              deferred = true;
              const unsigned long long slot = atomicAdd(&A.wl_cnt[0], 1ULL);
              const unsigned long long off = atomicAdd(&A.wl_cnt[1], (unsigned long long)len);
              if ((int64_t)slot < A.wl_cap) {
                LongEntry le; le.unit = j; le.len = len; le.scratch_off = (int64_t)off;
                le.lstart = (cutmode && S.cont[seg]) ? -1 - (gbase + st + ls) : ls + (vs ? 0 : (st - sl + 1));     // (a part: absolute position in the cleaned text)
                A.wl[slot] = le;
              }
              return 0u;
            });
            cnt[w] = 0;
            if (cutmode && A.wtok) A.wtok[gbase + st] = (uint8_t)((deferred || cc >= 255u) ? 255u : cc);
            atomicAdd(&S.xtok[seg], (int32_t)cc);
          }
        }
        __syncthreads();
        if (tid < ns && mylen > 0) {                   // per segment: sum of its words
          int32_t t = S.xtok[tid];
          const int32_t w1 = woff + mynw;
          for (int32_t w = woff; w < w1; ++w) t += cnt[w];
          mytok = t;
          const int32_t v = S.spmin[tid];
          mysp = v == 0x7fffffff ? -1 : (v & 15);
        }
        FUSE_PH(10);
      } else {
        // too many words for the list: write the tile out and count per segment
        if (cutmode && tid == 0) atomicAdd(&A.wstat[4], 1ULL);          // (no per-word counts from this tile: the sentence stage counts itself)
        if (wid == 0) {
          long long gb, kb;
          f5_resolve2(A.desc_bytes, A.desc_kept, tile, tot_out, tot_kept, lane, &gb, &kb);
          if (lane == 0) { S.gbase = gb; S.kbase = kb; }
        }
        __syncthreads();
        gbase = S.gbase; kbase = S.kbase;
        f3_copy_out(D.C + gbase, S.out, tot_out, tid, F5_T);
        if (tid < ns && mylen > 0) {
          const int64_t j = kbase + koff - mycont;
          const uint8_t* p = S.out + loff;
          const int32_t len0 = mylen;
          const uint32_t cc = trc_count_unit(T, F, byte_id, p, len0, mycont ? 0 : 1, [&](int64_t ls, int64_t len) -> uint32_t {
            const unsigned long long slot = atomicAdd(&A.wl_cnt[0], 1ULL);
            const unsigned long long off = atomicAdd(&A.wl_cnt[1], (unsigned long long)len);
            if ((int64_t)slot < A.wl_cap) { LongEntry le; le.unit = j; le.lstart = mycont ? -1 - (gbase + loff + ls) : ls; le.len = len; le.scratch_off = (int64_t)off; A.wl[slot] = le; }
            return 0u;
          });
          mytok = (int32_t)cc;
          mysp = trc_find_special(T, p, len0);
        }
      }
    } else {
      // ================= slow lane: sequential per-segment routines on global memory; giants come pre-cleaned =====
      // (never in cut mode: its tiles always fit and no oversized segment is left; the host checks this counter)
      if (cutmode && tid == 0) atomicAdd(&A.wstat[3], 1ULL);
      const uint8_t* src = D.text + base;
      int64_t a = 0, b = 0;
      if (tid < ns) {
        a = D.seg_off[s0 + tid] - base; b = D.seg_off[s0 + tid + 1] - base;
        gq = A.G.of_seg ? A.G.of_seg[s0 + tid] : -1;
        if (gq >= 0) { mylen = (int32_t)(A.G.toff[gq + 1] - A.G.toff[gq]); mychars = A.G.chars[gq]; mykeep = mylen > 0; }
        else {
          TrcCleanOut rr = trc_clean_segment<false>(T, src + a, b - a, nullptr);
          mylen = (int32_t)rr.bytes; mychars = (int32_t)rr.chars; mykeep = rr.bytes > 0;
        }
      }
      int32_t d1;
      f5_scan3(S, lane, wid, mylen, mykeep, 0, &loff, &koff, &d1);
      tot_out = S.tot[0]; tot_kept = S.tot[1];
      if (wid == 0) {                           // both aggregates go out before either prefix is looked up
        if (lane == 0) { fuse_publish(A.desc_bytes, tile, tot_out); fuse_publish(A.desc_kept, tile, tot_kept); }
        __syncwarp();
        long long gb, kb;
        f5_resolve2(A.desc_bytes, A.desc_kept, tile, tot_out, tot_kept, lane, &gb, &kb);
        if (lane == 0) { S.gbase = gb; S.kbase = kb; }
      }
      __syncthreads();
      gbase = S.gbase; kbase = S.kbase;
      if (tid < ns && mylen > 0 && gq < 0) trc_clean_segment<true>(T, src + a, b - a, D.C + gbase + loff);
      if (tid < F5_SEGS) { S.xtok[tid] = loff; S.spmin[tid] = tid < ns ? gq : -1; }
      __syncthreads();
      for (int i = 0; i < ns; ++i) {          // giants: the whole CTA copies their cleaned text
        const int32_t q = S.spmin[i];
        if (q < 0) continue;
        const uint8_t* gsrc = A.G.text + A.G.toff[q];
        uint8_t* gdst = D.C + gbase + S.xtok[i];
        const int64_t gl = A.G.toff[q + 1] - A.G.toff[q];
        for (int64_t k = tid; k < gl; k += F5_T) gdst[k] = gsrc[k];
      }
      __syncthreads();
      if (tid < ns && mylen > 0 && gq >= 0) {
        mytok = (int32_t)A.G.tok[gq];
        const long long v = A.G.spw[gq];
        mysp = v == 0x7fffffffffffffffLL ? -1 : (int32_t)(v & 15);
      } else if (tid < ns && mylen > 0) {
        const int64_t j = kbase + koff;
        const uint8_t* p = D.C + gbase + loff;
        const int32_t len0 = mylen;
        const uint32_t cc = trc_count_unit(T, F, byte_id, p, len0, 1, [&](int64_t ls, int64_t len) -> uint32_t {
          const unsigned long long slot = atomicAdd(&A.wl_cnt[0], 1ULL);
          const unsigned long long off = atomicAdd(&A.wl_cnt[1], (unsigned long long)len);
          if ((int64_t)slot < A.wl_cap) { LongEntry le; le.unit = j; le.lstart = ls; le.len = len; le.scratch_off = (int64_t)off; A.wl[slot] = le; }
          return 0u;
        });
        mytok = (int32_t)cc;
        mysp = trc_find_special(T, p, len0);
      }
    }
    FUSE_PH(7);
    // ---- outputs
    if (tid < ns && cutmode) {                 // counts are sums over the parts of a segment: everything was zeroed before the launch
      const int64_t io = A.vorig[s0 + tid];
      if (mychars) atomicAdd(&D.cchars[io], mychars);
      if (mylen > 0) {
        const int64_t jo = kbase + koff - mycont;
        if (mytok) atomicAdd(&D.ctok[jo], (uint32_t)mytok);
        if (mysp >= 0) atomicMin(&A.special_pos[jo], (unsigned long long)(((gbase + loff) << 4) | mysp));
      }
    }
    if (tid < ns) {
      if (!cutmode) D.cchars[s0 + tid] = mychars;
      if (mykeep) {
        const int64_t j = kbase + koff;
        const int64_t i = cutmode ? (int64_t)A.vorig[s0 + tid] : s0 + tid;
        D.k2r[j] = (int32_t)i; D.coff[j] = gbase + loff;
        if (!cutmode) { D.ctok[j] = (uint32_t)mytok; A.seg_special[j] = mysp; }
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


// ---- the generic matcher once per DISTINCT word ------------------------------------------------------------------------
// The fused kernel hands over every OCCURRENCE of a word its fast paths cannot count; the launch memo knows these words
// (entries with the "generic matcher" mark).  Pass A: the first occurrence to claim a word's memo entry counts it with the
// generic matcher + byte-pair merges and writes the count into the entry; the other occurrences queue (occurrence, entry).
// Pass B: the queued occurrences read the finished counts.  On the bench lexicon ~400 k occurrences are ~2 k distinct
// words; on an open vocabulary of 500 k forms 1.47 M occurrences are ~0.35 M.  Words that are not in the memo (longer
// than 15 bytes, or crowded out of their two slots) are counted per occurrence as before.
__device__ __forceinline__ uint32_t f5_count_slow_word(const TrcDev& D, const TrcTables& T, const TrcFast& F, const uint32_t* byte_id, const SlowWord& w,
                                                       LongEntry* wl, unsigned long long* wl_cnt, int64_t wl_cap, bool* deferred = nullptr) {
  const int vs = (int)((uint32_t)w.len_vs >> 31);
  const int64_t len = w.len_vs & 0x7fffffff;
  const int64_t j = w.unit;
  const int64_t rel = w.pos - D.coff[j];
  return trc_count_unit(T, F, byte_id, D.C + w.pos, len, vs, [&](int64_t ls, int64_t slen) -> uint32_t {
    if (deferred) *deferred = true;
    unsigned long long slot = atomicAdd(&wl_cnt[0], 1ULL);
    unsigned long long off = atomicAdd(&wl_cnt[1], (unsigned long long)slen);
    if ((int64_t)slot < wl_cap) { LongEntry e; e.unit = j; e.lstart = ls + (vs ? 0 : rel + 1); e.len = slen; e.scratch_off = (int64_t)off; wl[slot] = e; }
    return 0u;
  });
}
// Pass A: claim only.  work[]: occurrences that must run the generic matcher (first occurrence of a memo word: entry in
// the high bits; or a word that is not in the memo: entry = all ones); deferred[]: the other occurrences of memo words.
// The lists are dense, so the expensive pass below has a word for every lane.
#define F5_NOSLOT ((1ULL << 24) - 1)
__global__ void k_slow_words_claim(TrcDev D, FuseArgs A, const SlowWord* __restrict__ sl, int64_t n, unsigned long long* work,
                                   unsigned long long* deferred, unsigned long long* counters /* [0] work, [1] deferred */) {
  GSTRIDE(q, n) {
    const SlowWord w = sl[q];
    const int vs = (int)((uint32_t)w.len_vs >> 31);
    const int64_t len = w.len_vs & 0x7fffffff;
    const int Lw = (int)(len + vs);
    long long slot = -1;
    if (len > 0 && Lw <= 15) {
      uint64_t lo = 0, hi = 0;
      const uint8_t* p = D.C + w.pos;
      for (int i = 0; i < Lw; ++i) {
        const uint64_t b = (vs && i == 0) ? (uint64_t)' ' : (uint64_t)p[i - vs];
        if (i < 8) lo |= b << (8 * i); else hi |= b << (8 * (i - 8));
      }
      if (lo != 0) {
        const uint32_t s0 = f3_memo_slot(lo, hi, A.memo_mask);
        for (int way = 0; way < 2 && slot < 0; ++way) {
          const uint32_t sw = s0 ^ (uint32_t)way;
          const unsigned long long ex = ld_relaxed(&A.memo[2 * (size_t)sw]), ey = ld_relaxed(&A.memo[2 * (size_t)sw + 1]);
          if (ex == lo && (ey & 0x00ffffffffffffffULL) == hi && (ey >> 56)) slot = sw;
        }
      }
    }
    if (slot >= 0 && atomicCAS(&A.memo_lock[slot], 1u, 2u) != 1u) {
      deferred[atomicAdd(&counters[1], 1ULL)] = ((unsigned long long)slot << 40) | (unsigned long long)q;
    } else {
      work[atomicAdd(&counters[0], 1ULL)] = ((slot >= 0 ? (unsigned long long)slot : F5_NOSLOT) << 40) | (unsigned long long)q;
    }
  }
}
// Pass A2: the generic matcher + byte-pair merges, a listed occurrence per thread; counts of memo words go into their entry.
__global__ void k_slow_words_count(TrcDev D, TrcTables T, const uint32_t* __restrict__ byte_id, FuseArgs A, const SlowWord* __restrict__ sl,
                                   const unsigned long long* __restrict__ work, const unsigned long long* __restrict__ counters) {
  const TrcFast F = trc_fast_default(T);
  const int64_t n = (int64_t)counters[0];
  GSTRIDE(k, n) {
    const unsigned long long d = work[k];
    const int64_t q = (int64_t)(d & ((1ULL << 40) - 1));
    const unsigned long long slot = d >> 40;
    const SlowWord w = sl[q];
    bool deferred = false;
    const uint32_t c = f5_count_slow_word(D, T, F, byte_id, w, A.wl, A.wl_cnt, A.wl_cap, &deferred);
    atomicAdd(&D.ctok[w.unit], c);
    if (A.wtok) A.wtok[w.pos] = (uint8_t)((deferred || c >= 255u) ? 255u : c);
    if (slot != F5_NOSLOT) {
      const bool whole = c > 0 && c < 250u;             // (words of <= 15 bytes have no span for the long-span kernel)
      if (whole) {
        const unsigned long long ey = ld_relaxed(&A.memo[2 * (size_t)slot + 1]);
        st_relaxed(&A.memo[2 * (size_t)slot + 1], (ey & 0x00ffffffffffffffULL) | ((uint64_t)(c + 1u) << 56));
      }
      A.memo_lock[slot] = whole ? 3u : 4u;
    }
  }
}
// Pass B: the other occurrences read the finished counts.
__global__ void k_slow_words_follow(TrcDev D, TrcTables T, const uint32_t* __restrict__ byte_id, FuseArgs A, const SlowWord* __restrict__ sl,
                                    const unsigned long long* __restrict__ deferred, const unsigned long long* __restrict__ counters) {
  const TrcFast F = trc_fast_default(T);
  const int64_t n = (int64_t)counters[1];
  GSTRIDE(k, n) {
    const unsigned long long d = deferred[k];
    const int64_t q = (int64_t)(d & ((1ULL << 40) - 1));
    const size_t slot = (size_t)(d >> 40);
    const SlowWord w = sl[q];
    bool deferred = false;
    const uint32_t c = A.memo_lock[slot] == 3u ? (uint32_t)(A.memo[2 * slot + 1] >> 56) - 1u
                                               : f5_count_slow_word(D, T, F, byte_id, w, A.wl, A.wl_cnt, A.wl_cap, &deferred);
    atomicAdd(&D.ctok[w.unit], c);
    if (A.wtok) A.wtok[w.pos] = (uint8_t)((deferred || c >= 255u) ? 255u : c);
  }
}
