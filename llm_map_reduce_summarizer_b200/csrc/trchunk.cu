// libtrchunk.so - kernels + C ABI (include/trchunk.h).  Build: see llm_map_reduce_summarizer_b200/build.py
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo --fmad=false -shared -Xcompiler -fPIC
// --fmad=false: the merge cap, sub-chunk times and the percentage are fp64 mul/add chains that CPython
// evaluates without contraction (preprocessor.py:134-153, big_chunkeroosky.py:182-184,351-366,475-506).
#ifdef TRC_HOSTSIM
#include "trc_hostshim.h"
#else
#include <cuda_runtime.h>
#endif
#include <stdio.h>
#include <string>
#include <vector>
#include "../../include/trchunk.h"
#include "trc_common.cuh"
#include "trc_format.cuh"
#include "trc_clean.cuh"
#include "trc_clean_par.cuh"
#include "trc_swar.cuh"
#include "trc_w32.cuh"
#include "trc_tok.cuh"
#include "trc_stages.cuh"
#include "trc_sent.cuh"
#include "trc_vocab_host.h"
#include "trc_ingest_host.h"

static_assert(sizeof(TrcSub) == sizeof(trc_sub), "trc_sub layout");
static_assert(sizeof(TrcChunk) == sizeof(trc_chunk_rec), "trc_chunk_rec layout");
static_assert(sizeof(TrcSent) == sizeof(trc_sent), "trc_sent layout");

// =====================================================================================================
// context
// =====================================================================================================
struct Arena {
  struct Block { char* p; size_t cap; };
  std::vector<Block> blocks;
  size_t used = 0;       // in the last block
  size_t total_req = 0;  // bytes requested since reset
  cudaError_t err = cudaSuccess;
  void* alloc(size_t bytes) {
    bytes = (bytes + 255) & ~(size_t)255;
    if (bytes == 0) bytes = 256;
    total_req += bytes;
    if (blocks.empty() || used + bytes > blocks.back().cap) {
      size_t cap = bytes > (size_t)(64u << 20) ? bytes : (size_t)(64u << 20);
      if (!blocks.empty() && cap < blocks.back().cap) cap = blocks.back().cap;
      char* p = nullptr;
      cudaError_t e = cudaMalloc(&p, cap);
      if (e != cudaSuccess) { err = e; return nullptr; }
      blocks.push_back({p, cap});
      used = 0;
    }
    void* r = blocks.back().p + used;
    used += bytes;
    return r;
  }
  // Start a new batch.  If the previous one spilled into several blocks, fuse them so the steady state
  // is a single block and no cudaMalloc happens inside a timed region.
  void reset() {
    if (blocks.size() > 1) {
      size_t need = total_req + total_req / 4;
      for (auto& b : blocks) cudaFree(b.p);
      blocks.clear();
      char* p = nullptr;
      if (cudaMalloc(&p, need) == cudaSuccess) blocks.push_back({p, need});
    }
    used = 0; total_req = 0; err = cudaSuccess;
  }
  void release() { for (auto& b : blocks) cudaFree(b.p); blocks.clear(); used = 0; }
};

struct LongEntry { int64_t unit, lstart, len, scratch_off; };

struct trc_ctx {
  int device = 0;
  cudaStream_t own_stream = nullptr;
  cudaStream_t aux_stream = nullptr;   // side stream: the slow-word kernel runs beside the merge kernels
  cudaEvent_t ev_aux[2] = {nullptr, nullptr};
  std::string err;
  bool have_tables = false, have_vocab = false;
  TrcTables T{};
  uint8_t* d_props = nullptr;
  uint8_t* d_acls = nullptr;
  uint32_t* d_byte_id = nullptr;
  std::vector<void*> vocab_allocs;
  uint32_t n_special = 0;
  Arena arena;
  TrcDev D{};
  bool have_pre = false, have_chunk = false;
  bool use_fused = true;          // single-pass clean+tokenise kernel (TRCHUNK_NO_FUSED=1 selects the v1 kernels)
  int fused_gen = 5;              // TRCHUNK_FUSED=2 byte-per-lane kernel, 3 8-byte cells, 4 warp-private tiles, 5 (default) CTA tile with 32-byte chunks
  int64_t counts[TRC_N_COUNTERS] = {0};
  int64_t* h_pin = nullptr;        // mapped pinned scratch for small counter reads (64 int64)
  void* d_pin = nullptr;           // its device address
  unsigned long long* d_memo = nullptr;   // word memo of the fused kernel (FuseArgs::memo): entries, then claim words
  bool use_memo = true;
  int64_t launches = 0;
  cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};
  bool ev_pre = false, ev_chunk = false;
  // optional per-kernel timing (trc_set_profiling): events around every launch
  bool prof = false;
  struct ProfRec { const char* name; cudaEvent_t a, b; };
  std::vector<ProfRec> prof_recs;
  std::vector<cudaEvent_t> prof_pool;
  struct ProfAcc { std::string name; double ms; int64_t launches; };
  std::vector<ProfAcc> prof_acc;
  // sentence path
  TrcSentState SS{};
  TrcWordSums WS{};
  // result array table
  struct Arr { const void* p; int64_t bytes; } arr[TRC_A__COUNT] = {};
};

#define CK(call)                                                                                   \
  do {                                                                                             \
    cudaError_t e_ = (call);                                                                       \
    if (e_ != cudaSuccess) {                                                                       \
      ctx->err = std::string(#call) + ": " + cudaGetErrorString(e_);                               \
      return TRC_ECUDA;                                                                            \
    }                                                                                              \
  } while (0)

#define ALLOC(ptr, type, count)                                                                    \
  do {                                                                                             \
    ptr = (type*)ctx->arena.alloc(sizeof(type) * (size_t)((count) > 0 ? (count) : 1));             \
    if (!ptr) { ctx->err = std::string("device allocation failed: ") + cudaGetErrorString(ctx->arena.err); return TRC_ECUDA; } \
  } while (0)

static inline int grid_for(int64_t n, int block) {
  int64_t g = (n + block - 1) / block;
  if (g < 1) g = 1;
  const int64_t cap = 148LL * 64;   // grid-stride beyond 64 waves of 148 SMs
  return (int)(g > cap ? cap : g);
}
#ifdef TRC_HOSTSIM
#define GSTRIDE(i, n) for (int64_t i = 0; i < (n); ++i)
#define WSTRIDE(w, n) for (int64_t w = 0; w < (n); ++w)
#define TRC_LANE 0
#define TRC_NL 1
#else
#define GSTRIDE(i, n) for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < (n); i += (int64_t)gridDim.x * blockDim.x)
#define WSTRIDE(w, n) for (int64_t w = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5; w < (n); w += ((int64_t)gridDim.x * blockDim.x) >> 5)
#define TRC_LANE (threadIdx.x & 31)
#define TRC_NL 32
#endif

// =====================================================================================================
// K3: single-pass exclusive prefix sum with decoupled look-back (hand written; no CUB).
// Tiles take their index from an atomic counter, so a tile only ever waits on tiles that already run.
// Descriptor word: status (2 bits: 0 empty, 1 aggregate, 2 inclusive prefix) | value (62 bits).
// =====================================================================================================
#ifndef TRC_HOSTSIM
#define SCAN_T 256
#define SCAN_I 8
#define SCAN_TILE (SCAN_T * SCAN_I)

__device__ __forceinline__ unsigned long long ld_relaxed(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_relaxed(unsigned long long* p, unsigned long long v) {
  asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

template <class TIn>
__global__ void __launch_bounds__(SCAN_T) k_scan_lookback(const TIn* __restrict__ in, int64_t* __restrict__ out, int64_t n,
                                                          int64_t addc, unsigned long long* desc /* [0]=tile counter, [1..]=tiles */) {
  __shared__ unsigned int s_tile;
  __shared__ long long s_wsum[SCAN_T / 32];
  __shared__ long long s_prefix;
  if (threadIdx.x == 0) s_tile = (unsigned int)atomicAdd(desc, 1ULL);
  __syncthreads();
  const unsigned tile = s_tile;
  unsigned long long* d = desc + 1;
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const int64_t base = (int64_t)tile * SCAN_TILE + (int64_t)threadIdx.x * SCAN_I;
  long long v[SCAN_I];
  long long tsum = 0;
#pragma unroll
  for (int k = 0; k < SCAN_I; ++k) {
    int64_t idx = base + k;
    v[k] = idx < n ? (long long)in[idx] + addc : 0;
    tsum += v[k];
  }
  long long x = tsum;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) { long long y = __shfl_up_sync(0xffffffffu, x, o); if (lane >= o) x += y; }
  if (lane == 31) s_wsum[wid] = x;
  __syncthreads();
  if (wid == 0) {
    long long w = lane < SCAN_T / 32 ? s_wsum[lane] : 0;
#pragma unroll
    for (int o = 1; o < SCAN_T / 32; o <<= 1) { long long y = __shfl_up_sync(0xffffffffu, w, o); if (lane >= o) w += y; }
    if (lane < SCAN_T / 32) s_wsum[lane] = w;
  }
  __syncthreads();
  const long long block_total = s_wsum[SCAN_T / 32 - 1];
  const long long texcl = (wid ? s_wsum[wid - 1] : 0) + x - tsum;
  if (wid == 0) {
    long long excl = 0;
    if (tile == 0) {
      if (lane == 0) st_relaxed(&d[0], (2ULL << 62) | (unsigned long long)block_total);
    } else {
      if (lane == 0) st_relaxed(&d[tile], (1ULL << 62) | (unsigned long long)block_total);
      long long idx = (long long)tile - 1;
      for (;;) {
        long long my = idx - lane;
        unsigned long long dv = (2ULL << 62);
        if (my >= 0) {
          dv = ld_relaxed(&d[my]);
          while ((dv >> 62) == 0) dv = ld_relaxed(&d[my]);
        }
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
      if (lane == 0) st_relaxed(&d[tile], (2ULL << 62) | (unsigned long long)(excl + block_total));
    }
    if (lane == 0) s_prefix = excl;
  }
  __syncthreads();
  long long p = s_prefix + texcl;
#pragma unroll
  for (int k = 0; k < SCAN_I; ++k) {
    int64_t idx = base + k;
    if (idx < n) { out[idx] = p; p += v[k]; if (idx == n - 1) out[n] = p; }
  }
}

#endif
struct ScanWs { unsigned long long* desc; size_t cap_tiles; };

static void prof_begin(trc_ctx* ctx, cudaStream_t st, const char* name);
static void prof_end(trc_ctx* ctx, cudaStream_t st);
template <class TIn>
static int scan_excl(trc_ctx* ctx, cudaStream_t st, const TIn* in, int64_t* out, int64_t n, int64_t addc = 0) {
  if (n <= 0) { CK(cudaMemsetAsync(out, 0, sizeof(int64_t), st)); return TRC_OK; }
#ifdef TRC_HOSTSIM
  { int64_t acc = 0; for (int64_t i = 0; i < n; ++i) { out[i] = acc; acc += (int64_t)in[i] + addc; } out[n] = acc; (void)st; return TRC_OK; }
#else
  int64_t tiles = (n + SCAN_TILE - 1) / SCAN_TILE;
  unsigned long long* desc;
  ALLOC(desc, unsigned long long, tiles + 1);
  CK(cudaMemsetAsync(desc, 0, sizeof(unsigned long long) * (size_t)(tiles + 1), st));
  prof_begin(ctx, st, "k_scan_lookback");
  k_scan_lookback<TIn><<<(unsigned)tiles, SCAN_T, 0, st>>>(in, out, n, addc, desc);
  prof_end(ctx, st);
  ctx->launches++;
  CK(cudaGetLastError());
  return TRC_OK;
#endif
}

// =====================================================================================================
// kernels (thin wrappers around the per-unit routines)
// =====================================================================================================
__global__ void k_clean_count(TrcDev D, TrcTables T, int32_t* keep) {
  GSTRIDE(i, D.n) {
    int64_t a = D.seg_off[i], b = D.seg_off[i + 1];
    TrcCleanOut r = trc_clean_segment<false>(T, D.text + a, b - a, nullptr);
    D.clen[i] = r.bytes; D.cchars[i] = (int32_t)r.chars; keep[i] = r.bytes > 0;
  }
}

#ifdef TRC_HOSTSIM
// Host test build only: the byte-parallel formulation of clean_text (trc_clean_par.cuh), one segment at a
// time, so the CPU suite checks the rules the fused CUDA kernel applies per tile.
static void hostsim_clean_par_count(TrcDev D, TrcTables T, int32_t* keep, uint8_t* cls, uint8_t* code) {
  for (int64_t i = 0; i < D.n; ++i) {
    int64_t a = D.seg_off[i];
    int32_t n = (int32_t)(D.seg_off[i + 1] - a);
    const uint8_t* in = D.text + a;
    for (int32_t k = 0; k < n; ++k) cp_classify(T, in, cls + a, code + a, k, n);
    for (int32_t k = 0; k < n; ++k) if (cp_dup_candidate(cls + a, k, 0)) cp_mark_dup(in, cls + a, code + a, k, 0, n);
    int64_t len = 0; int32_t chars = 0;
    const int32_t fns = cp_first_nonspace(cls + a, 0, n);
    for (int32_t k = 0; k < n; ++k) {
      uint32_t r = cp_code(in, cls + a, code + a, k, 0, n, fns);
      len += cp_outlen(r);
      chars += (r == 3) ? 2 : ((r == 2 || (r == 1 && !(cls[a + k] & CP_CONT))) ? 1 : 0);
    }
    D.clen[i] = len; D.cchars[i] = chars; keep[i] = len > 0;
  }
}
// The 8-byte-cell (SWAR) formulation, tiled exactly like k_clean_tok_fused3 (tiles of `tile_segs` segments
// starting at a 16-byte aligned base, 16-byte loads that pull in bytes of the neighbouring tiles, mask bytes
// the kernel never writes poisoned with 0xff): validates the bit logic of trc_swar.cuh and its tile edges.
// tile_segs <= 0: the whole batch as one tile.
static void hostsim_clean_swar(TrcDev D, TrcTables T, int32_t* keep, uint8_t* code, int64_t braw, int64_t tile_segs) {
  if (tile_segs <= 0) tile_segs = D.n > 0 ? D.n : 1;
  std::vector<uint8_t> buf, arr, leadv;
  for (int64_t s0 = 0; s0 < D.n; s0 += tile_segs) {
    const int64_t ns = (D.n - s0) < tile_segs ? (D.n - s0) : tile_segs;
    const int64_t raw0 = D.seg_off[s0], raw1 = D.seg_off[s0 + ns];
    const int32_t al = (int32_t)(raw0 & 15);
    const int64_t base = raw0 - al;
    const int32_t span = (int32_t)(raw1 - base);
    const int32_t ncell = (span + 7) >> 3, nstep = (span + 255) >> 8, stride = nstep * 32 + 8;
    buf.assign((size_t)nstep * 256 + 32, 0);
    for (int32_t k = 0; k < span; k += 16)
      for (int q = 0; q < 16; ++q) buf[k + q] = base + k + q < braw ? D.text[base + k + q] : 0;
    arr.assign((size_t)stride * 8, 0xff);
    leadv.assign((size_t)nstep * 32, 0);
    SwMasks M; M.ncell = ncell;
    M.sp = arr.data(); M.w = M.sp + stride; M.ct = M.w + stride; M.lt = M.ct + stride; M.ep = M.lt + stride;
    M.bd = M.ep + stride; M.drop = M.bd + stride; M.ds = M.drop + stride;
    memset(M.bd, 0, (size_t)stride * 3);
    for (int q = 0; q < 8; ++q) { M.w[nstep * 32 + q] = 0; M.sp[nstep * 32 + q] = 0; }
    for (int64_t i = s0; i < s0 + ns; ++i) if (D.seg_off[i + 1] > D.seg_off[i]) sw_or_bit(M.bd, (int32_t)(D.seg_off[i] - base));
    for (int32_t c = 0; c < nstep * 32; ++c) {
      uint64_t x; memcpy(&x, buf.data() + 8 * (size_t)c, 8);
      SwCell cl = sw_classify(x);
      M.sp[c] = (uint8_t)cl.sp; M.w[c] = (uint8_t)cl.w; M.ct[c] = (uint8_t)cl.ct; M.lt[c] = (uint8_t)cl.lt; M.ep[c] = (uint8_t)cl.ep;
      leadv[c] = (uint8_t)cl.lead;
    }
    for (int32_t c = 0; c < nstep * 32; ++c) if (leadv[c]) sw_fix_multibyte(T, buf.data(), M, c, leadv[c], span);
    int64_t seg = s0;
    for (int32_t c = 0; c < ncell; ++c) {
      const int32_t k0 = c << 3;
      uint32_t valid = 0xffu;
      if (k0 < al) valid &= (al - k0 >= 8) ? 0u : (0xffu << (al - k0));
      if (k0 + 8 > span) valid &= (span <= k0) ? 0u : (0xffu >> (k0 + 8 - span));
      uint32_t cand = sw_dup_candidates(M, c, valid & 0xffu);
      while (cand) {
        int bit = 0; while (!((cand >> bit) & 1u)) ++bit;
        cand &= cand - 1;
        const int32_t k = k0 + bit;
        if (sw_dup_quick_reject(buf.data(), M, k, span)) continue;
        while (seg + 1 < s0 + ns && D.seg_off[seg + 1] - base <= k) ++seg;
        sw_mark_dup(buf.data(), M, k, (int32_t)(D.seg_off[seg] - base), (int32_t)(D.seg_off[seg + 1] - base));
      }
    }
    for (int32_t c = 0; c < ncell; ++c) {
      SwCodes r = sw_cell_codes(buf.data(), M, c, al, span);
      for (int bit = 0; bit < 8; ++bit) {
        const int32_t k = c * 8 + bit;
        if (k < al) continue;
        if (k >= span) break;
        code[base + k] = ((r.ins >> bit) & 1u) ? 3 : (((r.space >> bit) & 1u) ? 2 : (((r.copy >> bit) & 1u) ? 1 : 0));
      }
    }
  }
  for (int64_t i = 0; i < D.n; ++i) {
    int64_t len = 0; int32_t chars = 0;
    for (int64_t k = D.seg_off[i]; k < D.seg_off[i + 1]; ++k) {
      uint8_t r = code[k];
      len += cp_outlen(r);
      chars += (r == 3) ? 2 : ((r == 2 || (r == 1 && (D.text[k] & 0xC0) != 0x80)) ? 1 : 0);
    }
    D.clen[i] = len; D.cchars[i] = chars; keep[i] = len > 0;
  }
}

// The 32-byte-chunk formulation (trc_w32.cuh), tiled exactly like k_clean_tok_w32 (trc_fused4.cuh): tiles of
// `tile_segs` segments that start at a 16-byte aligned base, the neighbouring tiles' bytes visible past the span,
// class masks cut to the tile's own bytes, the classification through the same 256-entry table.
static void hostsim_clean_w32(TrcDev D, TrcTables T, int32_t* keep, uint8_t* code, int64_t braw, int64_t tile_segs) {
  if (tile_segs <= 0) tile_segs = D.n > 0 ? D.n : 1;
  std::vector<uint64_t> lut(256);
  for (uint32_t b = 0; b < 256; ++b) lut[b] = w32_lut_entry(T, b);
  std::vector<uint8_t> buf;
  std::vector<uint32_t> arr;
  for (int64_t s0 = 0; s0 < D.n; s0 += tile_segs) {
    const int64_t ns = (D.n - s0) < tile_segs ? (D.n - s0) : tile_segs;
    const int64_t raw0 = D.seg_off[s0], raw1 = D.seg_off[s0 + ns];
    const int32_t al = (int32_t)(raw0 & 15);
    const int64_t base = raw0 - al;
    const int32_t span = (int32_t)(raw1 - base);
    const int32_t nch = (span + 31) >> 5, nstep = (span + 1023) >> 10, nall = nstep * 32;
    buf.assign((size_t)nall * 32 + 64, 0);
    for (int32_t k = 0; k < nall * 32; ++k) buf[k] = base + k < braw ? D.text[base + k] : 0;
    const size_t stride = (size_t)nall + 2;
    arr.assign(stride * 8, 0u);
    W32 M; M.nch = nch;
    M.sp = arr.data() + 1; M.w = M.sp + stride; M.ct = M.w + stride; M.bd = M.ct + stride; M.ep = M.bd + stride;
    M.lt = M.ep + stride; M.drop = M.lt + stride; M.ds = M.drop + stride;
    for (int64_t i = s0; i < s0 + ns; ++i) if (D.seg_off[i + 1] > D.seg_off[i]) w32_or_bit(M.bd, (int32_t)(D.seg_off[i] - base));
    std::vector<uint32_t> leadv((size_t)nall, 0u);
    for (int32_t c = 0; c < nall; ++c) {
      uint32_t x[8];
      memcpy(x, buf.data() + 32 * (size_t)c, 32);
      const W32Cls cl = w32_classify(lut.data(), x);
      const uint32_t v = w32_valid(c, al, span);
      M.sp[c] = cl.sp & v; M.w[c] = cl.w & v; M.ct[c] = cl.ct & v; M.ep[c] = cl.ep & v; M.lt[c] = cl.lt & v;
      leadv[c] = cl.na & ~cl.ct & v;
    }
    for (int32_t c = 0; c < nall; ++c) if (leadv[c]) w32_fix_multibyte(T, buf.data(), M, c, leadv[c], span);
    std::vector<int32_t> relv((size_t)ns + 1);
    for (int64_t i = 0; i <= ns; ++i) relv[(size_t)i] = (int32_t)(D.seg_off[s0 + i] - base);
    auto seg_of = [&](int32_t k) { int lo = 0, hi = (int)ns; while (hi - lo > 1) { const int mid = (lo + hi) >> 1; if (relv[mid] <= k) lo = mid; else hi = mid; } return lo; };
    for (int32_t c = 0; c < nch; ++c) w32_dup_chunk(buf.data(), M, c, w32_valid(c, al, span), span, relv.data(), seg_of);
    for (int32_t c = 0; c < nch; ++c) {
      const W32Codes r = w32_chunk_codes(buf.data(), M, c, al, span);
      for (int bit = 0; bit < 32; ++bit) {
        const int32_t k = c * 32 + bit;
        if (k < al) continue;
        if (k >= span) break;
        code[base + k] = ((r.ins >> bit) & 1u) ? 3 : (((r.space >> bit) & 1u) ? 2 : (((r.copy >> bit) & 1u) ? 1 : 0));
      }
    }
  }
  for (int64_t i = 0; i < D.n; ++i) {
    int64_t len = 0; int32_t chars = 0;
    for (int64_t k = D.seg_off[i]; k < D.seg_off[i + 1]; ++k) {
      uint8_t r = code[k];
      len += cp_outlen(r);
      chars += (r == 3) ? 2 : ((r == 2 || (r == 1 && (D.text[k] & 0xC0) != 0x80)) ? 1 : 0);
    }
    D.clen[i] = len; D.cchars[i] = chars; keep[i] = len > 0;
  }
}

static void hostsim_clean_par_write(TrcDev D, const uint8_t* code) {
  for (int64_t j = 0; j < D.n_kept; ++j) {
    int32_t r = D.k2r[j];
    int64_t a = D.seg_off[r], b = D.seg_off[r + 1];
    uint8_t* o = D.C + D.coff[j];
    for (int64_t k = a; k < b; ++k) {
      uint8_t c = code[k];
      if (c == 1) *o++ = D.text[k];
      else if (c == 2) *o++ = ' ';
      else if (c == 3) { *o++ = ' '; *o++ = D.text[k]; }
    }
  }
}
#endif

__global__ void k_compact(TrcDev D, const int32_t* keep) {
  GSTRIDE(i, D.n) {
    if (keep[i]) {
      int64_t j = D.kscan[i];
      D.k2r[j] = (int32_t)i; D.coff[j] = D.cscan[i];
      int64_t lo = 0, hi = D.n_doc;            // document of segment i: last x with doc_off[x] <= i
      while (hi - lo > 1) { int64_t mid = (lo + hi) >> 1; if (D.doc_off[mid] <= i) lo = mid; else hi = mid; }
      D.sdoc[j] = (int32_t)lo;
    }
    if (i == D.n - 1) D.coff[D.kscan[D.n]] = D.cscan[D.n];
  }
}

__global__ void k_map_off(const int64_t* __restrict__ src_off, const int64_t* __restrict__ scan, int64_t* __restrict__ dst, int64_t n1) {
  GSTRIDE(x, n1) dst[x] = scan[src_off[x]];
}

__global__ void k_clean_write(TrcDev D, TrcTables T) {
  GSTRIDE(j, D.n_kept) {
    int32_t r = D.k2r[j];
    int64_t a = D.seg_off[r], b = D.seg_off[r + 1];
    trc_clean_segment<true>(T, D.text + a, b - a, D.C + D.coff[j]);
  }
}

#ifdef TRC_HOSTSIM
static uint32_t hostsim_bpe_long(const TrcTables& T, const uint32_t* byte_id, const TrcStr& S, int64_t a, int64_t len) {
  std::vector<uint32_t> id((size_t)len);
  for (int64_t i = 0; i < len; ++i) id[i] = byte_id[S.at(a + i)];
  for (;;) {
    uint32_t best = TRC_NO_RANK; int64_t bi = -1;
    for (int64_t i = 0; i + 1 < (int64_t)id.size(); ++i) { uint32_t r = trc_pair_rank(T, id[i], id[i + 1]); if (r < best) { best = r; bi = i; } }
    if (bi < 0) break;
    id[bi] = best; id.erase(id.begin() + bi + 1);
  }
  return (uint32_t)id.size();
}
#endif
__global__ void k_count_units(TrcTables T, const uint32_t* __restrict__ byte_id, const uint8_t* __restrict__ bytes,
                              const int64_t* __restrict__ ustart, const int64_t* __restrict__ uend, int64_t n, int vs,
                              uint32_t* __restrict__ counts, int32_t* __restrict__ special, LongEntry* wl,
                              unsigned long long* wl_cnt, int64_t wl_cap, int64_t coop_min, unsigned long long* coop_cnt,
                              int32_t* coop_owner, int64_t* coop_base) {
  GSTRIDE(j, n) {
    const uint8_t* p = bytes + ustart[j];
    int64_t nb = uend[j] - ustart[j];
    if (coop_min > 0 && nb >= coop_min) {                  // a long unit of cleaned text: its 32-byte chunks go to k_count_units_coop
      const int64_t nch = (nb + 31) >> 5;
      const int64_t off = (int64_t)atomicAdd(coop_cnt, (unsigned long long)nch);
      coop_base[j] = off;
      for (int64_t k = 0; k < nch; ++k) coop_owner[off + k] = (int32_t)j;
      counts[j] = 0;
      continue;
    }
    uint32_t c = trc_count_unit(T, trc_fast_default(T), byte_id, p, nb, vs, [&](int64_t ls, int64_t len) -> uint32_t {
#ifdef TRC_HOSTSIM
      TrcStr S; S.p = p; S.n = nb + vs; S.vs = vs;
      atomicAdd(&wl_cnt[0], 1ULL);
      if (true) return hostsim_bpe_long(T, byte_id, S, ls, len);
#endif
      unsigned long long slot = atomicAdd(&wl_cnt[0], 1ULL);
      unsigned long long off = atomicAdd(&wl_cnt[1], (unsigned long long)len);
      if ((int64_t)slot < wl_cap) { LongEntry e; e.unit = j; e.lstart = ls; e.len = len; e.scratch_off = (int64_t)off; wl[slot] = e; }
      return 0u;
    });
    counts[j] = c;
    if (special) special[j] = trc_find_special(T, p, nb);
  }
}

#ifndef TRC_HOSTSIM
// Long units of CLEANED text (a sentence of several KB in an over-budget monologue): a thread walked them byte by byte, a
// load latency each.  Cleaned text has single U+0020 between words and the pre-tokeniser never joins across one
// (DESIGN.md §3), so the count of a unit is the sum over its words: k_count_units lists the 32-byte chunks of every unit
// of coop_min bytes or more, and here a thread per chunk counts the words that START in its chunk.
__global__ void k_count_units_coop(TrcTables T, const uint32_t* __restrict__ byte_id, const uint8_t* __restrict__ bytes,
                                   const int64_t* __restrict__ ustart, const int64_t* __restrict__ uend, int vs,
                                   uint32_t* __restrict__ counts, LongEntry* wl, unsigned long long* wl_cnt, int64_t wl_cap,
                                   const unsigned long long* __restrict__ coop_cnt, const int32_t* __restrict__ coop_owner,
                                   const int64_t* __restrict__ coop_base) {
  const TrcFast F = trc_fast_default(T);
  const int64_t nchunks = (int64_t)*coop_cnt;
  GSTRIDE(g, nchunks) {
    const int64_t j = coop_owner[g];
    const int64_t a = ustart[j], b = uend[j];
    const int64_t c0 = a + ((g - coop_base[j]) << 5), c1 = c0 + 32 < b ? c0 + 32 : b;
    uint32_t total = 0;
    for (int64_t i = c0; i < c1; ++i) {
      if (i != a && bytes[i] != ' ') continue;
      int64_t e = i + 1;
      while (e < b && bytes[e] != ' ') ++e;
      const int wvs = i == a ? vs : 0;
      const int64_t shift = i == a ? 0 : (i - a) + vs;     // logical index of the word's first byte inside the unit
      total += trc_count_unit(T, F, byte_id, bytes + i, e - i, wvs, [&](int64_t ls, int64_t len) -> uint32_t {
        unsigned long long slot = atomicAdd(&wl_cnt[0], 1ULL);
        unsigned long long off = atomicAdd(&wl_cnt[1], (unsigned long long)len);
        if ((int64_t)slot < wl_cap) { LongEntry le; le.unit = j; le.lstart = shift + ls; le.len = len; le.scratch_off = (int64_t)off; wl[slot] = le; }
        return 0u;
      });
    }
    if (total) atomicAdd(&counts[j], total);
  }
}
// Byte-pair merge of one long span per block; the four work arrays live in global scratch.
#define LONG_T 256
#define LONGB_T 1024                     // threads of the long-span block = most chunks a span is cut into
#define LONGB_CAND 2048                  // rank-r pairs one round can take through the short list
__global__ void __launch_bounds__(LONGB_T) k_bpe_long(TrcTables T, const uint32_t* __restrict__ byte_id, const uint8_t* __restrict__ bytes,
                                                      const int64_t* __restrict__ ustart, const int64_t* __restrict__ uend, int vs,
                                                      uint32_t* counts, const LongEntry* __restrict__ wl, int64_t n_entries,
                                                      uint32_t* __restrict__ scratch /* 4 words per span byte */, int64_t scratch_total) {
  __shared__ unsigned long long s_red[LONGB_T / 32];
  __shared__ unsigned long long s_best;
  __shared__ unsigned long long s_cm[LONGB_T];          // per chunk: smallest (rank << 32 | position)
  __shared__ int32_t s_cand[LONGB_CAND];
  __shared__ uint8_t s_dirty[LONGB_T];
  __shared__ int s_ncand, s_cnt, s_bad;
  const uint32_t SEL = 0x80000000u, IDM = 0x7fffffffu;  // token ids stay below 2^21: the top bit marks a pair taken this round
  const int tid = threadIdx.x;
  for (int64_t w = blockIdx.x; w < n_entries; w += gridDim.x) {
    const LongEntry e = wl[w];
    const int64_t len = e.len;
    TrcStr S; S.p = bytes + ustart[e.unit]; S.vs = vs; S.n = uend[e.unit] - ustart[e.unit] + vs;
    int64_t l0 = e.lstart;
    if (l0 < 0) { S.p = bytes + (-1 - l0); S.vs = 0; S.n = len; l0 = 0; }      // absolute position (parts of a cut segment, trc_fused5.cuh)
    uint32_t* id = scratch + e.scratch_off;
    uint32_t* rk = scratch + scratch_total + e.scratch_off;
    int32_t* nx = (int32_t*)(scratch + 2 * scratch_total + e.scratch_off);
    int32_t* pv = (int32_t*)(scratch + 3 * scratch_total + e.scratch_off);
    // One ROUND per distinct rank instead of one per merge: with r the smallest rank present, tiktoken's order merges
    // every pair of rank r, left to right, before anything else - PROVIDED no merge creates a pair of rank below r.  That
    // holds for rank tables that are merge lists (a token is ranked after its parts) and is checked after every round: a
    // table that violates it sends the span back to the start with one merge per round.  Overlapping occurrences ("aaa":
    // two pairs that share the middle token) go leftmost first: a pair is taken iff an even number of rank-r pairs chain up
    // to it from the left.  The span is cut into at most 1 024 chunks whose minima live in shared memory, so a round
    // reduces the chunk minima, lists the rank-r pairs of the chunks that hold one, merges them and rescans only the
    // chunks it touched: a few thousand cheap rounds for a 64 KB span instead of ~60 000 reductions over the whole span.
    const int64_t csz = ((len + LONGB_T - 1) / LONGB_T + 31) & ~(int64_t)31;     // chunk size (multiple of 32)
    const int nchunk = (int)((len + csz - 1) / csz);
    auto chunk_min = [&](int c) {
      unsigned long long best = ~0ULL;
      const int64_t a0 = (int64_t)c * csz, a1 = a0 + csz < len ? a0 + csz : len;
      for (int64_t i = a0; i < a1; ++i) { const unsigned long long k = ((unsigned long long)rk[i] << 32) | (unsigned long long)i; best = k < best ? k : best; }
      return best;
    };
    bool batched = true;
    int64_t m = len;
    for (;;) {
      for (int64_t i = tid; i < len; i += LONGB_T) { id[i] = byte_id[S.at(l0 + i)]; nx[i] = (int32_t)(i + 1); pv[i] = (int32_t)(i - 1); }
      __syncthreads();
      for (int64_t i = tid; i < len; i += LONGB_T) rk[i] = i + 1 < len ? trc_pair_rank(T, id[i], id[i + 1]) : TRC_NO_RANK;
      __syncthreads();
      s_cm[tid] = tid < nchunk ? chunk_min(tid) : ~0ULL;
      s_dirty[tid] = 0;
      m = len;
      bool bad = false;
      __syncthreads();
      for (;;) {
        unsigned long long best = s_cm[tid];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) { unsigned long long y = __shfl_xor_sync(0xffffffffu, best, o); best = y < best ? y : best; }
        if ((tid & 31) == 0) s_red[tid >> 5] = best;
        __syncthreads();
        if (tid < 32) {
          unsigned long long bb = s_red[tid];
#pragma unroll
          for (int o = 16; o > 0; o >>= 1) { unsigned long long y = __shfl_xor_sync(0xffffffffu, bb, o); bb = y < bb ? y : bb; }
          if (tid == 0) { s_best = bb; s_ncand = 0; s_cnt = 0; s_bad = 0; }
        }
        __syncthreads();
        const uint32_t r = (uint32_t)(s_best >> 32);
        if (r == TRC_NO_RANK) break;
        if (!batched) {                                 // one merge per round
          if (tid == 0) {
            const int32_t i = (int32_t)(s_best & 0xffffffffu);
            const int32_t j = nx[i], nj = nx[j];
            id[i] = r; nx[i] = nj; rk[j] = TRC_NO_RANK;
            if (nj < len) { pv[nj] = i; rk[i] = trc_pair_rank(T, r, id[nj]); } else rk[i] = TRC_NO_RANK;
            const int32_t p = pv[i];
            if (p >= 0) rk[p] = trc_pair_rank(T, id[p], r);
            s_dirty[i / csz] = 1; s_dirty[j / csz] = 1; if (p >= 0) s_dirty[p / csz] = 1;
          }
          --m;
          __syncthreads();
          if (tid < nchunk && s_dirty[tid]) { s_cm[tid] = chunk_min(tid); s_dirty[tid] = 0; }
          __syncthreads();
          continue;
        }
        // list the pairs of rank r (only chunks whose minimum is r can hold one)
        if (tid < nchunk && (uint32_t)(s_cm[tid] >> 32) == r) {
          const int64_t a0 = (int64_t)tid * csz, a1 = a0 + csz < len ? a0 + csz : len;
          for (int64_t i = a0; i < a1; ++i) if (rk[i] == r) { const int q = atomicAdd(&s_ncand, 1); if (q < LONGB_CAND) s_cand[q] = (int32_t)i; }
        }
        __syncthreads();
        const int ncand = s_ncand;
        const bool listed = ncand <= LONGB_CAND;        // otherwise every thread walks its share of the span
        // A: choose (reads rk / pv only, marks in the id word)
        int taken = 0;
        auto choose = [&](int64_t i) {
          int back = 0;
          for (int32_t p = pv[i]; p >= 0 && rk[p] == r; p = pv[p]) ++back;
          if (!(back & 1)) { id[i] |= SEL; ++taken; }
        };
        if (listed) { for (int q = tid; q < ncand; q += LONGB_T) choose(s_cand[q]); }
        else { for (int64_t i = tid; i < len; i += LONGB_T) if (rk[i] == r) choose(i); }
        if (taken) atomicAdd(&s_cnt, taken);
        __syncthreads();
        // B: merge: the right token of every taken pair dies, links skip it
        auto merge = [&](int64_t i) {
          if (!(id[i] & SEL)) return;
          const int32_t j = nx[i], nj = nx[j];
          rk[j] = TRC_NO_RANK;
          nx[i] = nj;
          if (nj < len) pv[nj] = (int32_t)i;
          id[i] = r | SEL;
          s_dirty[i / csz] = 1; s_dirty[j / csz] = 1;
        };
        if (listed) { for (int q = tid; q < ncand; q += LONGB_T) merge(s_cand[q]); }
        else { for (int64_t i = tid; i < len; i += LONGB_T) merge(i); }
        __syncthreads();
        // C: ranks of the pairs that changed; none may fall below r
        int worse = 0;
        auto rerank = [&](int64_t i) {
          if (!(id[i] & SEL)) return;
          const int32_t right = nx[i], left = pv[i];
          const uint32_t a = right < len ? trc_pair_rank(T, r, id[right] & IDM) : TRC_NO_RANK;
          rk[i] = a;
          worse |= a < r;
          if (left >= 0) { const uint32_t b2 = trc_pair_rank(T, id[left] & IDM, r); rk[left] = b2; worse |= b2 < r; s_dirty[left / csz] = 1; }
        };
        if (listed) { for (int q = tid; q < ncand; q += LONGB_T) rerank(s_cand[q]); }
        else { for (int64_t i = tid; i < len; i += LONGB_T) rerank(i); }
        if (worse) s_bad = 1;
        __syncthreads();
        // D: clear the marks, refresh the minima of the chunks that changed
        if (listed) { for (int q = tid; q < ncand; q += LONGB_T) { const int32_t i = s_cand[q]; if (id[i] & SEL) id[i] = r; } }
        else { for (int64_t i = tid; i < len; i += LONGB_T) if (id[i] & SEL) id[i] = r; }
        m -= s_cnt;
        bad = s_bad != 0;
        __syncthreads();
        if (tid < nchunk && s_dirty[tid]) { s_cm[tid] = chunk_min(tid); s_dirty[tid] = 0; }
        __syncthreads();
        if (bad) break;
      }
      if (!bad) break;
      batched = false;                                  // not a merge list: start over, one merge per round
    }
    if (tid == 0) atomicAdd(&counts[e.unit], (uint32_t)m);
    __syncthreads();
  }
}

#endif
#ifndef TRC_HOSTSIM
#define TRC_MEMO_LOG2 20
#include "trc_giant.cuh"
#include "trc_fused.cuh"
#include "trc_fused3.cuh"
#include "trc_fused4.cuh"
#include "trc_fused5.cuh"
#endif

__global__ void k_special_docs(TrcDev D, const int32_t* __restrict__ seg_special, const int32_t* __restrict__ spk_special) {
  GSTRIDE(j, D.n_kept) {
    const int32_t r = D.k2r[j];
    const int32_t sp = D.spk[r];
    if (D.kspk) { D.kspk[j] = sp; D.kdur[j] = D.end[r] - D.start[r]; }     // per piece, for the forward scans of trc_merge_next
    if (seg_special[j] >= 0 || spk_special[sp] >= 0)
      atomicMin((unsigned long long*)&D.doc_err_seg[D.sdoc[j]], (unsigned long long)j);
  }
}
__global__ void k_doc_err(TrcDev D, const int32_t* __restrict__ seg_special, const int32_t* __restrict__ spk_special) {
  GSTRIDE(x, D.n_doc) {
    int64_t j = D.doc_err_seg[x];
    if (j != INT64_MAX) {
      int32_t s = spk_special[D.spk[D.k2r[j]]];
      D.doc_err[x] = 1 + (s >= 0 ? s : seg_special[j]);
    }
  }
}

__global__ void k_build_spk_units(const uint8_t* __restrict__ spk_bytes, const int64_t* __restrict__ spk_off, int32_t n_spk,
                                  uint8_t* __restrict__ ubytes, int64_t* __restrict__ uoff) {
  GSTRIDE(s, n_spk) {
    int64_t a = spk_off[s], b = spk_off[s + 1], o = a + s;
    for (int64_t k = a; k < b; ++k) ubytes[o + (k - a)] = spk_bytes[k];
    ubytes[o + (b - a)] = ':';
    uoff[s] = o;
    if (s == n_spk - 1) uoff[n_spk] = b + n_spk;
  }
}

__global__ void k_fill_i64(int64_t* p, int64_t n, int64_t v) { GSTRIDE(i, n) p[i] = v; }
// ---- already-processed input (include/trchunk.h: trc_batch.already_processed)
__global__ void k_identity_pieces(TrcDev D) {
  GSTRIDE(j, D.n_kept) {
    D.k2r[j] = (int32_t)j;
    int64_t lo = 0, hi = D.n_doc;
    while (hi - lo > 1) { int64_t mid = (lo + hi) >> 1; if (D.doc_off[mid] <= j) lo = mid; else hi = mid; }
    D.sdoc[j] = (int32_t)lo;
    int32_t c = 0;
    for (int64_t k = D.coff[j]; k < D.coff[j + 1]; ++k) c += (D.C[k] & 0xC0) != 0x80;
    D.cchars[j] = c;
  }
}
__global__ void k_flag_groups(const int64_t* __restrict__ group_off, int64_t m, int32_t* gflag) { GSTRIDE(g, m) gflag[group_off[g]] = 1; }
__global__ void k_verify_groups(TrcDev D, TrcTables T, const double* __restrict__ gs, const double* __restrict__ ge,
                                const uint8_t* __restrict__ full, const int64_t* __restrict__ full_off, const uint32_t* __restrict__ ftok) {
  GSTRIDE(g, D.m) {
    int32_t a = D.gstart[g], b = D.gstart[g + 1];
    bool ok = D.start[D.k2r[a]] == gs[g] && D.end[D.k2r[b - 1]] == ge[g];
    int64_t len = D.poff[b] - D.poff[a];
    ok = ok && len == full_off[g + 1] - full_off[g];
    if (ok) { const uint8_t* x = D.M + D.poff[a]; const uint8_t* y = full + full_off[g]; for (int64_t k = 0; k < len; ++k) if (x[k] != y[k]) { ok = false; break; } }
    if (!ok) D.doc_err[D.sdoc[a]] = -1;
    // tokens of the segment as the reference counts them: on its OWN text (big_chunkeroosky.py:82-83)
    D.gtok[g] = 2 + trc_ts_tokens(T, D.start[D.k2r[a]]) + D.spk_tok[D.spk[D.k2r[a]]] + ftok[g];
  }
}

__global__ void k_clen32(const int64_t* __restrict__ coff, int64_t nk, int32_t* out) { GSTRIDE(j, nk) out[j] = (int32_t)(coff[j + 1] - coff[j]); }
__global__ void k_merge_next(TrcDev D) { GSTRIDE(j, D.n_kept) trc_merge_next(D, j); }
__global__ void k_group_flag_nocap(TrcDev D) { GSTRIDE(j, D.n_kept) trc_group_flag_nocap(D, j); }
__global__ void k_orbit(const int64_t* __restrict__ range_off, const int32_t* __restrict__ nxt, int32_t* flag, int64_t n_doc) {
  GSTRIDE(x, n_doc) trc_orbit_walk(range_off, nxt, flag, x);
}
#ifndef TRC_HOSTSIM
// Orbit of ONE long chain (a single huge document: config 4 walks 2 k merged-segment starts over 1 M pieces, then 10 k chunk
// starts over 10 k items).  The hops are a dependent chain, so the links are staged in shared memory by the whole block:
//  * up to ORBIT_DBL links: pointer doubling.  reach = {f^k(first) : k < 2^r} and J = f^(2^r) after r rounds; a round marks
//    J(reach) and squares J, all threads at once; it ends when J(first) has left the range.  10 k hops take 14 rounds of a
//    few hundred cycles instead of 10 k shared-memory round trips by one thread (308 us -> ~10 us);
//  * longer ranges: the links only point forward, so the range is staged window by window and one thread walks each window
//    in shared memory (~30 cycles per hop instead of an L2 round trip).
#define ORBIT_SMEM 49152
#define ORBIT_DBL 20480
__global__ void __launch_bounds__(1024) k_orbit_staged(const int64_t* __restrict__ range_off, const int32_t* __restrict__ nxt, int32_t* flag, int64_t n_doc) {
  extern __shared__ int32_t s_nxt[];
  __shared__ int64_t s_pos;
  for (int64_t x = blockIdx.x; x < n_doc; x += gridDim.x) {
    const int64_t a = range_off[x], b = range_off[x + 1];
    const int64_t n = b - a;
    if (n <= 0) continue;
    if (n <= ORBIT_DBL) {
      const int32_t nn = (int32_t)n;
      int32_t* Jc = s_nxt; int32_t* Jn = s_nxt + nn;
      uint8_t* reach = reinterpret_cast<uint8_t*>(s_nxt + 2 * nn);
      for (int32_t k = threadIdx.x; k < nn; k += blockDim.x) {
        const int64_t t = (int64_t)nxt[a + k] - a;
        Jc[k] = t < nn ? (int32_t)t : nn;
        reach[k] = k == 0;
      }
      __syncthreads();
      for (;;) {
        for (int32_t k = threadIdx.x; k < nn; k += blockDim.x) {
          const int32_t t = Jc[k];
          if (t < nn) { if (reach[k]) reach[t] = 1; Jn[k] = Jc[t]; } else Jn[k] = nn;
        }
        __syncthreads();
        int32_t* tmp = Jc; Jc = Jn; Jn = tmp;
        if (Jc[0] >= nn) break;
        __syncthreads();
      }
      __syncthreads();
      for (int32_t k = threadIdx.x; k < nn; k += blockDim.x) if (reach[k]) flag[a + k] = 1;
      __syncthreads();
      continue;
    }
    if (threadIdx.x == 0) s_pos = 0;
    __syncthreads();
    for (;;) {
      const int64_t w0 = s_pos;
      if (w0 >= n) break;
      const int64_t cnt = (n - w0) < ORBIT_SMEM ? (n - w0) : ORBIT_SMEM;
      for (int64_t k = threadIdx.x; k < cnt; k += blockDim.x) s_nxt[k] = (int32_t)(nxt[a + w0 + k] - a - w0);   // relative to the window
      __syncthreads();
      if (threadIdx.x == 0) {
        int64_t p = 0;
        while (p < cnt) { flag[a + w0 + p] = 1; p = s_nxt[p]; }
        s_pos = w0 + p;
      }
      __syncthreads();
    }
    __syncthreads();
  }
}
#endif
#ifndef TRC_HOSTSIM
// Many documents of moderate length (config 3: 10 k documents of ~220 pieces, ~16 hops each): a WARP per document stages
// the document's links in shared memory with coalesced loads and lane 0 walks them there; a thread per document
// paid an L2 round trip per hop and the longest document set the kernel's time.  Longer ranges are walked in place.
#define ORBIT_W_LINKS 1024
__global__ void __launch_bounds__(256) k_orbit_warp(const int64_t* __restrict__ range_off, const int32_t* __restrict__ nxt, int32_t* flag, int64_t n_doc) {
  __shared__ int32_t s_nxt[8][ORBIT_W_LINKS];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  for (int64_t x = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5; x < n_doc; x += ((int64_t)gridDim.x * blockDim.x) >> 5) {
    const int64_t a = range_off[x], b = range_off[x + 1];
    const int64_t n = b - a;
    if (n <= 0) continue;
    if (n > ORBIT_W_LINKS) { if (lane == 0) trc_orbit_walk(range_off, nxt, flag, x); continue; }
    for (int64_t k = lane; k < n; k += 32) s_nxt[w][k] = (int32_t)(nxt[a + k] - a);
    __syncwarp();
    if (lane == 0) { int32_t p = 0; while (p < n) { flag[a + p] = 1; p = s_nxt[w][p]; } }
    __syncwarp();
  }
}
#endif
__global__ void k_scatter_starts(const int32_t* __restrict__ flag, const int64_t* __restrict__ scan, int32_t* starts, int64_t n) {
  GSTRIDE(j, n) {
    if (flag[j]) starts[scan[j]] = (int32_t)j;
    if (j == n - 1) starts[scan[n]] = (int32_t)n;
  }
}
__global__ void k_piece_len(TrcDev D, int32_t* plen_sub) {
  GSTRIDE(j, D.n_kept) {
    trc_piece_len(D, j);
    if (plen_sub) {   // layout WITH timestamps, used for sub-chunk text when preserve_timestamps is off
      int64_t g = D.gscan[j + 1] - 1;
      int32_t a = D.gstart[g], b = D.gstart[g + 1];
      int64_t len = D.coff[j + 1] - D.coff[j];
      if (b - a > 1) { len += 3 + trc_ts_len(D.start[D.k2r[j]]); if (j + 1 < b) len += 1; }
      plen_sub[j] = (int32_t)len;
    }
  }
}
__global__ void k_emit_merged(TrcDev D) { WSTRIDE(j, D.n_kept) trc_emit_merged(D, j, TRC_LANE, TRC_NL); }
__global__ void k_group_stats(TrcDev D, TrcTables T) { WSTRIDE(g, D.m) trc_group_stats(D, T, g, TRC_LANE, TRC_NL); }
__global__ void k_group_stats_t(TrcDev D, TrcTables T) { GSTRIDE(g, D.m) trc_group_stats(D, T, g, 0, 1); }   // short groups: a thread each
#ifndef TRC_HOSTSIM
// groups of a dozen pieces (config 3): eight lanes per group, four groups per warp
__global__ void k_group_stats_8(TrcDev D, TrcTables T) {
  const int lane = threadIdx.x & 31;
  for (int64_t base = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) - lane; (base >> 3) < D.m; base += (int64_t)gridDim.x * blockDim.x) {
    const int64_t g = (base + lane) >> 3;
    const uint32_t mask = __ballot_sync(0xffffffffu, g < D.m);
    if (g < D.m) trc_group_stats(D, T, g, lane & 7, 8, mask);
  }
}
#endif

// ---- chunk stage
// warp per merged segment (the sub-chunk greedy of an over-budget combined segment walks thousands of pieces)
#ifdef TRC_HOSTSIM
#define TRC_PIECE_TOKENS(hi_) PieceTokDirect{&D, &T}
#else
#define TRC_PIECE_TOKENS(hi_) PieceTokWarp{&D, &T, (hi_), (int)TRC_LANE, -64, 0}
#endif
__global__ void k_large_count(TrcDev D, TrcTables T, int32_t* nsub_g, int32_t* n_raw_large) {
  WSTRIDE(g, D.m) {
    int32_t a = D.gstart[g];
    bool large = D.gtok[g] > (int64_t)D.doc_E[D.sdoc[a]];
    int32_t ns = 0;
    if (large) {
      if (D.gstart[g + 1] - a > 1) ns = trc_split_combined<false>(D, g, nullptr, TRC_PIECE_TOKENS(D.gstart[g + 1]), TRC_LANE == 0);
      else { ns = -1; if (TRC_LANE == 0) atomicAdd(n_raw_large, 1); }   // sentence path: counted by the sentence stage
    }
    if (TRC_LANE == 0) nsub_g[g] = ns;
  }
}
__global__ void k_large_count_t(TrcDev D, TrcTables T, int32_t* nsub_g, int32_t* n_raw_large) {   // short groups: a thread each
  GSTRIDE(g, D.m) {
    int32_t a = D.gstart[g];
    bool large = D.gtok[g] > (int64_t)D.doc_E[D.sdoc[a]];
    int32_t ns = 0;
    if (large) {
      if (D.gstart[g + 1] - a > 1) ns = trc_split_combined<false>(D, g, nullptr, PieceTokDirect{&D, &T}, true);
      else { ns = -1; atomicAdd(n_raw_large, 1); }
    }
    nsub_g[g] = ns;
  }
}
__global__ void k_large_write_t(TrcDev D, TrcTables T, const int32_t* __restrict__ nsub_g) {
  GSTRIDE(g, D.m) {
    int32_t a = D.gstart[g];
    if (nsub_g[g] > 0 && D.gstart[g + 1] - a > 1) trc_split_combined<true>(D, g, D.subs + D.sub_off[g], PieceTokDirect{&D, &T}, true);
  }
}
__global__ void k_nitems(TrcDev D, const int32_t* __restrict__ nsub_g) {
  GSTRIDE(g, D.m) {
    int32_t a = D.gstart[g];
    bool large = D.gtok[g] > (int64_t)D.doc_E[D.sdoc[a]];
    D.nitems_g[g] = large ? nsub_g[g] : 1;
  }
}
__global__ void k_large_write(TrcDev D, TrcTables T, const int32_t* __restrict__ nsub_g) {
  WSTRIDE(g, D.m) {
    int32_t a = D.gstart[g];
    if (nsub_g[g] > 0 && D.gstart[g + 1] - a > 1)
      trc_split_combined<true>(D, g, D.subs + D.sub_off[g], TRC_PIECE_TOKENS(D.gstart[g + 1]), TRC_LANE == 0);
  }
}
__global__ void k_build_items(TrcDev D) {
  GSTRIDE(g, D.m) {
    trc_build_items(D, g);
    int64_t t0 = D.item_off[g];
    int32_t a = D.gstart[g], b = D.gstart[g + 1];
    if (D.it_sub[t0] < 0) { for (int32_t j = a; j < b; ++j) D.pc_item[j] = (int32_t)t0; }
    else {
      int32_t cnt = D.nitems_g[g];
      for (int32_t j = a; j < b; ++j) D.pc_item[j] = -1;
      for (int32_t k = 0; k < cnt; ++k) {
        const TrcSub& s = D.subs[D.it_sub[t0 + k]];
        if (s.kind == 0) for (int32_t j = s.r0; j <= s.r1; ++j) D.pc_item[j] = (int32_t)(t0 + k);
      }
    }
  }
}
#ifndef TRC_HOSTSIM
// the same with a warp per merged segment (few segments of hundreds of pieces: config 4): the lanes fill pc_item
__global__ void k_build_items_w(TrcDev D) {
  const int lane = threadIdx.x & 31;
  WSTRIDE(g, D.m) {
    if (lane == 0) trc_build_items(D, g);
    __syncwarp();
    const int64_t t0 = D.item_off[g];
    const int32_t a = D.gstart[g], b = D.gstart[g + 1];
    if (D.it_sub[t0] < 0) { for (int32_t j = a + lane; j < b; j += 32) D.pc_item[j] = (int32_t)t0; }
    else {
      const int32_t cnt = D.nitems_g[g];
      for (int32_t j = a + lane; j < b; j += 32) D.pc_item[j] = -1;
      __syncwarp();
      for (int32_t k = 0; k < cnt; ++k) {
        const TrcSub& sb = D.subs[D.it_sub[t0 + k]];
        if (sb.kind == 0) for (int32_t j = sb.r0 + lane; j <= sb.r1; j += 32) D.pc_item[j] = (int32_t)(t0 + k);
      }
    }
  }
}
#endif
__global__ void k_forced_list(TrcDev D, const int32_t* __restrict__ large_flag, const int64_t* __restrict__ large_scan, int64_t* forced_pos) {
  GSTRIDE(g, D.m) if (large_flag[g]) forced_pos[large_scan[g]] = D.item_off[g];
}
__global__ void k_large_flag(TrcDev D, int32_t* large_flag) {
  GSTRIDE(g, D.m) large_flag[g] = D.gtok[g] > (int64_t)D.doc_E[D.sdoc[D.gstart[g]]];
}
__global__ void k_pack_next(TrcDev D, const int64_t* __restrict__ forced_pos, const int64_t* __restrict__ n_forced) {
  GSTRIDE(a, D.n_items) trc_pack_next(D, forced_pos, n_forced, a);
}
__global__ void k_chunk_record(TrcDev D) { GSTRIDE(c, D.n_chunks) trc_chunk_record(D, c); }
__global__ void k_chunk_header(TrcDev D) { GSTRIDE(c, D.n_chunks) trc_chunk_header_write(D, c); }
__global__ void k_emit_chunk_pieces(TrcDev D) { WSTRIDE(j, D.n_kept) trc_emit_chunk_piece(D, j, TRC_LANE, TRC_NL); }
__global__ void k_item_layout(TrcDev D) { GSTRIDE(t, D.n_items) trc_item_layout(D, t); }
#ifdef TRC_HOSTSIM
__global__ void k_emit_pieces_flat(TrcDev D) { GSTRIDE(j, D.n_kept) trc_emit_piece_flat(D, j, 0, 1); }
#else
// ---- chunk records, headers and item prefixes, CUDA build ---------------------------------------------------------------
// The per-chunk and per-item routines of trc_stages.cuh are chains of dependent loads (item -> group -> piece -> raw segment
// -> value); a thread per chunk walked its ~15 items one after the other (profiles/r2s: 68 + 43 + 69 us for 11 k chunks and
// 165 k items).  Here a WARP owns a chunk and its lanes take the items, and the loop over an item's pieces is gone: the
// item kernel leaves one 16-byte layout record per item and the copy kernel derives a piece's destination from it.
struct TrcItemLay {
  int64_t base;        // destination of piece j of the item = base + po[j]   (po = poff, or poffs with TRC_LAY_SUBS)
  int32_t p1;          // last piece of the item; -1: the item's text is written by the sentence stage
  uint32_t flags;
};
#define TRC_LAY_TS 1u
#define TRC_LAY_COMB 2u
#define TRC_LAY_SUBS 4u

__global__ void k_chunk_record_w(TrcDev D) {
  const int lane = threadIdx.x & 31;
  for (int64_t c = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5; c < D.n_chunks; c += ((int64_t)gridDim.x * blockDim.x) >> 5) {
    const int64_t a = D.cstart[c], b = D.cstart[c + 1];
    if (lane == 0) trc_chunk_record_head(D, c);
    // sorted distinct speaker ids: every round finds the smallest id above the last one found.  A lane keeps the
    // speakers of its first four items in registers (chunks of up to 128 items never look an item up twice).
    int32_t v[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) { const int64_t t = a + lane + 32 * q; v[q] = t < b ? trc_item_speaker(D, t) : 0x7fffffff; }
    int32_t* lst = D.spk_list + a;
    int32_t ns = 0, cur = -1;
    for (;;) {
      uint32_t best = 0x7fffffffu;
#pragma unroll
      for (int q = 0; q < 4; ++q) if (v[q] > cur && (uint32_t)v[q] < best) best = (uint32_t)v[q];
      for (int64_t t = a + lane + 128; t < b; t += 32) { const int32_t w = trc_item_speaker(D, t); if (w > cur && (uint32_t)w < best) best = (uint32_t)w; }
      best = __reduce_min_sync(0xffffffffu, best);
      if (best == 0x7fffffffu) break;
      if (lane == 0) lst[ns] = (int32_t)best;
      ++ns; cur = (int32_t)best;
    }
    __syncwarp();
    if (lane == 0) trc_chunk_record_tail(D, c, ns);
  }
}

// One launch, two roles.  Blocks [0, hb): a warp per chunk writes the chunk's header.  Blocks [hb, ...): a thread per item
// works out where the item's text goes (it_dst, from the chunk's offset, header length and the item-length prefix sums:
// nothing the header role writes), writes the "\n\n" in front of it and the "[ts] speaker: " prefix of whole segments,
// and leaves the layout record for k_emit_pieces_flat.
__global__ void k_headers_items(TrcDev D, TrcItemLay* __restrict__ lay, int hb) {
  if ((int)blockIdx.x < hb) {
    const int lane = threadIdx.x & 31;
    const int64_t c = (blockIdx.x * (int64_t)blockDim.x + threadIdx.x) >> 5;
    if (c >= D.n_chunks || lane != 0) return;
    TrcChunk& ch = D.chunks[c];
    ch.twc_off = D.twc_scan[c];
    if (D.add_context) {
      uint8_t* dst = D.twc + ch.twc_off;
      const int64_t o = trc_header<true>(D, ch, dst);
      dst[o] = '\n'; dst[o + 1] = '\n';
    }
    return;
  }
  const int64_t t = ((int64_t)blockIdx.x - hb) * blockDim.x + threadIdx.x;
  if (t >= D.n_items) return;
  const int64_t c = D.cscanf[t + 1] - 1;
  const TrcChunk& ch = D.chunks[c];
  const int64_t dst_off = D.twc_scan[c] + (D.add_context ? ch.hdr_len + 2 : 0) + (D.L[t] - D.L[ch.item0]);
  D.it_dst[t] = dst_off;
  const int32_t g = D.it_group[t];
  const int32_t s = D.it_sub[t];
  TrcItemLay r; r.base = 0; r.p1 = -1; r.flags = 0;
  if (s >= 0 && D.subs[s].kind != 0) { lay[t] = r; return; }             // sentence / clause items: k_emit_sent_recs
  const int32_t ga = D.gstart[g], gb = D.gstart[g + 1];
  const int32_t p0 = s >= 0 ? D.subs[s].r0 : ga;
  const int32_t p1 = s >= 0 ? D.subs[s].r1 : gb - 1;
  const bool comb = gb - ga > 1;
  uint8_t* base = D.twc + dst_off;
  int64_t pre = 0;
  if (s < 0) {
    const int32_t rr = D.k2r[ga];
    const int32_t sp = D.spk[rr];
    const int64_t so = D.spk_off[sp], sl = D.spk_off[sp + 1] - so;
    uint8_t buf[TRC_TS_MAX + 4];
    const int tl = trc_fmt_ts(D.start[rr], buf + 1);
    buf[0] = '['; buf[tl + 1] = ']'; buf[tl + 2] = ' ';
    for (int k = 0; k < tl + 3; ++k) base[k] = buf[k];
    for (int64_t k = 0; k < sl; ++k) base[tl + 3 + k] = D.spk_bytes[so + k];
    base[tl + 3 + sl] = ':'; base[tl + 4 + sl] = ' ';
    pre = tl + 5 + sl;
  }
  if (t > ch.item0) { base[-2] = '\n'; base[-1] = '\n'; }
  const int64_t* po = s >= 0 ? D.poffs : D.poff;
  r.base = dst_off + pre - po[p0];
  r.p1 = p1;
  r.flags = ((comb && (D.preserve || s >= 0)) ? TRC_LAY_TS : 0u) | (comb ? TRC_LAY_COMB : 0u) | (s >= 0 ? TRC_LAY_SUBS : 0u);
  lay[t] = r;
}

// 8 lanes per piece.  The body is copied in 4-byte words aligned to the DESTINATION (two aligned source
// words + a funnel shift per destination word); head, tail and the "[ts] " prefix go byte by byte.
__global__ void k_emit_pieces_flat(TrcDev D, const TrcItemLay* __restrict__ lay) {
  for (int64_t q = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; (q >> 3) < D.n_kept; q += (int64_t)gridDim.x * blockDim.x) {
    const int64_t j = q >> 3;
    const int lane = (int)(q & 7);
    // everything that only depends on j is requested at once (the kernel is bound by dependent-load latency)
    const int32_t t = D.pc_item[j];
    const int64_t c0 = D.coff[j], c1 = D.coff[j + 1];
    const double st = D.start[D.k2r[j]];
    const int64_t pa = D.poff[j], pb = D.poffs == D.poff ? pa : D.poffs[j];
    if (t < 0) continue;                                     // (pdst stays -1: the piece belongs to a sentence-path item)
    const int4 lr = __ldg(reinterpret_cast<const int4*>(lay + t));
    const int32_t p1 = lr.z;
    if (p1 < 0) continue;
    const uint32_t lf = (uint32_t)lr.w;
    const int64_t d = (((int64_t)(uint32_t)lr.y << 32) | (int64_t)(uint32_t)lr.x) + ((lf & TRC_LAY_SUBS) ? pb : pa);
    const uint8_t f = (uint8_t)(((lf & TRC_LAY_TS) ? 1 : 0) | (((lf & TRC_LAY_COMB) && j < p1) ? 2 : 0));
    if (lane == 0) D.pdst[j] = d;
    uint8_t* dst = D.twc + d;
    int64_t o = 0;
    if (f & 1) {
      int64_t h; int m, sec;
      trc_ts_split(st, &h, &m, &sec);
      if (h <= 0) {                          // "[MM:SS] " is exactly eight bytes: every lane of the group writes one of them
        const unsigned long long pre = (unsigned long long)'[' | ((unsigned long long)('0' + m / 10) << 8) | ((unsigned long long)('0' + m % 10) << 16) |
                                       ((unsigned long long)':' << 24) | ((unsigned long long)('0' + sec / 10) << 32) |
                                       ((unsigned long long)('0' + sec % 10) << 40) | ((unsigned long long)']' << 48) | ((unsigned long long)' ' << 56);
        dst[lane] = (uint8_t)(pre >> (8 * lane));
        o = 8;
      } else {
        o = trc_ts_len(st) + 3;
        if (lane == 0) {
          uint8_t buf[TRC_TS_MAX + 4];
          int tl = trc_fmt_ts(st, buf + 1);
          buf[0] = '['; buf[tl + 1] = ']'; buf[tl + 2] = ' ';
          for (int k = 0; k < tl + 3; ++k) dst[k] = buf[k];
        }
      }
    }
    const int64_t len = c1 - c0;
    const uint8_t* src = D.C + c0;
    uint8_t* db = dst + o;
    if ((f & 2) && lane == 7) db[len] = ' ';
    if (len < 24) { for (int64_t k = lane; k < len; k += 8) db[k] = src[k]; continue; }
    const int head = (int)((4 - ((uintptr_t)db & 3)) & 3);            // bytes until db is 4-byte aligned
    if (lane < head) db[lane] = src[lane];
    const int64_t nw = (len - head) >> 2;                              // whole destination words
    const uint8_t* s0 = src + head;
    const unsigned sh = (unsigned)((uintptr_t)s0 & 3) * 8u;
    const uint32_t* sw = reinterpret_cast<const uint32_t*>((uintptr_t)s0 & ~(uintptr_t)3);
    uint32_t* dw = reinterpret_cast<uint32_t*>(db + head);
    for (int64_t k = lane; k < nw; k += 8) {
      const uint32_t a = sw[k], b = sh ? sw[k + 1] : 0u;
      dw[k] = __funnelshift_r(a, b, sh);
    }
    const int64_t done = head + (nw << 2);
    if (lane < len - done) db[done + lane] = src[done + lane];
  }
}
#endif

// =====================================================================================================
// host helpers
// =====================================================================================================
static cudaEvent_t prof_event(trc_ctx* ctx) {
  cudaEvent_t e = nullptr;
  if (!ctx->prof_pool.empty()) { e = ctx->prof_pool.back(); ctx->prof_pool.pop_back(); }
  else cudaEventCreate(&e);
  return e;
}
static void prof_begin(trc_ctx* ctx, cudaStream_t st, const char* name) {
  if (!ctx->prof) return;
  trc_ctx::ProfRec r; r.name = name; r.a = prof_event(ctx); r.b = prof_event(ctx);
  cudaEventRecord(r.a, st);
  ctx->prof_recs.push_back(r);
}
static void prof_end(trc_ctx* ctx, cudaStream_t st) {
  if (!ctx->prof) return;
  cudaEventRecord(ctx->prof_recs.back().b, st);
}
static void prof_collect(trc_ctx* ctx) {
  for (auto& r : ctx->prof_recs) {
    float ms = 0;
    cudaEventSynchronize(r.b);
    cudaEventElapsedTime(&ms, r.a, r.b);
    bool found = false;
    for (auto& a : ctx->prof_acc) if (a.name == r.name) { a.ms += ms; a.launches++; found = true; break; }
    if (!found) ctx->prof_acc.push_back({r.name, (double)ms, 1});
    ctx->prof_pool.push_back(r.a); ctx->prof_pool.push_back(r.b);
  }
  ctx->prof_recs.clear();
}
#ifdef TRC_HOSTSIM
#define LAUNCH(kern, n, block, ...) do { if ((n) > 0) { kern(__VA_ARGS__); ctx->launches++; } } while (0)
#else
#define LAUNCH(kern, n, block, ...)                                                  \
  do {                                                                               \
    if ((n) > 0) {                                                                   \
      prof_begin(ctx, st, #kern);                                                    \
      kern<<<grid_for((n), (block)), (block), 0, st>>>(__VA_ARGS__);                 \
      prof_end(ctx, st);                                                             \
      ctx->launches++;                                                               \
      CK(cudaGetLastError());                                                        \
    }                                                                                \
  } while (0)
#endif
#define LAUNCH_W(kern, nwarps, ...) LAUNCH(kern, (nwarps) * 32, 256, __VA_ARGS__)

// Orbit marking: a thread per document, or - few documents with long chains - a block per document that stages the
// links in shared memory (k_orbit_staged).
static int orbit_marks(trc_ctx* ctx, cudaStream_t st, const int64_t* range_off, const int32_t* nxt, int32_t* flag, int64_t n_doc, int64_t n_links) {
#ifndef TRC_HOSTSIM
  if (n_doc > 0 && n_doc <= 64 && n_links >= 2048 * n_doc) {
    static bool attr_done[64] = {false};
    if (!attr_done[ctx->device & 63]) {
      CK(cudaFuncSetAttribute(k_orbit_staged, cudaFuncAttributeMaxDynamicSharedMemorySize, ORBIT_SMEM * 4));
      attr_done[ctx->device & 63] = true;
    }
    prof_begin(ctx, st, "k_orbit_staged");
    k_orbit_staged<<<(unsigned)n_doc, 1024, ORBIT_SMEM * 4, st>>>(range_off, nxt, flag, n_doc);
    prof_end(ctx, st);
    ctx->launches++;
    CK(cudaGetLastError());
    return TRC_OK;
  }
#endif
#ifndef TRC_HOSTSIM
  if (n_links >= 8 * n_doc) { LAUNCH_W(k_orbit_warp, n_doc, range_off, nxt, flag, n_doc); return TRC_OK; }
#endif
  (void)n_links;
  LAUNCH(k_orbit, n_doc, 64, range_off, nxt, flag, n_doc);
  return TRC_OK;
}

// Small counter reads.  A tiny kernel stores the words into mapped pinned memory instead of a D2H copy: a copy
// would queue on the copy engine behind the bulk result transfers of other contexts working on the same GPU
// (several contexts take batches in turn in a serving loop, bench.py's end-to-end leg) and stall this one.
#ifndef TRC_HOSTSIM
__global__ void k_read_back(const uint32_t* __restrict__ src, uint32_t* __restrict__ dst, int nwords) {
  for (int k = threadIdx.x; k < nwords; k += blockDim.x) dst[k] = src[k];
}
#endif
static int read_back(trc_ctx* ctx, cudaStream_t st, const void* dsrc, size_t bytes, void* hdst) {
  if (bytes > 64 * sizeof(int64_t) || (bytes & 3)) { ctx->err = "read_back size"; return TRC_EINVAL; }
#ifdef TRC_HOSTSIM
  memcpy(hdst, dsrc, bytes);
#else
  k_read_back<<<1, 64, 0, st>>>((const uint32_t*)dsrc, (uint32_t*)ctx->d_pin, (int)(bytes >> 2));
  ctx->launches++;
  CK(cudaGetLastError());
  CK(cudaStreamSynchronize(st));
  memcpy(hdst, ctx->h_pin, bytes);
#endif
  return TRC_OK;
}

// Two 8-byte counters in one round trip (either pointer may be null: its value is then 0).
#ifndef TRC_HOSTSIM
__global__ void k_read_back2(const unsigned long long* a, const unsigned long long* b, unsigned long long* dst) {
  if (threadIdx.x == 0) dst[0] = a ? *a : 0ULL;
  if (threadIdx.x == 1) dst[1] = b ? *b : 0ULL;
}
#endif
static int read_back2(trc_ctx* ctx, cudaStream_t st, const void* a, const void* b, int64_t out[2]) {
#ifdef TRC_HOSTSIM
  out[0] = a ? *(const int64_t*)a : 0; out[1] = b ? *(const int64_t*)b : 0;
#else
  k_read_back2<<<1, 32, 0, st>>>((const unsigned long long*)a, (const unsigned long long*)b, (unsigned long long*)ctx->d_pin);
  ctx->launches++;
  CK(cudaGetLastError());
  CK(cudaStreamSynchronize(st));
  out[0] = ctx->h_pin[0]; out[1] = ctx->h_pin[1];
#endif
  return TRC_OK;
}

// Two 8-byte values and a pair of them in one round trip (the group / merged-text totals and the side stream's counters).
#ifndef TRC_HOSTSIM
__global__ void k_read_back4(const unsigned long long* a, const unsigned long long* b, const unsigned long long* c2, unsigned long long* dst) {
  if (threadIdx.x == 0) dst[0] = a ? *a : 0ULL;
  if (threadIdx.x == 1) dst[1] = b ? *b : 0ULL;
  if (threadIdx.x == 2) dst[2] = c2 ? c2[0] : 0ULL;
  if (threadIdx.x == 3) dst[3] = c2 ? c2[1] : 0ULL;
}
static int read_back4(trc_ctx* ctx, cudaStream_t st, const void* a, const void* b, const void* c2, int64_t out[4]) {
  k_read_back4<<<1, 32, 0, st>>>((const unsigned long long*)a, (const unsigned long long*)b, (const unsigned long long*)c2, (unsigned long long*)ctx->d_pin);
  ctx->launches++;
  CK(cudaGetLastError());
  CK(cudaStreamSynchronize(st));
  for (int k = 0; k < 4; ++k) out[k] = ctx->h_pin[k];
  return TRC_OK;
}
#endif
// Three counters in one round trip: a 4-byte and two 8-byte values (any pointer may be null: 0).
#ifndef TRC_HOSTSIM
__global__ void k_read_back3(const int32_t* a, const unsigned long long* b, const unsigned long long* c, unsigned long long* dst) {
  if (threadIdx.x == 0) dst[0] = a ? (unsigned long long)(long long)*a : 0ULL;
  if (threadIdx.x == 1) dst[1] = b ? *b : 0ULL;
  if (threadIdx.x == 2) dst[2] = c ? *c : 0ULL;
}
#endif
static int read_back3(trc_ctx* ctx, cudaStream_t st, const int32_t* a, const void* b, const void* c, int64_t out[3]) {
#ifdef TRC_HOSTSIM
  out[0] = a ? *a : 0; out[1] = b ? *(const int64_t*)b : 0; out[2] = c ? *(const int64_t*)c : 0;
#else
  k_read_back3<<<1, 32, 0, st>>>(a, (const unsigned long long*)b, (const unsigned long long*)c, (unsigned long long*)ctx->d_pin);
  ctx->launches++;
  CK(cudaGetLastError());
  CK(cudaStreamSynchronize(st));
  out[0] = ctx->h_pin[0]; out[1] = ctx->h_pin[1]; out[2] = ctx->h_pin[2];
#endif
  return TRC_OK;
}

template <class Tp>
static int to_device(trc_ctx* ctx, cudaStream_t st, const Tp* src, int64_t count, int memory, const Tp** out) {
  if (memory == TRC_MEM_DEVICE) { *out = src; return TRC_OK; }
  Tp* d;
  ALLOC(d, Tp, count);
  if (count > 0) CK(cudaMemcpyAsync(d, src, sizeof(Tp) * (size_t)count, cudaMemcpyHostToDevice, st));
  *out = d;
  return TRC_OK;
}

// Token counts of n units (device pointers).  Long spans are finished by k_bpe_long after one sync - or, with `defer`, by
// finish_long() once the caller has read the two counters (defer->ext_cnt, zeroed by the caller) with a round trip of its own.
struct DeferLong {
  unsigned long long* ext_cnt = nullptr; bool pending = false;
  LongEntry* wl = nullptr; int64_t cap = 0; const uint8_t* bytes = nullptr; const int64_t* ustart = nullptr; const int64_t* uend = nullptr;
  int vs = 0; uint32_t* counts = nullptr;
};
static int finish_long(trc_ctx* ctx, cudaStream_t st, DeferLong& dl, unsigned long long h0, unsigned long long h1);
static int count_units(trc_ctx* ctx, cudaStream_t st, const uint8_t* bytes, const int64_t* ustart, const int64_t* uend,
                       int64_t n, int64_t total_bytes, int vs, uint32_t* counts, int32_t* special, bool allow_sync_free, int64_t coop_min = 0,
                       DeferLong* defer = nullptr) {
  if (n <= 0) return TRC_OK;
  if (special) coop_min = 0;                   // (the cooperative counter does not look for special-token literals)
  if (coop_min > 0) { if (const char* e = getenv("TRCHUNK_COOP_MIN")) { const long v = atol(e); if (v > 0) coop_min = v; } }   // experiments
#ifdef TRC_HOSTSIM
  coop_min = 0;
#endif
  // a long span has more than TRC_SHORT_SPAN LOGICAL bytes, and with vs every unit is one byte longer than it is stored
  const int64_t cap = (total_bytes + (vs ? n : 0)) / (TRC_SHORT_SPAN + 1) + 1;
  LongEntry* wl; unsigned long long* wl_cnt;
  ALLOC(wl, LongEntry, cap);
#ifndef TRC_HOSTSIM
  if (defer && defer->ext_cnt) wl_cnt = defer->ext_cnt;
  else
#endif
  {
    defer = nullptr;
    ALLOC(wl_cnt, unsigned long long, 2);
    CK(cudaMemsetAsync(wl_cnt, 0, 16, st));
  }
  unsigned long long* coop_cnt = nullptr; int32_t* coop_owner = nullptr; int64_t* coop_base = nullptr;
  const int64_t coop_bound = total_bytes / 32 + n + 1;      // most 32-byte chunks the long units can have
  if (coop_min > 0) {
    ALLOC(coop_cnt, unsigned long long, 1); ALLOC(coop_owner, int32_t, coop_bound); ALLOC(coop_base, int64_t, n);
    CK(cudaMemsetAsync(coop_cnt, 0, 8, st));
  }
  LAUNCH(k_count_units, n, 128, ctx->T, ctx->d_byte_id, bytes, ustart, uend, n, vs, counts, special, wl, wl_cnt, cap, coop_min, coop_cnt, coop_owner, coop_base);
#ifndef TRC_HOSTSIM
  if (coop_min > 0) LAUNCH(k_count_units_coop, coop_bound, 128, ctx->T, ctx->d_byte_id, bytes, ustart, uend, vs, counts, wl, wl_cnt, cap, coop_cnt, coop_owner, coop_base);
#endif
  if (defer) {
    defer->pending = true; defer->wl = wl; defer->cap = cap; defer->bytes = bytes; defer->ustart = ustart; defer->uend = uend;
    defer->vs = vs; defer->counts = counts;
    return TRC_OK;
  }
  unsigned long long h[2];
  int rc = read_back(ctx, st, wl_cnt, 16, h);
  if (rc) return rc;
  (void)allow_sync_free;
#ifdef TRC_HOSTSIM
  ctx->counts[TRC_N_LONG_SPANS] += (int64_t)h[0];
  h[0] = 0;
#endif
  if ((int64_t)h[0] > cap) { ctx->err = "long-span work list overflow"; return TRC_EINVAL; }   // cannot happen: cap is the bound
  if (h[0] > 0) {
#ifndef TRC_HOSTSIM
    uint32_t* scratch;
    ALLOC(scratch, uint32_t, 4 * (int64_t)h[1]);
    int blocks = (int)(h[0] < 148 * 8 ? h[0] : 148 * 8);
    prof_begin(ctx, st, "k_bpe_long");
    k_bpe_long<<<blocks, LONGB_T, 0, st>>>(ctx->T, ctx->d_byte_id, bytes, ustart, uend, vs, counts, wl, (int64_t)h[0], scratch, (int64_t)h[1]);
    prof_end(ctx, st);
    ctx->launches++;
    CK(cudaGetLastError());
    ctx->counts[TRC_N_LONG_SPANS] += (int64_t)h[0];
#endif
  }
  return TRC_OK;
}
static int finish_long(trc_ctx* ctx, cudaStream_t st, DeferLong& dl, unsigned long long h0, unsigned long long h1) {
  dl.pending = false;
#ifndef TRC_HOSTSIM
  if ((int64_t)h0 > dl.cap) { ctx->err = "long-span work list overflow"; return TRC_EINVAL; }   // cannot happen: cap is the bound
  if (h0 > 0) {
    uint32_t* scratch;
    ALLOC(scratch, uint32_t, 4 * (int64_t)h1);
    const int blocks = (int)(h0 < 148 * 8 ? h0 : 148 * 8);
    prof_begin(ctx, st, "k_bpe_long");
    k_bpe_long<<<blocks, LONGB_T, 0, st>>>(ctx->T, ctx->d_byte_id, dl.bytes, dl.ustart, dl.uend, dl.vs, dl.counts, dl.wl, (int64_t)h0, scratch, (int64_t)h1);
    prof_end(ctx, st);
    ctx->launches++;
    CK(cudaGetLastError());
    ctx->counts[TRC_N_LONG_SPANS] += (int64_t)h0;
  }
#else
  (void)ctx; (void)st; (void)h0; (void)h1;
#endif
  return TRC_OK;
}

#include "trc_sent_kernels.cuh"

#ifndef TRC_HOSTSIM
// Pre-pass for segments larger than a fused tile (trc_giant.cuh).  Leaves *V empty when there are none.
static int giant_stage(trc_ctx* ctx, cudaStream_t st, int64_t braw, GiantView* V, TileCand cand, int64_t tile_cap, int64_t* stats, CutView* cut,
                       unsigned long long* cblock = nullptr) {
  TrcDev& D = ctx->D;
  const int64_t n = D.n;
  int rc;
  {                                            // most batches have none: one counting pass decides
    // (cblock: seven words of the caller, [5..6] = the long-span counters of the speaker strings, read in the same round trip)
    unsigned long long* gcnt = cblock;
    if (!gcnt) ALLOC(gcnt, unsigned long long, 7);
    CK(cudaMemsetAsync(gcnt, 0, cblock ? 40 : 56, st));
    LAUNCH(k_giant_any, n, 256, D, gcnt, cand, tile_cap);
    if ((rc = read_back(ctx, st, gcnt, 56, stats))) return rc;      // [0] oversized segments, [1..4] tiles that would not fit, per candidate
    if (stats[0] == 0) return TRC_OK;
  }
  int32_t* gflag; int64_t* gscan;
  ALLOC(gflag, int32_t, n); ALLOC(gscan, int64_t, n + 1);
  LAUNCH(k_giant_flag, n, 256, D, gflag);
  if ((rc = scan_excl(ctx, st, gflag, gscan, n))) return rc;
  const int64_t ng = stats[0];                  // (k_giant_any counted them: no second round trip)
  GiantState G{};
  int32_t* list; int64_t* len; int32_t* of_seg; int64_t* boff;
  ALLOC(list, int32_t, ng); ALLOC(len, int64_t, ng); ALLOC(of_seg, int32_t, n); ALLOC(boff, int64_t, ng + 1);
  LAUNCH(k_giant_list, n, 256, D, gflag, gscan, list, len, of_seg);
  if ((rc = scan_excl(ctx, st, len, boff, ng))) return rc;
  int64_t gb = 0, nt = 0;
  int32_t* tcnt = nullptr; int64_t* tpre = nullptr;
  if (cut) {                                   // (their bytes and the number of cut targets come back together)
    ALLOC(tcnt, int32_t, ng); ALLOC(tpre, int64_t, ng + 1);
    LAUNCH(k_cut_counts, ng, 128, len, ng, tcnt);
    if ((rc = scan_excl(ctx, st, tcnt, tpre, ng))) return rc;
    int64_t h2[2];
    if ((rc = read_back2(ctx, st, boff + ng, tpre + ng, h2))) return rc;
    gb = h2[0]; nt = h2[1];
  } else if ((rc = read_back(ctx, st, boff + ng, 8, &gb))) return rc;
  if (gb >= (1LL << 31)) { ctx->err = "more than 2 GiB of oversized segments in one batch"; return TRC_EUNSUPPORTED; }
  ctx->counts[TRC_N_GIANTS] = ng;
  if (cut) {                                   // generation 5: parts for the fused kernel instead of the byte-parallel pre-pass
    int64_t* cutpos;
    if (n + nt < (1LL << 31)) {
      int64_t* vseg_off; int32_t* vorig; uint8_t* vcont;
      ALLOC(cutpos, int64_t, nt); ALLOC(cut->bad, unsigned long long, 1);
      CK(cudaMemsetAsync(cut->bad, 0, 8, st));
      LAUNCH(k_cut_find, nt, 128, D, ctx->T, list, len, tpre, ng, nt, cutpos);
      ALLOC(vseg_off, int64_t, n + nt + 1); ALLOC(vorig, int32_t, n + nt); ALLOC(vcont, uint8_t, n + nt);
      LAUNCH(k_vseg_build, n, 256, D, gflag, gscan, tpre, cutpos, nt, vseg_off, vorig, vcont, tile_cap - 64, cut->bad);
      cut->nv = n + nt; cut->vseg_off = vseg_off; cut->vorig = vorig; cut->vcont = vcont;
      return TRC_OK;
    }
  }
  G.list = list; G.boff = boff; G.ng = ng; G.gb = gb; G.of_seg = of_seg;
  ALLOC(G.cls, uint8_t, gb); ALLOC(G.code, uint8_t, gb); ALLOC(G.fns, int32_t, ng); ALLOC(G.pk, int64_t, gb); ALLOC(G.opk, int64_t, gb + 1);
  ALLOC(G.toff, int64_t, ng + 1); ALLOC(G.chars, int32_t, ng); ALLOC(G.tok, uint32_t, ng); ALLOC(G.spw, long long, ng);
  LAUNCH(k_giant_classify, gb, 256, D, ctx->T, G);
  LAUNCH(k_giant_fns, ng, 64, G);
  LAUNCH(k_giant_markdup, gb, 256, D, G);
  LAUNCH(k_giant_code, gb, 256, D, G);
  if ((rc = scan_excl(ctx, st, G.pk, G.opk, gb))) return rc;
  int64_t packed = 0;
  if ((rc = read_back(ctx, st, G.opk + gb, 8, &packed))) return rc;
  const int64_t gc = packed & 0xffffffffLL;
  ALLOC(G.text, uint8_t, gc);
  LAUNCH(k_giant_offsets, ng + 1, 128, G);
  LAUNCH(k_giant_scatter, gb, 256, D, G);
  // words of the cleaned giants -> token counts
  int32_t* wflag; int64_t* wscan;
  ALLOC(wflag, int32_t, gc); ALLOC(wscan, int64_t, gc + 1);
  LAUNCH(k_giant_wflag, gc, 256, G, gc, wflag, G.toff);
  LAUNCH(k_giant_first, ng, 64, G, wflag);
  if ((rc = scan_excl(ctx, st, wflag, wscan, gc))) return rc;
  int64_t nw = 0;
  if (gc) { if ((rc = read_back(ctx, st, wscan + gc, 8, &nw))) return rc; }
  int64_t* ustart; int64_t* uend; int32_t* wgiant; uint8_t* wfirst; uint32_t* c0; int32_t* sp; int64_t* fs; int64_t* fe; uint32_t* c1;
  ALLOC(ustart, int64_t, nw); ALLOC(uend, int64_t, nw); ALLOC(wgiant, int32_t, nw); ALLOC(wfirst, uint8_t, nw);
  ALLOC(c0, uint32_t, nw); ALLOC(sp, int32_t, nw); ALLOC(fs, int64_t, ng); ALLOC(fe, int64_t, ng); ALLOC(c1, uint32_t, ng);
  LAUNCH(k_giant_words, gc, 256, G, gc, wflag, wscan, ustart);
  LAUNCH(k_giant_wends, nw, 256, G, nw, gc, ustart, uend, wgiant, wfirst);
  LAUNCH(k_giant_firstunits, ng, 64, G, wscan, uend, fs, fe);
  if ((rc = count_units(ctx, st, G.text, ustart, uend, nw, gc, 0, c0, sp, false))) return rc;
  if ((rc = count_units(ctx, st, G.text, fs, fe, ng, gc, 1, c1, nullptr, false))) return rc;
  // running sum of the per-word counts: totals per giant are differences, and so are the sentences of the chunk stage
  int32_t* wc; int64_t* wpre;
  ALLOC(wc, int32_t, nw); ALLOC(wpre, int64_t, nw + 1);
  LAUNCH(k_giant_wcount, nw, 256, G, nw, wgiant, c0, c1, wfirst, sp, wc);
  if ((rc = scan_excl(ctx, st, wc, wpre, nw))) return rc;
  LAUNCH(k_giant_totals, ng, 64, G, wscan, wpre);
  ctx->WS.of_seg = of_seg; ctx->WS.toff = G.toff; ctx->WS.wscan = wscan; ctx->WS.wpre = wpre;
  V->of_seg = of_seg; V->toff = G.toff; V->text = G.text; V->chars = G.chars; V->tok = G.tok; V->spw = G.spw;
  return TRC_OK;
}
#endif

// =====================================================================================================
// C ABI
// =====================================================================================================
extern "C" int trc_create(int device, trc_ctx** out) {
  if (!out) return TRC_EINVAL;
  *out = nullptr;
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev <= 0 || device < 0 || device >= ndev) return TRC_ECUDA;
  cudaDeviceProp prop;
  if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) return TRC_ECUDA;
  if (prop.major != 10) return TRC_ECUDA;   // built for sm_100a only
  if (cudaSetDevice(device) != cudaSuccess) return TRC_ECUDA;
  trc_ctx* ctx = new trc_ctx();
  ctx->device = device;
  if (cudaStreamCreateWithFlags(&ctx->own_stream, cudaStreamNonBlocking) != cudaSuccess ||
      cudaStreamCreateWithFlags(&ctx->aux_stream, cudaStreamNonBlocking) != cudaSuccess ||
      cudaEventCreateWithFlags(&ctx->ev_aux[0], cudaEventDisableTiming) != cudaSuccess ||
      cudaEventCreateWithFlags(&ctx->ev_aux[1], cudaEventDisableTiming) != cudaSuccess ||
      cudaHostAlloc((void**)&ctx->h_pin, 64 * sizeof(int64_t), cudaHostAllocMapped) != cudaSuccess ||
      cudaHostGetDevicePointer(&ctx->d_pin, ctx->h_pin, 0) != cudaSuccess) { delete ctx; return TRC_ECUDA; }
  for (int k = 0; k < 4; ++k) cudaEventCreate(&ctx->ev[k]);
  { const char* e = getenv("TRCHUNK_NO_FUSED"); ctx->use_fused = !(e && e[0] == '1'); }
  { const char* e = getenv("TRCHUNK_FUSED"); if (e && e[0] >= '2' && e[0] <= '5') ctx->fused_gen = e[0] - '0'; }
  { const char* e = getenv("TRCHUNK_NO_MEMO"); ctx->use_memo = !(e && e[0] == '1'); }
  *out = ctx;
  return TRC_OK;
}

extern "C" void trc_destroy(trc_ctx* ctx) {
  if (!ctx) return;
  cudaSetDevice(ctx->device);
  cudaDeviceSynchronize();
  ctx->arena.release();
  for (void* p : ctx->vocab_allocs) cudaFree(p);
  if (ctx->d_props) cudaFree(ctx->d_props);
  if (ctx->d_acls) cudaFree(ctx->d_acls);
  if (ctx->h_pin) cudaFreeHost(ctx->h_pin);
  if (ctx->d_memo) cudaFree(ctx->d_memo);
  for (int k = 0; k < 4; ++k) if (ctx->ev[k]) cudaEventDestroy(ctx->ev[k]);
  if (ctx->own_stream) cudaStreamDestroy(ctx->own_stream);
  if (ctx->aux_stream) cudaStreamDestroy(ctx->aux_stream);
  for (int k = 0; k < 2; ++k) if (ctx->ev_aux[k]) cudaEventDestroy(ctx->ev_aux[k]);
  delete ctx;
}

extern "C" const char* trc_last_error(trc_ctx* ctx) { return ctx ? ctx->err.c_str() : "null context"; }

extern "C" int trc_load_tables(trc_ctx* ctx, const uint8_t* props) {
  if (!ctx || !props) return TRC_EINVAL;
  CK(cudaSetDevice(ctx->device));
  if (!ctx->d_props) CK(cudaMalloc(&ctx->d_props, 0x110000));
  CK(cudaMemcpy(ctx->d_props, props, 0x110000, cudaMemcpyHostToDevice));
  ctx->T.props = ctx->d_props;
  {
    uint8_t acls[128];
    for (int c = 0; c < 128; ++c) acls[c] = (uint8_t)TRC_RS(props[c]);
    if (!ctx->d_acls) CK(cudaMalloc(&ctx->d_acls, 128));
    CK(cudaMemcpy(ctx->d_acls, acls, 128, cudaMemcpyHostToDevice));
    ctx->T.acls = ctx->d_acls;
  }
  ctx->have_tables = true;
  return TRC_OK;
}

template <class Tp>
static int upload_vec(trc_ctx* ctx, const std::vector<Tp>& v, const Tp** out) {
  void* d = nullptr;
  CK(cudaMalloc(&d, sizeof(Tp) * (v.size() ? v.size() : 1)));
  ctx->vocab_allocs.push_back(d);
  if (!v.empty()) CK(cudaMemcpy(d, v.data(), sizeof(Tp) * v.size(), cudaMemcpyHostToDevice));
  *out = (const Tp*)d;
  return TRC_OK;
}

extern "C" int trc_load_vocab(trc_ctx* ctx, const uint8_t* token_bytes, const int64_t* token_off, const int32_t* ranks,
                              int32_t n_tokens, const uint8_t* special_bytes, const int64_t* special_off, int32_t n_special) {
  if (!ctx || !token_bytes || !token_off || !ranks) return TRC_EINVAL;
  if (!ctx->have_tables) { ctx->err = "trc_load_tables must be called first"; return TRC_ENOVOCAB; }
  CK(cudaSetDevice(ctx->device));
  for (void* p : ctx->vocab_allocs) cudaFree(p);
  ctx->vocab_allocs.clear();
  ctx->have_vocab = false;
  TrcHostVocab V;
  std::string e = trc_build_host_vocab(token_bytes, token_off, ranks, n_tokens, special_bytes, special_off, n_special, V);
  if (!e.empty()) { ctx->err = e; return TRC_EINVAL; }
  int rc;
  const uint32_t* byte_id;
  if ((rc = upload_vec(ctx, V.tok_bytes, &ctx->T.tok_bytes))) return rc;
  if ((rc = upload_vec(ctx, V.tok_off, &ctx->T.tok_off))) return rc;
  if ((rc = upload_vec(ctx, V.tokhash, &ctx->T.tokhash))) return rc;
  if ((rc = upload_vec(ctx, V.pairhash, &ctx->T.pairhash))) return rc;
  if ((rc = upload_vec(ctx, V.byte_id, &byte_id))) return rc;
  if ((rc = upload_vec(ctx, V.sp_bytes, &ctx->T.sp_bytes))) return rc;
  if ((rc = upload_vec(ctx, V.sp_off, &ctx->T.sp_off))) return rc;
  if ((rc = upload_vec(ctx, V.tok8, &ctx->T.tok8))) return rc;
  if ((rc = upload_vec(ctx, V.tok16, &ctx->T.tok16))) return rc;
  if ((rc = upload_vec(ctx, V.hot8, &ctx->T.hot8_src))) return rc;
  ctx->T.tok8_mask = (uint32_t)V.tok8.size() - 1; ctx->T.tok16_mask = (uint32_t)(V.tok16.size() / 2) - 1;
  ctx->T.hot8_size = (uint32_t)V.hot8.size();
  for (int k = 0; k < 4; ++k) ctx->T.sp_first[k] = V.sp_first[k];
  {
    int nset = 0, which = 0x100;
    for (int b = 0; b < 256; ++b) if ((V.sp_first[b >> 6] >> (b & 63)) & 1ULL) { ++nset; which = b; }
    ctx->T.sp_single = (nset == 1 && which >= 1 && which < 0x7f) ? (uint32_t)which : 0x100u;
  }
  ctx->d_byte_id = (uint32_t*)byte_id;
  ctx->T.n_tok = V.n_tok; ctx->T.n_sp = V.n_sp;
  ctx->T.tok_max_len = 0;
  for (uint32_t r = 0; r < V.n_tok; ++r) { const uint32_t l = V.tok_off[r + 1] - V.tok_off[r]; if (l > ctx->T.tok_max_len) ctx->T.tok_max_len = l; }
  ctx->T.tokhash_mask = (uint32_t)V.tokhash.size() - 1;
  ctx->T.pairhash_mask = (uint32_t)V.pairhash.size() - 1;
  ctx->n_special = V.n_sp;
  // token counts of the 1110 digit groups and of " [" with the kernels themselves
  std::vector<uint8_t> pb; std::vector<int64_t> po;
  trc_digit_probe_strings(pb, po);
  int64_t np = (int64_t)po.size() - 1;
  std::vector<uint8_t> zero(1111, 0);
  const uint8_t* d_dc;
  if ((rc = upload_vec(ctx, zero, &d_dc))) return rc;
  ctx->T.dcount = d_dc; ctx->T.cnt_space_lbracket = 0; ctx->T.dcount_ready = 0;
  ctx->have_vocab = true;
  std::vector<uint32_t> cnt((size_t)np);
  rc = trc_count_tokens(ctx, pb.data(), po.data(), np, 0, TRC_MEM_HOST, cnt.data(), nullptr, nullptr);
  if (rc) { ctx->have_vocab = false; return rc; }
  std::vector<uint8_t> dc(1111);
  for (int64_t k = 0; k < 1110; ++k) dc[k] = (uint8_t)cnt[k];
  CK(cudaMemcpy((void*)d_dc, dc.data(), 1110, cudaMemcpyHostToDevice));
  ctx->T.cnt_space_lbracket = cnt[1110];
  ctx->T.dcount_ready = 1;
  return TRC_OK;
}

extern "C" int trc_count_tokens(trc_ctx* ctx, const uint8_t* bytes, const int64_t* off, int64_t n, int32_t leading_space,
                                int32_t memory, uint32_t* counts, int32_t* special, void* stream) {
  if (!ctx || !off || !counts || n < 0) return TRC_EINVAL;
  if (!ctx->have_vocab) { ctx->err = "vocabulary not loaded"; return TRC_ENOVOCAB; }
  CK(cudaSetDevice(ctx->device));
  cudaStream_t st = stream ? (cudaStream_t)stream : ctx->own_stream;
  if (n == 0) return TRC_OK;
  // this service uses its own short-lived allocations so that it never disturbs a resident batch
  Arena saved = ctx->arena;
  ctx->arena = Arena();
  int rc = TRC_OK;
  do {
    int64_t total = 0;
    const int64_t* d_off = nullptr; const uint8_t* d_bytes = nullptr;
    if (memory == TRC_MEM_HOST) total = off[n];
    else { int64_t tmp[1]; if ((rc = read_back(ctx, st, off + n, 8, tmp))) break; total = tmp[0]; }
    if ((rc = to_device(ctx, st, off, n + 1, memory, &d_off))) break;
    if ((rc = to_device(ctx, st, bytes, total, memory, &d_bytes))) break;
    uint32_t* d_cnt; int32_t* d_sp = nullptr;
    d_cnt = (uint32_t*)ctx->arena.alloc(sizeof(uint32_t) * (size_t)n);
    if (special) d_sp = (int32_t*)ctx->arena.alloc(sizeof(int32_t) * (size_t)n);
    if (!d_cnt || (special && !d_sp)) { ctx->err = "device allocation failed"; rc = TRC_ECUDA; break; }
    if ((rc = count_units(ctx, st, d_bytes, d_off, d_off + 1, n, total, leading_space ? 1 : 0, d_cnt, d_sp, false))) break;
    cudaMemcpyKind kind = memory == TRC_MEM_HOST ? cudaMemcpyDeviceToHost : cudaMemcpyDeviceToDevice;
    if (cudaMemcpyAsync(counts, d_cnt, sizeof(uint32_t) * (size_t)n, kind, st) != cudaSuccess ||
        (special && cudaMemcpyAsync(special, d_sp, sizeof(int32_t) * (size_t)n, kind, st) != cudaSuccess) ||
        cudaStreamSynchronize(st) != cudaSuccess) { ctx->err = "copy of token counts failed"; rc = TRC_ECUDA; break; }
  } while (0);
  ctx->arena.release();
  ctx->arena = saved;
  return rc;
}

static void set_arr(trc_ctx* ctx, int which, const void* p, int64_t bytes) { ctx->arr[which].p = p; ctx->arr[which].bytes = bytes; }

extern "C" int trc_preprocess(trc_ctx* ctx, const trc_batch* B, void* stream) {
  if (!ctx || !B) return TRC_EINVAL;
  if (!ctx->have_vocab) { ctx->err = "vocabulary not loaded"; return TRC_ENOVOCAB; }
  if (B->n_seg < 0 || B->n_doc < 0 || B->n_spk < 1 || B->n_seg >= (1LL << 31)) { ctx->err = "bad batch sizes"; return TRC_EINVAL; }
  CK(cudaSetDevice(ctx->device));
  cudaStream_t st = stream ? (cudaStream_t)stream : ctx->own_stream;
  ctx->arena.reset();
  ctx->have_pre = ctx->have_chunk = false;
  ctx->WS = TrcWordSums{};
  for (auto& a : ctx->arr) a = {nullptr, 0};
  for (auto& c : ctx->counts) c = 0;
  TrcDev& D = ctx->D;
  D = TrcDev{};
  int rc;
  CK(cudaEventRecord(ctx->ev[0], st));
  const int64_t n = B->n_seg, nd = B->n_doc;
  int64_t braw = 0, bspk = 0;
  const bool len32 = B->seg_len32 != nullptr && B->seg_off == nullptr;
  if (!len32 && !B->seg_off) { ctx->err = "seg_off (or seg_len32 + text_bytes) is required"; return TRC_EINVAL; }
  if (len32 && B->already_processed) { ctx->err = "seg_len32 is not accepted together with already_processed"; return TRC_EINVAL; }
  if (len32 && B->memory == TRC_MEM_HOST) { braw = B->text_bytes; bspk = B->spk_off[B->n_spk]; }
  else if (B->memory == TRC_MEM_HOST) { braw = n ? B->seg_off[n] : 0; bspk = B->spk_off[B->n_spk]; }
  else if (len32) {
    int64_t t2[2];
    if ((rc = read_back2(ctx, st, nullptr, B->spk_off + B->n_spk, t2))) return rc;
    braw = B->text_bytes; bspk = t2[1];
  } else {
    int64_t t2[2];
    if ((rc = read_back2(ctx, st, n ? B->seg_off + n : nullptr, B->spk_off + B->n_spk, t2))) return rc;
    braw = t2[0]; bspk = t2[1];
  }
  if ((rc = to_device(ctx, st, B->text, braw, B->memory, &D.text))) return rc;
  const int32_t* d_len32 = nullptr;
  if (len32) { if ((rc = to_device(ctx, st, B->seg_len32, n, B->memory, &d_len32))) return rc; }
  else if ((rc = to_device(ctx, st, B->seg_off, n + 1, B->memory, &D.seg_off))) return rc;
  if ((rc = to_device(ctx, st, B->start, n, B->memory, &D.start))) return rc;
  if ((rc = to_device(ctx, st, B->end, n, B->memory, &D.end))) return rc;
  if ((rc = to_device(ctx, st, B->speaker, n, B->memory, &D.spk))) return rc;
  if ((rc = to_device(ctx, st, B->doc_off, nd + 1, B->memory, &D.doc_off))) return rc;
  if ((rc = to_device(ctx, st, B->spk_bytes, bspk, B->memory, &D.spk_bytes))) return rc;
  if ((rc = to_device(ctx, st, B->spk_off, (int64_t)B->n_spk + 1, B->memory, &D.spk_off))) return rc;
  if (len32) {     // lengths in, offsets by a prefix sum - after the last upload, so that the copies stay back to back
    int64_t* soff;
    ALLOC(soff, int64_t, n + 1);
    if ((rc = scan_excl(ctx, st, d_len32, soff, n))) return rc;
    D.seg_off = soff;
  }
  D.n = n; D.n_doc = nd; D.n_spk = B->n_spk;
  D.merge = B->merge_same_speaker; D.has_cap = B->has_max_duration; D.cap = B->max_duration; D.preserve = B->preserve_timestamps;
  ctx->counts[TRC_N_SEG] = n; ctx->counts[TRC_N_DOC] = nd; ctx->counts[TRC_B_RAW] = braw;

  // speaker units: count(" " + speaker + ":") and special-token check of the speaker strings
  uint8_t* spk_ub; int64_t* spk_uo; uint32_t* spk_tok; int32_t* spk_special;
  ALLOC(spk_ub, uint8_t, bspk + B->n_spk); ALLOC(spk_uo, int64_t, B->n_spk + 1);
  ALLOC(spk_tok, uint32_t, B->n_spk); ALLOC(spk_special, int32_t, B->n_spk);
  LAUNCH(k_build_spk_units, (int64_t)B->n_spk, 128, D.spk_bytes, D.spk_off, B->n_spk, spk_ub, spk_uo);
  // (the check for long spans in speaker strings waits for the next host round trip: the pre-pass of the fused kernel)
  DeferLong spk_long;
  unsigned long long* cblock = nullptr;
#ifndef TRC_HOSTSIM
  ALLOC(cblock, unsigned long long, 7);
  CK(cudaMemsetAsync(cblock, 0, 56, st));
  spk_long.ext_cnt = cblock + 5;
#endif
  if ((rc = count_units(ctx, st, spk_ub, spk_uo, spk_uo + 1, B->n_spk, bspk + B->n_spk, 1, spk_tok, spk_special, false, 0, &spk_long))) return rc;
  D.spk_tok = spk_tok;

  int64_t nk = 0, bclean = 0;
  bool slow_pending = false;                 // k_count_slow_words in flight on the side stream
  unsigned long long j_h0 = 0, j_h1 = 0; unsigned long long* j_fz = nullptr; LongEntry* j_wl = nullptr; int64_t j_wl_cap = 0;
  (void)slow_pending; (void)j_h0; (void)j_h1; (void)j_fz; (void)j_wl; (void)j_wl_cap;
  int32_t* seg_special = nullptr;
  ALLOC(D.doc_err, int32_t, nd); ALLOC(D.doc_err_seg, int64_t, nd);
  CK(cudaMemsetAsync(D.doc_err, 0, 4 * (size_t)(nd > 0 ? nd : 1), st));
  LAUNCH(k_fill_i64, nd, 256, D.doc_err_seg, nd, (int64_t)INT64_MAX);
  const bool modeB = B->already_processed != 0;
  if (modeB) {
    if (B->group_off && (!B->group_start || !B->group_end || !B->full_text || !B->full_off || B->n_groups < 0)) {
      ctx->err = "already_processed with group_off needs group_start, group_end, full_text and full_off"; return TRC_EINVAL;
    }
    nk = n; bclean = braw; D.n_kept = n;
    D.C = const_cast<uint8_t*>(D.text); D.coff = const_cast<int64_t*>(D.seg_off); D.kdoc_off = const_cast<int64_t*>(D.doc_off);
    ALLOC(D.cchars, int32_t, n); ALLOC(D.k2r, int32_t, n); ALLOC(D.sdoc, int32_t, n); ALLOC(D.ctok, uint32_t, n); ALLOC(seg_special, int32_t, n);
    LAUNCH(k_identity_pieces, n, 128, D);
    if ((rc = count_units(ctx, st, D.C, D.coff, D.coff + 1, n, braw, 1, D.ctok, seg_special, false))) return rc;
    ctx->counts[TRC_N_KEPT] = nk; ctx->counts[TRC_B_CLEAN] = bclean;
  } else
#ifndef TRC_HOSTSIM
  if (ctx->use_fused) {
    // P1+P2 fused: single pass, outputs sized by upper bounds (cleaned text <= 1.5 x raw, kept <= n)
    ALLOC(D.cchars, int32_t, n); ALLOC(D.k2r, int32_t, n); ALLOC(D.coff, int64_t, n + 1); ALLOC(D.sdoc, int32_t, n);
    ALLOC(D.ctok, uint32_t, n); ALLOC(seg_special, int32_t, n); ALLOC(D.kdoc_off, int64_t, nd + 1);
    const int64_t ccap = braw + braw / 2 + 64;
    ALLOC(D.C, uint8_t, ccap);
    if (n == 0) { CK(cudaMemsetAsync(D.coff, 0, 8, st)); CK(cudaMemsetAsync(D.kdoc_off, 0, 8 * (size_t)(nd + 1), st)); }
    else {
      // generation 4 needs a 16-byte aligned text buffer (bulk copies); its tiles hold as many segments as fit a
      // warp's buffers at this batch's mean segment length (at most 32, one per lane)
      int gen = ctx->fused_gen;
      if (gen >= 4 && ((uintptr_t)D.text & 15)) gen = 3;
      int tsegs4 = F4_SEGS;
      if (gen == 4) {
        const double avg = (double)braw / (double)n;
        const double fit = (double)(F4_IN_CAP - 16) / (avg * 1.12 + 1.0);
        tsegs4 = fit >= F4_SEGS ? F4_SEGS : (fit < 1.0 ? 1 : (int)fit);
      } else if (gen >= 5) {
        const double avg = (double)braw / (double)n;
        const double fit = (double)(F5_IN_CAP - 16) / (avg * 1.06 + 1.0);
        tsegs4 = fit >= F5_SEGS ? F5_SEGS : (fit < 1.0 ? 1 : (int)fit);
      }
      // The pre-pass for oversized segments also says how many tiles of a few candidate sizes would not fit the fused
      // kernel's tile buffer.  Such a tile takes the kernel's slow lane, and while its thread-per-segment cleaning runs
      // every later tile waits for its byte count in the look-back (measured: 87 such tiles of 5 900 cost 1.7 ms of
      // 2.2).  So generation 5 takes the largest candidate that fits everywhere, and when none does - oversized
      // segments, heavy-tailed lengths - tiles of varying size (k_tile_flags).
      FuseArgs A;
      A.G = GiantView{};
      A.tile_start = nullptr; A.vorig = nullptr; A.vcont = nullptr; A.special_pos = nullptr; A.wtok = nullptr;
      TrcDev Dk = D;                         // the batch as the fused kernel sees it (cut mode: parts in place of oversized segments)
      int64_t ntiles_var = -1;
      if (gen >= 3) {
        TileCand cand{{0, 0, 0, 0}, 0};
        int64_t stats[7] = {0, 0, 0, 0, 0, 0, 0};
        const char* forced = gen >= 5 ? getenv("TRCHUNK_TSEGS") : nullptr;          // experiments: a fixed tile size
        if (forced) { const int v = atoi(forced); if (v >= 1 && v <= F5_SEGS) tsegs4 = v; else forced = nullptr; }
        if (gen >= 5 && !forced) {
          for (int q = 0; q < 4; ++q) { cand.v[q] = tsegs4 * (8 - q) / 8; if (cand.v[q] < 1) cand.v[q] = 1; }
          if (tsegs4 >= 16) { cand.all8 = 1; for (int q = 0; q < 4; ++q) cand.v[q] &= ~7; }
        }
        const char* cut_env = getenv("TRCHUNK_GIANT_CUT");                           // "0": keep the byte-parallel pre-pass (tests)
        const bool want_cut = gen >= 5 && !forced && !(cut_env && atoi(cut_env) == 0);
        CutView CV{};
        if ((rc = giant_stage(ctx, st, braw, &A.G, cand, F5_IN_CAP, stats, want_cut ? &CV : nullptr, cblock))) return rc;
        if (spk_long.pending) { if ((rc = finish_long(ctx, st, spk_long, (unsigned long long)stats[5], (unsigned long long)stats[6]))) return rc; }
        if (gen >= 5 && !forced) {
          int pick = -1;
          if (stats[0] == 0) for (int q = 0; q < 4 && pick < 0; ++q) if (stats[1 + q] == 0) pick = q;
          if (pick >= 0) tsegs4 = cand.v[pick];
          else {
            static_assert(F5_IN_CAP > TRC_GIANT_MIN + 16, "a window of the varying tiles");
            auto var_tiles = [&](const TrcDev& Dx, int64_t iso, const unsigned long long* word, int64_t* word_out) -> int {
              int32_t* tflag; int64_t* tscan; int32_t* tstart;
              ALLOC(tflag, int32_t, Dx.n); ALLOC(tscan, int64_t, Dx.n + 1);
              LAUNCH(k_tile_flags, Dx.n, 256, Dx, tflag, (int64_t)F5_IN_CAP - iso - 16, (int)F5_SEGS, iso);
              int r2;
              if ((r2 = scan_excl(ctx, st, tflag, tscan, Dx.n))) return r2;
              int64_t h2[2] = {0, 0};
              if ((r2 = read_back2(ctx, st, tscan + Dx.n, word, h2))) return r2;
              ntiles_var = h2[0];
              if (word_out) *word_out = h2[1];
              ALLOC(tstart, int32_t, ntiles_var + 1);
              LAUNCH(k_scatter_starts, Dx.n, 256, tflag, tscan, tstart, Dx.n);
              A.tile_start = tstart;
              return TRC_OK;
            };
            bool cut_done = false;
            if (CV.nv > 0) {
              Dk.seg_off = CV.vseg_off; Dk.n = CV.nv;
              int64_t bad = 0;
              if ((rc = var_tiles(Dk, TRC_CUT_ISO, CV.bad, &bad))) return rc;
              if (getenv("TRCHUNK_CUT_DEBUG")) fprintf(stderr, "[trchunk] cut: %lld segments -> %lld, targets without a cut: %lld, tiles %lld\n", (long long)n, (long long)CV.nv, (long long)bad, (long long)ntiles_var);
              if (bad == 0) {
                cut_done = true;
                A.vorig = CV.vorig; A.vcont = CV.vcont;
                ALLOC(A.special_pos, unsigned long long, n);
                CK(cudaMemsetAsync(A.special_pos, 0xff, 8 * (size_t)n, st));
                CK(cudaMemsetAsync(D.ctok, 0, 4 * (size_t)n, st));
                CK(cudaMemsetAsync(D.cchars, 0, 4 * (size_t)n, st));
                if (ctx->use_memo) {         // per-word token counts for the sentence stage (the slow words' kernels fill theirs in)
                  ALLOC(A.wtok, uint8_t, ccap + 8);
                  CK(cudaMemsetAsync(A.wtok, 0, (size_t)ccap + 8, st));
                }
              } else {                       // a segment that cannot be cut: the byte-parallel pre-pass for this batch
                Dk = D;
                if ((rc = giant_stage(ctx, st, braw, &A.G, TileCand{{0, 0, 0, 0}, 0}, F5_IN_CAP, stats, nullptr))) return rc;
              }
            }
            if (!cut_done) { if ((rc = var_tiles(D, TRC_GIANT_MIN, nullptr, nullptr))) return rc; }
          }
        }
      }
      const int64_t tile_segs = gen >= 4 ? tsegs4 : (gen >= 3 ? F3_SEGS : FUSE_SEGS);
      const int64_t ntiles = ntiles_var >= 0 ? ntiles_var : (n + tile_segs - 1) / tile_segs;
      unsigned long long* fz;
      const int64_t ngroups = (ntiles + 31) / 32;
      ALLOC(fz, unsigned long long, 2 * ntiles + 3 * ngroups + 32);
      CK(cudaMemsetAsync(fz, 0, 8 * (size_t)(2 * ntiles + 3 * ngroups + 32), st));
      A.tile_counter = fz; A.wl_cnt = fz + 2; A.totals = (int64_t*)(fz + 4); A.desc_bytes = fz + 32; A.desc_kept = fz + 32 + ntiles;
      A.wstat = fz + 24;
      A.grp_cnt = fz + 32 + 2 * ntiles; A.grp_bytes = A.grp_cnt + ngroups; A.grp_kept = A.grp_bytes + ngroups;
      static const bool want_phases = getenv("TRCHUNK_PHASES") != nullptr;
      A.phase = want_phases ? fz + 8 : nullptr;
      A.text_bytes = braw; A.seg_special = seg_special;
      A.wl_cap = (ccap + n) / (TRC_SHORT_SPAN + 1) + 1;          // logical bytes: every kept segment carries a virtual leading space
      ALLOC(A.wl, LongEntry, A.wl_cap);
      // words that need the generic matcher: open-vocabulary text sends a quarter of its words there, so the list is sized
      // by the most words the text can hold (a word and its separator take at least three bytes on average); past it,
      // tiles count their words in place
      A.sl_cap = braw / 3 + 2 * n + 1024; A.sl_cnt = fz + 7;
      ALLOC(A.sl, SlowWord, A.sl_cap);
      A.memo = nullptr; A.memo_lock = nullptr; A.memo_mask = 0;
      if (ctx->use_memo && gen >= 3) {
        const size_t ents = (size_t)1 << TRC_MEMO_LOG2, bytes = ents * 16 + ents * 4;
        if (!ctx->d_memo) CK(cudaMalloc(&ctx->d_memo, bytes));
        CK(cudaMemsetAsync(ctx->d_memo, 0, bytes, st));
        A.memo = ctx->d_memo; A.memo_lock = (uint32_t*)(ctx->d_memo + 2 * ents); A.memo_mask = (uint32_t)(ents - 1);
      }
      static bool attr_done[64] = {false};
      if (!attr_done[ctx->device & 63]) {
        CK(cudaFuncSetAttribute(k_clean_tok_fused, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(FuseSmem)));
        CK(cudaFuncSetAttribute(k_clean_tok_fused3, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(F3Smem)));
        CK(cudaFuncSetAttribute(k_clean_tok_w32, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(F4Smem)));
        CK(cudaFuncSetAttribute(k_clean_tok_c32<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(F5Smem)));
        CK(cudaFuncSetAttribute(k_clean_tok_c32<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(F5Smem)));
        attr_done[ctx->device & 63] = true;
      }
      int blocks = (int)(ntiles < 148 * 2 ? ntiles : 148 * 2);
      if (gen >= 5) {
        blocks = (int)(ntiles < 148 * F5_CTAS ? ntiles : 148 * F5_CTAS);
        prof_begin(ctx, st, "k_clean_tok_c32");
        if (A.vcont) k_clean_tok_c32<true><<<blocks, F5_T, sizeof(F5Smem), st>>>(Dk, ctx->T, ctx->d_byte_id, A, ntiles, tsegs4);
        else k_clean_tok_c32<false><<<blocks, F5_T, sizeof(F5Smem), st>>>(Dk, ctx->T, ctx->d_byte_id, A, ntiles, tsegs4);
      } else if (gen >= 4) {
        const int64_t want = (ntiles + F4_WARPS - 1) / F4_WARPS;
        blocks = (int)(want < 148 ? want : 148);
        prof_begin(ctx, st, "k_clean_tok_w32");
        k_clean_tok_w32<<<blocks, F4_T, sizeof(F4Smem), st>>>(D, ctx->T, ctx->d_byte_id, A, ntiles, tsegs4);
      } else if (gen >= 3) {
        prof_begin(ctx, st, "k_clean_tok_fused3");
        k_clean_tok_fused3<<<blocks, F3_T, sizeof(F3Smem), st>>>(D, ctx->T, ctx->d_byte_id, A, ntiles);
      } else {
        prof_begin(ctx, st, "k_clean_tok_fused");
        k_clean_tok_fused<<<blocks, FUSE_T, sizeof(FuseSmem), st>>>(D, ctx->T, ctx->d_byte_id, A, ntiles);
      }
      prof_end(ctx, st);
      ctx->launches++;
      CK(cudaGetLastError());
      unsigned long long h[27];
      if ((rc = read_back(ctx, st, fz + 2, 216, h))) return rc;     // wl count, wl scratch, n_kept, bclean, -, slow words, ..., word statistics
      ctx->counts[TRC_N_WORDS] = (int64_t)h[22]; ctx->counts[TRC_N_WORD_FIRST] = (int64_t)h[23]; ctx->counts[TRC_N_WORD_SLOW] = (int64_t)h[24];
      nk = (int64_t)h[2]; bclean = (int64_t)h[3];
      D.n_kept = nk;
      if (A.vcont) {
        if (h[25]) { ctx->err = "internal: a tile of a batch with cut segments took the slow lane"; return TRC_EINVAL; }
        LAUNCH(k_special_from_pos, nk, 256, A.special_pos, nk, seg_special);
        if (A.wtok && h[26] == 0) ctx->WS.wtok = A.wtok;          // (h[26]: tiles that counted per segment and left no per-word counts)
      }
      {
        int64_t nslow = (int64_t)h[5] < A.sl_cap ? (int64_t)h[5] : A.sl_cap;   // entries past the capacity were counted in the kernel
        if (nslow > 0) {       // on the side stream: nothing before the per-group token totals needs these counts
          CK(cudaEventRecord(ctx->ev_aux[0], st));
          CK(cudaStreamWaitEvent(ctx->aux_stream, ctx->ev_aux[0], 0));
          if (gen >= 5 && A.memo) {              // the generic matcher once per distinct word (claim), then the other occurrences
            unsigned long long* work; unsigned long long* deferred; unsigned long long* lcnt;
            ALLOC(work, unsigned long long, nslow); ALLOC(deferred, unsigned long long, nslow); ALLOC(lcnt, unsigned long long, 2);
            CK(cudaMemsetAsync(lcnt, 0, 16, ctx->aux_stream));
            prof_begin(ctx, ctx->aux_stream, "k_slow_words_claim");
            k_slow_words_claim<<<grid_for(nslow, 256), 256, 0, ctx->aux_stream>>>(D, A, A.sl, nslow, work, deferred, lcnt);
            prof_end(ctx, ctx->aux_stream);
            prof_begin(ctx, ctx->aux_stream, "k_slow_words_count");
            k_slow_words_count<<<grid_for(nslow, 128), 128, 0, ctx->aux_stream>>>(D, ctx->T, ctx->d_byte_id, A, A.sl, work, lcnt);
            prof_end(ctx, ctx->aux_stream);
            prof_begin(ctx, ctx->aux_stream, "k_slow_words_follow");
            k_slow_words_follow<<<grid_for(nslow, 256), 256, 0, ctx->aux_stream>>>(D, ctx->T, ctx->d_byte_id, A, A.sl, deferred, lcnt);
            prof_end(ctx, ctx->aux_stream);
            ctx->launches += 3;
            ctx->launches += 2;
          } else {
            prof_begin(ctx, ctx->aux_stream, "k_count_slow_words");
            k_count_slow_words<<<grid_for(nslow, 128), 128, 0, ctx->aux_stream>>>(D, ctx->T, ctx->d_byte_id, A.sl, nslow, A.wl, A.wl_cnt, A.wl_cap);
            prof_end(ctx, ctx->aux_stream);
            ctx->launches++;
          }
          CK(cudaGetLastError());
          slow_pending = true;
        }
      }
      if (want_phases && gen == 4) {
        unsigned long long ph[16];
        if ((rc = read_back(ctx, st, fz + 8, 16 * 8, ph))) return rc;
        unsigned long long tot = 0; for (int k = 0; k < 13; ++k) tot += ph[k];
        static const char* nm[13] = {"wait+load", "classify", "dup", "codes+scan", "segments", "scatter+wordstarts", "special", "tok-cache", "tok-match",
                                     "resolve", "copy-out", "slow-handoff+sums", "outputs"};
        fprintf(stderr, "[trchunk w32 phases %%]");
        for (int k = 0; k < 13; ++k) fprintf(stderr, " %s %.1f", nm[k], 100.0 * ph[k] / (tot ? tot : 1));
        fprintf(stderr, "\n");
      } else if (want_phases) {
        unsigned long long ph[16];
        if ((rc = read_back(ctx, st, fz + 8, 16 * 8, ph))) return rc;
        fprintf(stderr, "[trchunk words] %llu in word lists, %llu past the word cache, %llu to the slow list, %llu matcher words longer than 15 bytes\n", ph[12], ph[13], ph[14], ph[15]);
        unsigned long long tot = 0; for (int k = 0; k < 12; ++k) tot += ph[k];
        fprintf(stderr, "[trchunk phases %%] load %.1f  classify %.1f  dup %.1f  codes+scan %.1f  publish %.1f  scatter %.1f  words %.1f  tok-cache %.1f  tok-match %.1f  resolve+copy %.1f  tok-slow %.1f  outputs %.1f\n",
                100.0 * ph[0] / tot, 100.0 * ph[1] / tot, 100.0 * ph[2] / tot, 100.0 * ph[3] / tot, 100.0 * ph[4] / tot, 100.0 * ph[5] / tot,
                100.0 * ph[6] / tot, 100.0 * ph[11] / tot, 100.0 * ph[9] / tot, 100.0 * ph[10] / tot, 100.0 * ph[7] / tot, 100.0 * ph[8] / tot);
      }
      j_h0 = h[0]; j_h1 = h[1]; j_fz = fz; j_wl = A.wl; j_wl_cap = A.wl_cap;      // long spans: launched at the join below
      LAUNCH(k_doc_ranges_stream, nk + 1, 256, D.sdoc, nk, D.kdoc_off, nd);
    }
    ctx->counts[TRC_N_KEPT] = nk; ctx->counts[TRC_B_CLEAN] = bclean;
  } else
#endif
  {
  // P1: clean (count -> scan -> write)
  int32_t* keep;
  ALLOC(D.clen, int64_t, n); ALLOC(D.cchars, int32_t, n); ALLOC(keep, int32_t, n);
  ALLOC(D.kscan, int64_t, n + 1); ALLOC(D.cscan, int64_t, n + 1);
#ifdef TRC_HOSTSIM
  uint8_t* hs_cls = nullptr; uint8_t* hs_code = nullptr;
  const char* hs_seq = getenv("TRCHUNK_HOSTSIM_SEQ");
  const char* hs_mode = getenv("TRCHUNK_HOSTSIM_CLEAN");          // "swar" (default) | "par" | "seq"
  const bool hs_par = !(hs_seq && hs_seq[0] == '1') && !(hs_mode && hs_mode[0] == 's' && hs_mode[1] == 'e');
  if (hs_par) {
    ALLOC(hs_cls, uint8_t, braw); ALLOC(hs_code, uint8_t, braw);
    if (hs_mode && hs_mode[0] == 'p') hostsim_clean_par_count(D, ctx->T, keep, hs_cls, hs_code);
    else if (hs_mode && hs_mode[0] == 'w') { const char* ts = getenv("TRCHUNK_HOSTSIM_TILE"); hostsim_clean_w32(D, ctx->T, keep, hs_code, braw, ts ? atoll(ts) : 32); }
    else { const char* ts = getenv("TRCHUNK_HOSTSIM_TILE"); hostsim_clean_swar(D, ctx->T, keep, hs_code, braw, ts ? atoll(ts) : 256); }   // 256 = F3_SEGS
  } else
#endif
  LAUNCH(k_clean_count, n, 128, D, ctx->T, keep);
  if ((rc = scan_excl(ctx, st, keep, D.kscan, n))) return rc;
  if ((rc = scan_excl(ctx, st, D.clen, D.cscan, n))) return rc;
  int64_t h2[2] = {0, 0};
  if (n) {
    if ((rc = read_back2(ctx, st, D.kscan + n, D.cscan + n, h2))) return rc;
  }
  nk = h2[0]; bclean = h2[1];
  D.n_kept = nk;
  ctx->counts[TRC_N_KEPT] = nk; ctx->counts[TRC_B_CLEAN] = bclean;
  ALLOC(D.k2r, int32_t, nk); ALLOC(D.coff, int64_t, nk + 1); ALLOC(D.C, uint8_t, bclean); ALLOC(D.sdoc, int32_t, nk);
  ALLOC(D.kdoc_off, int64_t, nd + 1);
  if (n == 0) CK(cudaMemsetAsync(D.coff, 0, 8, st));
  LAUNCH(k_compact, n, 256, D, keep);
  if (n) LAUNCH(k_map_off, nd + 1, 256, D.doc_off, D.kscan, D.kdoc_off, nd + 1);
  else CK(cudaMemsetAsync(D.kdoc_off, 0, 8 * (size_t)(nd + 1), st));
#ifdef TRC_HOSTSIM
  if (hs_par) hostsim_clean_par_write(D, hs_code); else
#endif
  LAUNCH(k_clean_write, nk, 128, D, ctx->T);

  // P2: tokens of " " + cleaned text, special-token scan
  ALLOC(D.ctok, uint32_t, nk); ALLOC(seg_special, int32_t, nk);
  if ((rc = count_units(ctx, st, D.C, D.coff, D.coff + 1, nk, bclean, 1, D.ctok, seg_special, true))) return rc;
  }
  if (spk_long.pending) {                      // (paths without the fused kernel's pre-pass: a round trip of its own)
    unsigned long long h[2];
    if ((rc = read_back(ctx, st, spk_long.ext_cnt, 16, h))) return rc;
    if ((rc = finish_long(ctx, st, spk_long, h[0], h[1]))) return rc;
  }
  D.kspk = nullptr; D.kdur = nullptr;
  if (D.merge && D.has_cap && !(modeB && B->group_off)) { ALLOC(D.kspk, int32_t, nk); ALLOC(D.kdur, double, nk); }
  LAUNCH(k_special_docs, nk, 256, D, seg_special, spk_special);
  LAUNCH(k_doc_err, nd, 128, D, seg_special, spk_special);

  // P3: merge
  ALLOC(D.mnxt, int32_t, nk); ALLOC(D.gflag, int32_t, nk); ALLOC(D.gscan, int64_t, nk + 1);
  CK(cudaMemsetAsync(D.gflag, 0, 4 * (size_t)(nk > 0 ? nk : 1), st));
  const int64_t* d_group_off = nullptr;
  if (modeB && B->group_off) {
    if ((rc = to_device(ctx, st, B->group_off, B->n_groups + 1, B->memory, &d_group_off))) return rc;
    LAUNCH(k_flag_groups, B->n_groups, 256, d_group_off, B->n_groups, D.gflag);
  } else if (!D.merge || !D.has_cap) {
    LAUNCH(k_group_flag_nocap, nk, 256, D);
  } else {
    LAUNCH(k_merge_next, nk, 256, D);
    if ((rc = orbit_marks(ctx, st, D.kdoc_off, D.mnxt, D.gflag, nd, nk))) return rc;
  }
  if ((rc = scan_excl(ctx, st, D.gflag, D.gscan, nk))) return rc;
  // (the number of merged segments is only needed on the host after the piece layout: gstart is sized by its
  // bound n_kept + 1 and one round trip returns both totals)
  ALLOC(D.gstart, int32_t, nk + 1); ALLOC(D.gdoc_off, int64_t, nd + 1);
  if (nk == 0) { CK(cudaMemsetAsync(D.gstart, 0, 4, st)); CK(cudaMemsetAsync(D.gdoc_off, 0, 8 * (size_t)(nd + 1), st)); }
  LAUNCH(k_scatter_starts, nk, 256, D.gflag, D.gscan, D.gstart, nk);
  if (nk) LAUNCH(k_map_off, nd + 1, 256, D.kdoc_off, D.gscan, D.gdoc_off, nd + 1);

  // P4: piece layout of the merged text
  int32_t* plen_sub = nullptr;
  ALLOC(D.plen, int32_t, nk); ALLOC(D.poff, int64_t, nk + 1);
  if (!D.preserve) { ALLOC(plen_sub, int32_t, nk); ALLOC(D.poffs, int64_t, nk + 1); } else D.poffs = D.poff;
  LAUNCH(k_piece_len, nk, 256, D, plen_sub);
  if ((rc = scan_excl(ctx, st, D.plen, D.poff, nk))) return rc;
  if (!D.preserve) { if ((rc = scan_excl(ctx, st, plen_sub, D.poffs, nk))) return rc; }
  int64_t m = 0, bm = 0;
#ifndef TRC_HOSTSIM
  if (nk && slow_pending) {                    // the join with the side stream rides on the same round trip (see below)
    CK(cudaEventRecord(ctx->ev_aux[1], ctx->aux_stream));
    CK(cudaStreamWaitEvent(st, ctx->ev_aux[1], 0));
    int64_t t4[4];
    if ((rc = read_back4(ctx, st, D.gscan + nk, D.poff + nk, j_fz + 2, t4))) return rc;
    m = t4[0]; bm = t4[1]; j_h0 = (unsigned long long)t4[2]; j_h1 = (unsigned long long)t4[3];
    slow_pending = false;
  } else
#endif
  if (nk) { int64_t t2[2]; if ((rc = read_back2(ctx, st, D.gscan + nk, D.poff + nk, t2))) return rc; m = t2[0]; bm = t2[1]; }
  D.m = m; ctx->counts[TRC_N_GROUPS] = m;
  ctx->counts[TRC_B_MERGED] = bm;
  if (B->want_merged_text || modeB) {
    ALLOC(D.M, uint8_t, bm);
    LAUNCH_W(k_emit_merged, nk, D);
    set_arr(ctx, TRC_A_MERGED, D.M, bm);
  }
#ifndef TRC_HOSTSIM
  // join: the slow words' counts (side stream) and the long spans both feed ctok, which the group totals read
  if (slow_pending) {
    unsigned long long h[2];
    if ((rc = read_back(ctx, ctx->aux_stream, j_fz + 2, 16, h))) return rc;    // long spans found by the slow words included
    j_h0 = h[0]; j_h1 = h[1];
    CK(cudaEventRecord(ctx->ev_aux[1], ctx->aux_stream));
    CK(cudaStreamWaitEvent(st, ctx->ev_aux[1], 0));
  }
  if ((int64_t)j_h0 > j_wl_cap) { ctx->err = "long-span work list overflow"; return TRC_EINVAL; }   // cannot happen: the cap is the bound
  if (j_h0 > 0) {
    uint32_t* scratch;
    ALLOC(scratch, uint32_t, 4 * (int64_t)j_h1);
    int lblocks = (int)(j_h0 < 148 * 8 ? j_h0 : 148 * 8);
    prof_begin(ctx, st, "k_bpe_long");
    k_bpe_long<<<lblocks, LONGB_T, 0, st>>>(ctx->T, ctx->d_byte_id, D.C, D.coff, D.coff + 1, 1, D.ctok, j_wl, (int64_t)j_h0, scratch, (int64_t)j_h1);
    prof_end(ctx, st);
    ctx->launches++;
    CK(cudaGetLastError());
    ctx->counts[TRC_N_LONG_SPANS] += (int64_t)j_h0;
  }
#endif
  // P5: group token totals, formatted lengths, formatted timestamps
  ALLOC(D.gts, uint8_t, m * 48); ALLOC(D.gtok, int64_t, m); ALLOC(D.glen, int64_t, m);
#ifdef TRC_HOSTSIM
  if (nk >= 64 * m) LAUNCH_W(k_group_stats, m, D, ctx->T); else LAUNCH(k_group_stats_t, m, 128, D, ctx->T);
#else
  if (nk >= 64 * m) LAUNCH_W(k_group_stats, m, D, ctx->T);
  else if (nk >= 4 * m) LAUNCH(k_group_stats_8, m * 8, 256, D, ctx->T);
  else LAUNCH(k_group_stats_t, m, 128, D, ctx->T);
#endif
  if (modeB && B->group_off) {
    if (m != B->n_groups) { ctx->err = "group_off does not describe the pieces"; return TRC_EINVAL; }
    const double* d_gs; const double* d_ge; const uint8_t* d_full; const int64_t* d_foff;
    int64_t bfull = 0;
    if (B->memory == TRC_MEM_HOST) bfull = B->full_off[m];
    else { int64_t t1[1]; if ((rc = read_back(ctx, st, B->full_off + m, 8, t1))) return rc; bfull = t1[0]; }
    if ((rc = to_device(ctx, st, B->group_start, m, B->memory, &d_gs))) return rc;
    if ((rc = to_device(ctx, st, B->group_end, m, B->memory, &d_ge))) return rc;
    if ((rc = to_device(ctx, st, B->full_text, bfull, B->memory, &d_full))) return rc;
    if ((rc = to_device(ctx, st, B->full_off, m + 1, B->memory, &d_foff))) return rc;
    uint32_t* ftok; int32_t* fsp;
    ALLOC(ftok, uint32_t, m); ALLOC(fsp, int32_t, m);
    if ((rc = count_units(ctx, st, d_full, d_foff, d_foff + 1, m, bfull, 1, ftok, fsp, false))) return rc;
    LAUNCH(k_verify_groups, m, 128, D, ctx->T, d_gs, d_ge, d_full, d_foff, ftok);
  }
  CK(cudaEventRecord(ctx->ev[1], st));
  ctx->ev_pre = true; ctx->ev_chunk = false;

  set_arr(ctx, TRC_A_K2R, D.k2r, 4 * nk); set_arr(ctx, TRC_A_COFF, D.coff, 8 * (nk + 1)); set_arr(ctx, TRC_A_CLEAN, D.C, bclean);
  set_arr(ctx, TRC_A_CTOK, D.ctok, 4 * nk); set_arr(ctx, TRC_A_GSTART, D.gstart, 4 * (m + 1));
  set_arr(ctx, TRC_A_POFF, D.poff, 8 * (nk + 1)); set_arr(ctx, TRC_A_GTS, D.gts, 48 * m); set_arr(ctx, TRC_A_GTOK, D.gtok, 8 * m);
  set_arr(ctx, TRC_A_DOC_ERR, D.doc_err, 4 * nd); set_arr(ctx, TRC_A_GDOC_OFF, D.gdoc_off, 8 * (nd + 1));
  set_arr(ctx, TRC_A_CCHARS, D.cchars, 4 * n);
  {
    int32_t* clen32;
    ALLOC(clen32, int32_t, nk);
    LAUNCH(k_clen32, nk, 256, D.coff, nk, clen32);
    set_arr(ctx, TRC_A_CLEN32, clen32, 4 * nk);
  }
  ctx->have_pre = true;
  return TRC_OK;
}

#include "trc_agg.cuh"
static_assert(sizeof(TrcAggRec) == sizeof(trc_agg) && sizeof(TrcAggMemRec) == sizeof(trc_agg_mem), "aggregate record layouts");

// ---- prompt assembly (llm_executor.py:187-190): lit[0] + twc(c) + lit[1] + twc(c) + ... + lit[n_lit-1] per chunk
__global__ void k_prompt_len(TrcDev D, int64_t lit_total, int32_t n_lit, int64_t* plen) {
  GSTRIDE(c, D.n_chunks) plen[c] = lit_total + (int64_t)(n_lit - 1) * D.twc_len[c];
}
__global__ void k_prompt_write(TrcDev D, const uint8_t* __restrict__ lit, const int64_t* __restrict__ lit_off, int32_t n_lit,
                               const int64_t* __restrict__ poff, uint8_t* out) {
  WSTRIDE(c, D.n_chunks) {
    const int lane = TRC_LANE, nl = TRC_NL;
    uint8_t* dst = out + poff[c];
    const uint8_t* src = D.twc + D.twc_scan[c];
    const int64_t tl = D.twc_len[c];
    for (int32_t k = 0; k < n_lit; ++k) {
      const int64_t a = lit_off[k], L = lit_off[k + 1] - a;
      for (int64_t q = lane; q < L; q += nl) dst[q] = lit[a + q];
      dst += L;
      if (k + 1 < n_lit) { for (int64_t q = lane; q < tl; q += nl) dst[q] = src[q]; dst += tl; }
    }
  }
}
extern "C" int trc_assemble_prompts(trc_ctx* ctx, const uint8_t* lit_bytes, const int64_t* lit_off, int32_t n_lit, void* stream) {
  if (!ctx || !lit_off || n_lit < 1 || (!lit_bytes && lit_off[n_lit] > 0)) return TRC_EINVAL;
  if (!ctx->have_chunk) { ctx->err = "trc_assemble_prompts needs a preceding trc_chunk"; return TRC_EINVAL; }
  CK(cudaSetDevice(ctx->device));
  cudaStream_t st = stream ? (cudaStream_t)stream : ctx->own_stream;
  TrcDev& D = ctx->D;
  int rc;
  const int64_t nc = D.n_chunks, lt = lit_off[n_lit] - lit_off[0];
  const uint8_t* d_lit; const int64_t* d_loff;
  if ((rc = to_device(ctx, st, lit_bytes ? lit_bytes + lit_off[0] : (const uint8_t*)"", lt > 0 ? lt : 1, TRC_MEM_HOST, &d_lit))) return rc;
  std::vector<int64_t> rel((size_t)n_lit + 1);
  for (int32_t k = 0; k <= n_lit; ++k) rel[(size_t)k] = lit_off[k] - lit_off[0];
  int64_t* loff;
  ALLOC(loff, int64_t, n_lit + 1);
  CK(cudaMemcpyAsync(loff, rel.data(), 8 * (size_t)(n_lit + 1), cudaMemcpyHostToDevice, st));
  CK(cudaStreamSynchronize(st));                        // rel is a local vector
  d_loff = loff;
  int64_t* plen; int64_t* poff;
  ALLOC(plen, int64_t, nc); ALLOC(poff, int64_t, nc + 1);
  LAUNCH(k_prompt_len, nc, 256, D, lt, n_lit, plen);
  if ((rc = scan_excl(ctx, st, plen, poff, nc))) return rc;
  int64_t total = 0;
  if (nc) { if ((rc = read_back(ctx, st, poff + nc, 8, &total))) return rc; }
  uint8_t* out;
  ALLOC(out, uint8_t, total);
  LAUNCH_W(k_prompt_write, nc, D, d_lit, d_loff, n_lit, poff, out);
  set_arr(ctx, TRC_A_PROMPTS, out, total);
  set_arr(ctx, TRC_A_PROMPT_OFF, poff, 8 * (nc + 1));
  return TRC_OK;
}

extern "C" int trc_aggregate(trc_ctx* ctx, double interval_seconds, void* stream) {
  if (!ctx) return TRC_EINVAL;
  if (!ctx->have_pre) { ctx->err = "trc_aggregate needs a preceding trc_preprocess"; return TRC_EINVAL; }
  if (!(interval_seconds == interval_seconds) || interval_seconds == 0.0) { ctx->err = "interval_seconds must be a non-zero number"; return TRC_EINVAL; }
  CK(cudaSetDevice(ctx->device));
  cudaStream_t st = stream ? (cudaStream_t)stream : ctx->own_stream;
  return aggregate_stage(ctx, st, interval_seconds);
}

extern "C" int trc_chunk(trc_ctx* ctx, const int32_t* doc_max_tokens, int32_t context_tokens, int32_t add_context, void* stream) {
  if (!ctx || !doc_max_tokens) return TRC_EINVAL;
  if (!ctx->have_pre) { ctx->err = "trc_chunk needs a preceding trc_preprocess"; return TRC_EINVAL; }
  CK(cudaSetDevice(ctx->device));
  cudaStream_t st = stream ? (cudaStream_t)stream : ctx->own_stream;
  TrcDev& D = ctx->D;
  int rc;
  CK(cudaEventRecord(ctx->ev[2], st));
  const int64_t nd = D.n_doc, nk = D.n_kept, m = D.m;
  D.add_context = add_context ? 1 : 0;
  {
    std::vector<int32_t> E((size_t)(nd > 0 ? nd : 1));
    for (int64_t x = 0; x < nd; ++x) E[x] = doc_max_tokens[x] - context_tokens;   // big_chunkeroosky.py:42
    int32_t* dE;
    ALLOC(dE, int32_t, nd);
    // (pageable source: the call returns once the bytes are staged, so E may go out of scope without a synchronisation)
    CK(cudaMemcpyAsync(dE, E.data(), 4 * (size_t)(nd > 0 ? nd : 1), cudaMemcpyHostToDevice, st));
    D.doc_E = dE;
  }
  // C1: over-budget segments -> sub-chunks
  int32_t* nsub_g; int32_t* d_nraw;
  ALLOC(nsub_g, int32_t, m); ALLOC(d_nraw, int32_t, 1); ALLOC(D.sub_off, int64_t, m + 1);
  ALLOC(D.nitems_g, int32_t, m); ALLOC(D.item_off, int64_t, m + 1);
  CK(cudaMemsetAsync(d_nraw, 0, 4, st));
  if (nk >= 64 * m) LAUNCH_W(k_large_count, m, D, ctx->T, nsub_g, d_nraw); else LAUNCH(k_large_count_t, m, 128, D, ctx->T, nsub_g, d_nraw);
  // Optimistic: most batches have no over-budget raw segment, so the sub-chunk and item scans run right away and ONE round
  // trip returns (raw large segments, sub-chunks, items); only when the sentence path is needed are the scans redone.
  int32_t nraw = 0;
  int64_t h2[2] = {0, 0};
  ctx->SS = TrcSentState{};
  if ((rc = scan_excl(ctx, st, nsub_g, D.sub_off, m))) return rc;
  LAUNCH(k_nitems, m, 256, D, nsub_g);
  if ((rc = scan_excl(ctx, st, D.nitems_g, D.item_off, m))) return rc;
  if (m) {
    int64_t h3[3];
    if ((rc = read_back3(ctx, st, d_nraw, D.sub_off + m, D.item_off + m, h3))) return rc;
    nraw = (int32_t)h3[0]; h2[0] = h3[1]; h2[1] = h3[2];
  }
  if (nraw > 0) {
    if ((rc = sentence_stage_count(ctx, st, nsub_g, nraw))) return rc;
    if ((rc = scan_excl(ctx, st, nsub_g, D.sub_off, m))) return rc;
    LAUNCH(k_nitems, m, 256, D, nsub_g);
    if ((rc = scan_excl(ctx, st, D.nitems_g, D.item_off, m))) return rc;
    if ((rc = read_back2(ctx, st, D.sub_off + m, D.item_off + m, h2))) return rc;
  }
  const int64_t nsub = h2[0], ni = h2[1];
  D.n_subs = nsub; D.n_items = ni;
  ctx->counts[TRC_N_SUBS] = nsub; ctx->counts[TRC_N_ITEMS] = ni;
  ALLOC(D.subs, TrcSub, nsub);
  if (nk >= 64 * m) LAUNCH_W(k_large_write, m, D, ctx->T, nsub_g); else LAUNCH(k_large_write_t, m, 128, D, ctx->T, nsub_g);
  if (nraw > 0) { if ((rc = sentence_stage_write(ctx, st, nsub_g))) return rc; }
  // C2: items
  ALLOC(D.it_group, int32_t, ni); ALLOC(D.it_sub, int32_t, ni); ALLOC(D.it_tok, int32_t, ni); ALLOC(D.it_len, int64_t, ni);
  ALLOC(D.it_forced, uint8_t, ni); ALLOC(D.pc_item, int32_t, nk);
  ALLOC(D.W, int64_t, ni + 1); ALLOC(D.L, int64_t, ni + 1); ALLOC(D.idoc_off, int64_t, nd + 1);
#ifdef TRC_HOSTSIM
  LAUNCH(k_build_items, m, 128, D);
#else
  if (nk >= 64 * m) LAUNCH_W(k_build_items_w, m, D); else LAUNCH(k_build_items, m, 128, D);
#endif
  if ((rc = scan_excl(ctx, st, D.it_tok, D.W, ni))) return rc;
  if ((rc = scan_excl(ctx, st, D.it_len, D.L, ni, 2))) return rc;
  if (m) LAUNCH(k_map_off, nd + 1, 256, D.gdoc_off, D.item_off, D.idoc_off, nd + 1);
  else CK(cudaMemsetAsync(D.idoc_off, 0, 8 * (size_t)(nd + 1), st));
  // forced breaks = first item of every over-budget segment
  int32_t* large_flag; int64_t* large_scan; int64_t* forced_pos;
  ALLOC(large_flag, int32_t, m); ALLOC(large_scan, int64_t, m + 1);
  LAUNCH(k_large_flag, m, 256, D, large_flag);
  if ((rc = scan_excl(ctx, st, large_flag, large_scan, m))) return rc;
  ALLOC(forced_pos, int64_t, m);                 // at most one forced break per merged segment; the count stays on the device
  LAUNCH(k_forced_list, m, 256, D, large_flag, large_scan, forced_pos);
  // C3: greedy packing as orbit of pnxt
  ALLOC(D.pnxt, int32_t, ni); ALLOC(D.cflag, int32_t, ni); ALLOC(D.cscanf, int64_t, ni + 1);
  LAUNCH(k_pack_next, ni, 256, D, forced_pos, large_scan + m);
  CK(cudaMemsetAsync(D.cflag, 0, 4 * (size_t)(ni > 0 ? ni : 1), st));
  if ((rc = orbit_marks(ctx, st, D.idoc_off, D.pnxt, D.cflag, nd, ni))) return rc;
  if ((rc = scan_excl(ctx, st, D.cflag, D.cscanf, ni))) return rc;
  int64_t nc = 0;
  if (ni) { if ((rc = read_back(ctx, st, D.cscanf + ni, 8, &nc))) return rc; }
  D.n_chunks = nc; ctx->counts[TRC_N_CHUNKS] = nc;
  ALLOC(D.cstart, int32_t, nc + 1); ALLOC(D.cdoc_off, int64_t, nd + 1);
  if (ni == 0) { CK(cudaMemsetAsync(D.cstart, 0, 4, st)); CK(cudaMemsetAsync(D.cdoc_off, 0, 8 * (size_t)(nd + 1), st)); }
  LAUNCH(k_scatter_starts, ni, 256, D.cflag, D.cscanf, D.cstart, ni);
  if (ni) LAUNCH(k_map_off, nd + 1, 256, D.idoc_off, D.cscanf, D.cdoc_off, nd + 1);
  // C4: chunk records, headers
  ALLOC(D.chunks, TrcChunk, nc); ALLOC(D.spk_list, int32_t, ni); ALLOC(D.twc_len, int64_t, nc); ALLOC(D.twc_scan, int64_t, nc + 1);
  ALLOC(D.it_dst, int64_t, ni);
#ifdef TRC_HOSTSIM
  LAUNCH(k_chunk_record, nc, 64, D);
#else
  LAUNCH_W(k_chunk_record_w, nc, D);
#endif
  if ((rc = scan_excl(ctx, st, D.twc_len, D.twc_scan, nc))) return rc;
  int64_t h3[2] = {0, 0};
  if (nc || ni) { if ((rc = read_back2(ctx, st, nc ? D.twc_scan + nc : nullptr, ni ? D.W + ni : nullptr, h3))) return rc; }
  const int64_t btwc = h3[0];
  ctx->counts[TRC_B_TWC] = btwc; ctx->counts[TRC_TOKENS] = h3[1];
  ALLOC(D.twc, uint8_t, btwc);
  // C5: item text (layout per item, then a flat copy per piece)
  ALLOC(D.pdst, int64_t, nk);
  CK(cudaMemsetAsync(D.pdst, 0xff, 8 * (size_t)(nk > 0 ? nk : 1), st));
#ifdef TRC_HOSTSIM
  ALLOC(D.pflag, uint8_t, nk);
  LAUNCH(k_chunk_header, nc, 64, D);
  LAUNCH(k_item_layout, ni, 128, D);
  LAUNCH(k_emit_pieces_flat, nk * 8, 256, D);
#else
  {
    TrcItemLay* lay;
    ALLOC(lay, TrcItemLay, ni);
    const int hb = (int)((nc * 32 + 255) / 256), ib = (int)((ni + 255) / 256);
    if (hb + ib > 0) {
      prof_begin(ctx, st, "k_headers_items");
      k_headers_items<<<hb + ib, 256, 0, st>>>(D, lay, hb);
      prof_end(ctx, st);
      ctx->launches++;
      CK(cudaGetLastError());
    }
    LAUNCH(k_emit_pieces_flat, nk * 8, 256, D, lay);
  }
#endif
  if (nraw > 0) { if ((rc = sentence_stage_emit(ctx, st))) return rc; }
  CK(cudaEventRecord(ctx->ev[3], st));
  ctx->ev_chunk = true;
  // error documents
  set_arr(ctx, TRC_A_ITEM_GROUP, D.it_group, 4 * ni); set_arr(ctx, TRC_A_ITEM_SUB, D.it_sub, 4 * ni);
  set_arr(ctx, TRC_A_ITEM_TOK, D.it_tok, 4 * ni); set_arr(ctx, TRC_A_ITEM_LEN, D.it_len, 8 * ni);
  set_arr(ctx, TRC_A_ITEM_DST, D.it_dst, 8 * ni); set_arr(ctx, TRC_A_SUBS, D.subs, (int64_t)sizeof(TrcSub) * nsub);
  set_arr(ctx, TRC_A_CHUNKS, D.chunks, (int64_t)sizeof(TrcChunk) * nc); set_arr(ctx, TRC_A_SPK_LIST, D.spk_list, 4 * ni);
  set_arr(ctx, TRC_A_TWC, D.twc, btwc); set_arr(ctx, TRC_A_CDOC_OFF, D.cdoc_off, 8 * (nd + 1));
  set_arr(ctx, TRC_A_SENT, ctx->SS.recs, (int64_t)sizeof(TrcSent) * ctx->SS.n_recs);
  set_arr(ctx, TRC_A_PDST, D.pdst, 8 * nk);
  ctx->counts[TRC_N_SENT] = ctx->SS.n_recs;
  ctx->have_chunk = true;
  return TRC_OK;
}

extern "C" int trc_ingest_json(const uint8_t* json, const int64_t* doc_off, int64_t n_doc, int32_t n_threads, trc_ingest** out,
                               char* err, int64_t err_cap) {
  if (!json || !doc_off || n_doc < 0 || !out) return TRC_EINVAL;
  *out = nullptr;
  trc_ingest* R = new trc_ingest();
  std::string e;
  try { e = trc_ingest_run(json, doc_off, n_doc, n_threads, R); } catch (const std::exception& x) { e = std::string("ingest: ") + x.what(); }
  if (!e.empty()) {
    delete R;
    if (err && err_cap > 0) { size_t k = e.size() < (size_t)err_cap - 1 ? e.size() : (size_t)err_cap - 1; memcpy(err, e.data(), k); err[k] = 0; }
    return TRC_EINVAL;
  }
  *out = R;
  return TRC_OK;
}
extern "C" int trc_ingest_view(const trc_ingest* R, trc_batch* B, const uint8_t** start_kind, const uint8_t** end_kind) {
  if (!R || !B) return TRC_EINVAL;
  B->text = R->text.data(); B->seg_off = R->seg_off.data(); B->start = R->start.data(); B->end = R->end.data();
  B->speaker = R->speaker.data(); B->n_seg = (int64_t)R->start.size(); B->doc_off = R->doc_off.data();
  B->n_doc = (int64_t)R->doc_off.size() - 1; B->spk_bytes = R->spk_bytes.data(); B->spk_off = R->spk_off.data();
  B->n_spk = (int32_t)R->n_spk; B->memory = TRC_MEM_HOST;
  if (start_kind) *start_kind = R->start_kind.data();
  if (end_kind) *end_kind = R->end_kind.data();
  return TRC_OK;
}
extern "C" void trc_ingest_free(trc_ingest* R) { delete R; }

extern "C" int trc_get_counts(trc_ctx* ctx, int64_t* counts) {
  if (!ctx || !counts) return TRC_EINVAL;
  for (int k = 0; k < TRC_N_COUNTERS; ++k) counts[k] = ctx->counts[k];
  return TRC_OK;
}

extern "C" int trc_get_array(trc_ctx* ctx, int which, void* dst, int64_t dst_bytes, int32_t memory, int64_t* nbytes, void* stream) {
  if (!ctx || which < 0 || which >= TRC_A__COUNT) return TRC_EINVAL;
  if (!ctx->have_pre) { ctx->err = "no batch resident"; return TRC_EINVAL; }
  const trc_ctx::Arr& a = ctx->arr[which];
  if (nbytes) *nbytes = a.bytes;
  if (!dst) return TRC_OK;
  if (!a.p && a.bytes > 0) { ctx->err = "array not produced by the stages run so far"; return TRC_EINVAL; }
  if (dst_bytes < a.bytes) { ctx->err = "destination too small"; return TRC_EINVAL; }
  if (a.bytes == 0) return TRC_OK;
  CK(cudaSetDevice(ctx->device));
  cudaStream_t st = stream ? (cudaStream_t)stream : ctx->own_stream;
  CK(cudaMemcpyAsync(dst, a.p, (size_t)a.bytes, memory == TRC_MEM_HOST ? cudaMemcpyDeviceToHost : cudaMemcpyDeviceToDevice, st));
  if (memory == TRC_MEM_HOST) CK(cudaStreamSynchronize(st));
  return TRC_OK;
}

// Several result arrays in ONE call: array which[k] lands at dst + offsets[k] (offsets are chosen here, 256-byte
// aligned, and returned); one cudaMemcpyAsync per array and a single stream synchronisation at the end when the
// destination is host memory.  With dst == NULL only the offsets and the total size are computed.
extern "C" int trc_get_arrays(trc_ctx* ctx, const int32_t* which, int32_t n_which, void* dst, int64_t dst_bytes, int32_t memory,
                              int64_t* offsets, int64_t* nbytes, int64_t* total_bytes, void* stream) {
  if (!ctx || !which || n_which < 0 || !offsets || !nbytes) return TRC_EINVAL;
  if (!ctx->have_pre) { ctx->err = "no batch resident"; return TRC_EINVAL; }
  int64_t off = 0;
  for (int32_t k = 0; k < n_which; ++k) {
    if (which[k] < 0 || which[k] >= TRC_A__COUNT) { ctx->err = "unknown array"; return TRC_EINVAL; }
    const trc_ctx::Arr& a = ctx->arr[which[k]];
    if (!a.p && a.bytes > 0) { ctx->err = "array not produced by the stages run so far"; return TRC_EINVAL; }
    offsets[k] = off; nbytes[k] = a.bytes;
    off += (a.bytes + 255) & ~(int64_t)255;
  }
  if (total_bytes) *total_bytes = off;
  if (!dst) return TRC_OK;
  if (dst_bytes < off) { ctx->err = "destination too small"; return TRC_EINVAL; }
  CK(cudaSetDevice(ctx->device));
  cudaStream_t st = stream ? (cudaStream_t)stream : ctx->own_stream;
  const cudaMemcpyKind kind = memory == TRC_MEM_HOST ? cudaMemcpyDeviceToHost : cudaMemcpyDeviceToDevice;
  for (int32_t k = 0; k < n_which; ++k)
    if (nbytes[k] > 0) CK(cudaMemcpyAsync((char*)dst + offsets[k], ctx->arr[which[k]].p, (size_t)nbytes[k], kind, st));
  if (memory == TRC_MEM_HOST) CK(cudaStreamSynchronize(st));
  return TRC_OK;
}

// format_timestamp / _format_time for n values with one small kernel (preprocessor.py:91-107 ==
// big_chunkeroosky.py:234-242).  out: n slots of 24 bytes, [length][characters].  Does not touch the resident batch.
__global__ void k_format_ts(const double* __restrict__ sec, int64_t n, uint8_t* __restrict__ out) {
  GSTRIDE(i, n) { uint8_t* o = out + 24 * i; o[0] = (uint8_t)trc_fmt_ts(sec[i], o + 1); }
}
extern "C" int trc_format_timestamps(trc_ctx* ctx, const double* seconds, int64_t n, uint8_t* out24, int32_t memory, void* stream) {
  if (!ctx || n < 0 || (n > 0 && (!seconds || !out24))) return TRC_EINVAL;
  if (n == 0) return TRC_OK;
  CK(cudaSetDevice(ctx->device));
  cudaStream_t st = stream ? (cudaStream_t)stream : ctx->own_stream;
  const double* d_sec = seconds; uint8_t* d_out = out24;
  void* tmp = nullptr;
  if (memory == TRC_MEM_HOST) {                       // a short-lived allocation of its own: the arena keeps the resident batch
    CK(cudaMalloc(&tmp, (size_t)n * 32));
    d_sec = (const double*)tmp; d_out = (uint8_t*)tmp + 8 * (size_t)n;
    if (cudaMemcpyAsync(tmp, seconds, 8 * (size_t)n, cudaMemcpyHostToDevice, st) != cudaSuccess) { cudaFree(tmp); ctx->err = "copy of the timestamps failed"; return TRC_ECUDA; }
  }
#ifdef TRC_HOSTSIM
  k_format_ts(d_sec, n, d_out);
#else
  k_format_ts<<<grid_for(n, 128), 128, 0, st>>>(d_sec, n, d_out);
#endif
  ctx->launches++;
  int rc = TRC_OK;
  if (memory == TRC_MEM_HOST) {
    if (cudaMemcpyAsync(out24, d_out, 24 * (size_t)n, cudaMemcpyDeviceToHost, st) != cudaSuccess || cudaStreamSynchronize(st) != cudaSuccess) {
      ctx->err = "copy of the formatted timestamps failed"; rc = TRC_ECUDA;
    }
    cudaFree(tmp);
  } else if (cudaGetLastError() != cudaSuccess) { ctx->err = "k_format_ts launch failed"; rc = TRC_ECUDA; }
  return rc;
}

extern "C" int trc_last_stage_ms(trc_ctx* ctx, float* pre, float* chk) {
  if (!ctx) return TRC_EINVAL;
  CK(cudaSetDevice(ctx->device));
  if (pre) { *pre = 0; if (ctx->ev_pre) { CK(cudaEventSynchronize(ctx->ev[1])); CK(cudaEventElapsedTime(pre, ctx->ev[0], ctx->ev[1])); } }
  if (chk) { *chk = 0; if (ctx->ev_chunk) { CK(cudaEventSynchronize(ctx->ev[3])); CK(cudaEventElapsedTime(chk, ctx->ev[2], ctx->ev[3])); } }
  return TRC_OK;
}

extern "C" int trc_set_profiling(trc_ctx* ctx, int32_t on) {
  if (!ctx) return TRC_EINVAL;
  prof_collect(ctx);
  ctx->prof = on != 0;
  ctx->prof_acc.clear();
  return TRC_OK;
}

extern "C" int trc_get_profile(trc_ctx* ctx, char* buf, int64_t buf_bytes) {
  if (!ctx || !buf || buf_bytes < 1) return TRC_EINVAL;
  CK(cudaSetDevice(ctx->device));
  prof_collect(ctx);
  std::string s;
  for (auto& a : ctx->prof_acc) { char line[256]; snprintf(line, sizeof line, "%s %.6f %lld\n", a.name.c_str(), a.ms, (long long)a.launches); s += line; }
  if ((int64_t)s.size() + 1 > buf_bytes) { ctx->err = "profile buffer too small"; return TRC_EINVAL; }
  memcpy(buf, s.c_str(), s.size() + 1);
  return TRC_OK;
}

extern "C" int trc_kernel_launches(trc_ctx* ctx, int64_t* launches) {
  if (!ctx || !launches) return TRC_EINVAL;
  *launches = ctx->launches;
  return TRC_OK;
}


#ifndef TRC_HOSTSIM
// =====================================================================================================
// Allocation-free span counter.
//
// Some embedders of this library are not allowed to allocate device memory once they are running: a
// serving process that has handed all of HBM to its own pool at start-up, a plugin host that forbids
// driver allocations after initialisation, or a latency-critical loop that cannot afford the implicit
// device-wide synchronisation a cudaMalloc/cudaFree brings.  The normal long-span merge (k_bpe_long)
// takes its four work arrays from the context arena, which may grow.  This variant does the same merge
// for ONE span with its work arrays in a fixed array that lives in the library image, so the call
// itself allocates nothing; the caller supplies the input bytes and the output word.
//
// The array is shared by every call on the device and keeps whatever the previous call left in it.
// The kernel writes each of the entries it will read before reading them, so no clearing pass is needed,
// but two calls that overlap on the GPU (different streams) would corrupt each other: serialise them.
// =====================================================================================================
#include <mutex>
#define TRC_STATIC_SPAN_MAX 65536                     /* bytes per span; 4 x 4 B per byte = 1 MiB per GPU */
static __device__ __align__(256) uint32_t g_span_scratch[4 * TRC_STATIC_SPAN_MAX];
static std::mutex g_span_mu;
static uint32_t* g_span_scratch_addr[64] = {nullptr};  // device address of the array, per device ordinal

__global__ void __launch_bounds__(LONG_T) k_bpe_span_static(TrcTables T, const uint32_t* __restrict__ byte_id,
                                                            const uint8_t* __restrict__ span, int64_t len,
                                                            uint32_t* __restrict__ out_count, uint32_t* scratch) {
  __shared__ unsigned long long s_red[LONG_T / 32];
  __shared__ unsigned long long s_best;
  __shared__ int s_whole;
  TrcStr S; S.p = span; S.vs = 0; S.n = len;
  if (threadIdx.x == 0) s_whole = (len == 1 || trc_is_token(T, trc_fast_default(T), S, 0, (int)len)) ? 1 : 0;
  __syncthreads();
  if (s_whole) { if (threadIdx.x == 0) *out_count = 1u; return; }
  uint32_t* id = scratch;
  uint32_t* rk = scratch + TRC_STATIC_SPAN_MAX;
  int32_t* nx = (int32_t*)(scratch + 2 * TRC_STATIC_SPAN_MAX);
  int32_t* pv = (int32_t*)(scratch + 3 * TRC_STATIC_SPAN_MAX);
  for (int64_t i = threadIdx.x; i < len; i += LONG_T) { id[i] = byte_id[span[i]]; nx[i] = (int32_t)(i + 1); pv[i] = (int32_t)(i - 1); }
  __syncthreads();
  for (int64_t i = threadIdx.x; i < len; i += LONG_T) rk[i] = i + 1 < len ? trc_pair_rank(T, id[i], id[i + 1]) : TRC_NO_RANK;
  __syncthreads();
  int64_t m = len;
  for (;;) {
    unsigned long long best = ~0ULL;
    for (int64_t i = threadIdx.x; i < len; i += LONG_T) {
      unsigned long long k = ((unsigned long long)rk[i] << 32) | (unsigned long long)i;
      best = k < best ? k : best;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { unsigned long long y = __shfl_xor_sync(0xffffffffu, best, o); best = y < best ? y : best; }
    if ((threadIdx.x & 31) == 0) s_red[threadIdx.x >> 5] = best;
    __syncthreads();
    if (threadIdx.x == 0) {
      unsigned long long b = s_red[0];
      for (int k = 1; k < LONG_T / 32; ++k) b = s_red[k] < b ? s_red[k] : b;
      s_best = b;
      uint32_t r = (uint32_t)(b >> 32);
      if (r != TRC_NO_RANK) {
        int32_t i = (int32_t)(b & 0xffffffffu);
        int32_t j = nx[i], nj = nx[j];
        id[i] = r; nx[i] = nj; rk[j] = TRC_NO_RANK;
        if (nj < len) { pv[nj] = i; rk[i] = trc_pair_rank(T, r, id[nj]); } else rk[i] = TRC_NO_RANK;
        int32_t p = pv[i];
        if (p >= 0) rk[p] = trc_pair_rank(T, id[p], r);
      }
    }
    __syncthreads();
    if ((uint32_t)(s_best >> 32) == TRC_NO_RANK) break;
    --m;
    __syncthreads();
  }
  if (threadIdx.x == 0) *out_count = (uint32_t)m;
}

extern "C" int trc_bpe_span_count_static(trc_ctx* ctx, const uint8_t* d_span, int64_t n_bytes, uint32_t* d_count, void* stream) {
  if (!ctx || !d_span || !d_count) return TRC_EINVAL;
  if (!ctx->have_vocab) { ctx->err = "vocabulary not loaded"; return TRC_ENOVOCAB; }
  if (n_bytes < 1) { ctx->err = "trc_bpe_span_count_static: the span must have at least one byte"; return TRC_EINVAL; }
  if (n_bytes > TRC_STATIC_SPAN_MAX) {
    ctx->err = "trc_bpe_span_count_static: span of " + std::to_string(n_bytes) + " bytes exceeds the built-in work area (limit " +
               std::to_string(TRC_STATIC_SPAN_MAX) + " bytes); use trc_count_tokens for longer spans";
    return TRC_EINVAL;
  }
  if (ctx->device < 0 || ctx->device >= 64) { ctx->err = "device ordinal out of range"; return TRC_EINVAL; }
  CK(cudaSetDevice(ctx->device));
  uint32_t* scratch;
  {
    // first call on this device: resolve the device address of the array (the host-side symbol is only
    // a handle).  This is a one-off; make that first call outside any stream capture.
    std::lock_guard<std::mutex> lk(g_span_mu);
    if (!g_span_scratch_addr[ctx->device]) {
      void* p = nullptr;
      CK(cudaGetSymbolAddress(&p, g_span_scratch));
      g_span_scratch_addr[ctx->device] = (uint32_t*)p;
    }
    scratch = g_span_scratch_addr[ctx->device];
  }
  cudaStream_t st = stream ? (cudaStream_t)stream : ctx->own_stream;
  k_bpe_span_static<<<1, LONG_T, 0, st>>>(ctx->T, ctx->d_byte_id, d_span, n_bytes, d_count, scratch);
  ctx->launches++;
  CK(cudaGetLastError());
  return TRC_OK;
}
#endif
