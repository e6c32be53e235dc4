// Shared definitions for the trchunk CUDA library (sm_100a).
//
// Every per-unit routine in trc_*.cuh is written as a TRC_HD function over raw pointers so that the
// SAME source is compiled (a) by nvcc into the kernels of libtrchunk.so - the product - and (b) by
// g++ into tests/hostsim (a test-only harness that lets the CPU test-suite check the bit-exact logic
// against the oracle without a GPU).  The product never runs these on the host.
#pragma once
#include <stdint.h>
#include <stddef.h>

#if defined(__CUDACC__)
#define TRC_HD __host__ __device__ __forceinline__
#define TRC_HDN __host__ __device__ __noinline__
#else
#define TRC_HD inline
#define TRC_HDN inline
#endif

// unicode_props.json bits (tools/gen_unicode_tables.py)
#define TRC_P_SPACE 1u   // str.isspace()            preprocessor.py:38,80
#define TRC_P_WORD 2u    // re \w                     preprocessor.py:84
#define TRC_P_UPPER 16u  // str.isupper()             Punkt first_upper
#define TRC_P_LOWER 32u  // str.islower()             Punkt first_lower
#define TRC_P_DIGIT 64u  // re \d                     Punkt ##number##, is_initial
#define TRC_RS(p) (((p) >> 2) & 3u)  // Rust-regex class: 0 other, 1 \p{L}, 2 \p{N}, 3 \s
#define TRC_RS_O 0u
#define TRC_RS_L 1u
#define TRC_RS_N 2u
#define TRC_RS_S 3u

#define TRC_MAX_SPECIALS 8
#define TRC_SHORT_SPAN 48
#define TRC_HOT8 2048     // slots of the direct-mapped hot set of short tokens (host image == shared-memory copy)  // spans up to this many bytes are merged in thread-local memory

struct TrcTables {
  const uint8_t* props;  // 0x110000 entries
  // vocabulary (rank == token id, as in tiktoken)
  const uint8_t* tok_bytes;
  const uint32_t* tok_off;  // n_tok+1
  uint32_t n_tok;
  uint32_t tok_max_len;     // longest vocabulary entry in bytes: longer spans are not tokens (no need to hash 64 KB to find out)
  const uint64_t* tokhash;  // (fingerprint32 << 32 | id), empty = ~0
  uint32_t tokhash_mask;
  const uint64_t* pairhash;  // (idL << 42 | idR << 21 | rank), empty = ~0
  uint32_t pairhash_mask;
  // special-token literals (tiktoken raises ValueError when the text contains one)
  const uint8_t* sp_bytes;
  const uint32_t* sp_off;  // n_sp+1
  uint32_t n_sp;
  // token counts of digit groups: index = value + (len==1 ? 0 : len==2 ? 10 : 110)
  const uint8_t* dcount;   // 1110 entries
  uint32_t cnt_space_lbracket;  // count(" [")
  uint32_t dcount_ready;        // 0 while the digit-group table itself is being computed
  // fast paths (exact): Rust class of the 128 ASCII bytes; sets of ALL vocabulary entries of <= 8 and
  // 9..16 bytes that contain no 0x00 byte, keyed by their zero-padded little-endian bytes
  const uint8_t* acls;          // 128 entries
  const uint64_t* tok8;  uint32_t tok8_mask;    // empty slot = 0 (no key is 0: keys have no NUL bytes, len >= 1)
  const uint64_t* tok16; uint32_t tok16_mask;   // pairs (lo, hi); empty = (0,0)
  const uint64_t* hot8_src; uint32_t hot8_size; // direct-mapped image of the lowest-rank <=8-byte tokens (kernels copy it to smem)
  uint64_t sp_first[4];         // bitmap of first bytes of the special-token literals
  uint32_t sp_single;           // the one ASCII byte every special literal starts with, or 0x100
};

TRC_HD uint64_t trc_mix64(uint64_t x);

// Per-kernel view of the small tables a kernel may have staged in shared memory.
struct TrcFast {
  const uint8_t* acls;     // ASCII class table (global or shared)
  const uint64_t* hot8;    // direct-mapped hot set in shared memory, or nullptr
  uint32_t hot8_mask;
};
TRC_HD TrcFast trc_fast_default(const TrcTables& T) { TrcFast F; F.acls = T.acls; F.hot8 = nullptr; F.hot8_mask = 0; return F; }
TRC_HD uint32_t trc_cls(const TrcTables& T, const TrcFast& F, uint32_t cp) { return cp < 128u ? (uint32_t)F.acls[cp] : TRC_RS(T.props[cp]); }
TRC_HD uint32_t trc_key8_slot(uint64_t key, uint32_t mask) {   // cheap 32-bit mix: keys are short byte strings
  uint32_t x = (uint32_t)key ^ ((uint32_t)(key >> 32) * 0x9E3779B1u);
  x *= 0x85EBCA6Bu; x ^= x >> 15; x *= 0xC2B2AE35u; x ^= x >> 13;
  return x & mask;
}

// Sum over the lanes of a warp (device); the host test build runs one "lane", so the value is already the sum.
TRC_HD int64_t trc_lanes_sum(int64_t v) {
#if defined(__CUDA_ARCH__)
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
#endif
  return v;
}
// the same over aligned groups of nl lanes (nl a power of two); `mask` names the lanes that take part
TRC_HD int64_t trc_lanes_sum_w(int64_t v, int nl, uint32_t mask) {
#if defined(__CUDA_ARCH__)
  for (int o = nl >> 1; o > 0; o >>= 1) v += __shfl_xor_sync(mask, v, o);
#else
  (void)nl; (void)mask;
#endif
  return v;
}
TRC_HD uint32_t trc_popc32(uint32_t x) {
#if defined(__CUDA_ARCH__)
  return (uint32_t)__popc(x);
#else
  return (uint32_t)__builtin_popcount(x);
#endif
}
TRC_HD uint32_t trc_ctz32(uint32_t x) {
#if defined(__CUDA_ARCH__)
  return (uint32_t)(__ffs((int)x) - 1);
#else
  return (uint32_t)__builtin_ctz(x);
#endif
}
// bit 7 of each byte of (a, b) -> 16-bit mask (a = bytes 0..7)
TRC_HD uint32_t sw_pack16(uint64_t a, uint64_t b) {
  return (uint32_t)(((a >> 7) * 0x0102040810204080ULL) >> 56) | ((uint32_t)(((b >> 7) * 0x0102040810204080ULL) >> 56) << 8);
}

// ---- UTF-8 (input is produced by Python's str.encode('utf-8'): always well formed) -----------------
TRC_HD int trc_u8len(uint8_t b) { return b < 0x80 ? 1 : (b < 0xE0 ? 2 : (b < 0xF0 ? 3 : 4)); }

TRC_HD uint32_t trc_decode(const uint8_t* p, int64_t i, int64_t n, int* len) {
  uint8_t b = p[i];
  if (b < 0x80) { *len = 1; return b; }
  if (b < 0xE0 && i + 1 < n) { *len = 2; return ((b & 0x1Fu) << 6) | (p[i + 1] & 0x3Fu); }
  if (b < 0xF0 && i + 2 < n) { *len = 3; return ((b & 0x0Fu) << 12) | ((p[i + 1] & 0x3Fu) << 6) | (p[i + 2] & 0x3Fu); }
  if (i + 3 < n) {
    *len = 4;
    uint32_t c = ((b & 0x07u) << 18) | ((p[i + 1] & 0x3Fu) << 12) | ((p[i + 2] & 0x3Fu) << 6) | (p[i + 3] & 0x3Fu);
    return c < 0x110000u ? c : 0xFFFDu;
  }
  *len = 1;
  return 0xFFFDu;
}

TRC_HD uint32_t trc_prop(const TrcTables& T, uint32_t cp) { return T.props[cp]; }

// ---- hashing ------------------------------------------------------------------------------------------
TRC_HD uint64_t trc_mix64(uint64_t x) {
  x ^= x >> 33; x *= 0xff51afd7ed558ccdULL; x ^= x >> 33; x *= 0xc4ceb9fe1a85ec53ULL; x ^= x >> 33;
  return x;
}

// FNV-1a style over bytes, finished with mix64 (host table builder uses the same function)
template <class ByteAt>
TRC_HD uint64_t trc_hash_bytes(ByteAt at, int n) {
  uint64_t h = 0xcbf29ce484222325ULL ^ (uint64_t)n;
  for (int i = 0; i < n; ++i) { h ^= (uint64_t)at(i); h *= 0x100000001b3ULL; }
  return trc_mix64(h);
}

#define TRC_PAIR_KEY(l, r) (((uint64_t)(l) << 21) | (uint64_t)(r))
#define TRC_NO_RANK 0xFFFFFFFFu

TRC_HD uint32_t trc_pair_rank(const TrcTables& T, uint32_t l, uint32_t r) {
  uint64_t key = TRC_PAIR_KEY(l, r);
  uint32_t h = (uint32_t)(trc_mix64(key)) & T.pairhash_mask;
  for (;;) {
    uint64_t e = T.pairhash[h];
    if (e == ~0ULL) return TRC_NO_RANK;
    if ((e >> 21) == key) return (uint32_t)(e & 0x1FFFFFu);
    h = (h + 1) & T.pairhash_mask;
  }
}
