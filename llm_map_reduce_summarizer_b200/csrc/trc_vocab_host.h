// Host-side construction of the lookup tables the kernels probe (set-up work, once per vocabulary).
// Shared by libtrchunk.so (which uploads the result) and tests/hostsim.
#pragma once
#include <stdint.h>
#include <string>
#include <string.h>
#include <unordered_map>
#include <vector>
#include "trc_common.cuh"

struct TrcHostVocab {
  std::vector<uint8_t> tok_bytes;
  std::vector<uint32_t> tok_off;   // n_tok + 1, indexed by token id (== rank)
  std::vector<uint64_t> tokhash;
  std::vector<uint64_t> pairhash;
  std::vector<uint32_t> byte_id;   // 256
  std::vector<uint8_t> sp_bytes;
  std::vector<uint32_t> sp_off;
  std::vector<uint64_t> tok8, tok16, hot8;
  uint64_t sp_first[4] = {0, 0, 0, 0};
  uint32_t n_tok = 0, n_sp = 0;
  uint64_t n_pairs = 0;
};

static inline uint32_t trc_pow2_at_least(uint64_t x) {
  uint32_t p = 1024;
  while ((uint64_t)p < x) p <<= 1;
  return p;
}

// Returns "" on success, else an error message.
static inline std::string trc_build_host_vocab(const uint8_t* token_bytes, const int64_t* token_off, const int32_t* ranks,
                                               int32_t n, const uint8_t* special_bytes, const int64_t* special_off,
                                               int32_t n_special, TrcHostVocab& V) {
  if (n < 256) return "vocabulary must contain at least the 256 single bytes";
  int64_t max_rank = -1;
  for (int32_t i = 0; i < n; ++i) { if (ranks[i] < 0) return "negative rank"; if (ranks[i] > max_rank) max_rank = ranks[i]; }
  if (max_rank >= (1 << 21)) return "rank does not fit in 21 bits";
  uint32_t ntok = (uint32_t)(max_rank + 1);
  V.n_tok = ntok;
  std::vector<int32_t> src(ntok, -1);
  for (int32_t i = 0; i < n; ++i) {
    if (src[ranks[i]] >= 0) return "duplicate rank";
    src[ranks[i]] = i;
  }
  V.tok_off.assign(ntok + 1, 0);
  for (uint32_t r = 0; r < ntok; ++r) {
    int64_t len = src[r] >= 0 ? token_off[src[r] + 1] - token_off[src[r]] : 0;
    V.tok_off[r + 1] = V.tok_off[r] + (uint32_t)len;
  }
  V.tok_bytes.resize(V.tok_off[ntok] + 1);
  std::unordered_map<std::string, uint32_t> ids;
  ids.reserve((size_t)n * 2);
  V.byte_id.assign(256, 0xFFFFFFFFu);
  for (uint32_t r = 0; r < ntok; ++r) {
    if (src[r] < 0) continue;
    const uint8_t* p = token_bytes + token_off[src[r]];
    uint32_t len = V.tok_off[r + 1] - V.tok_off[r];
    if (len == 0) return "empty token";
    memcpy(&V.tok_bytes[V.tok_off[r]], p, len);
    if (!ids.emplace(std::string((const char*)p, len), r).second) return "duplicate token bytes";
    if (len == 1) V.byte_id[p[0]] = r;
  }
  for (int b = 0; b < 256; ++b) if (V.byte_id[b] == 0xFFFFFFFFu) return "vocabulary lacks a single-byte token";
  // token set keyed by bytes
  uint32_t tsz = trc_pow2_at_least((uint64_t)n * 2);
  V.tokhash.assign(tsz, ~0ULL);
  for (uint32_t r = 0; r < ntok; ++r) {
    if (src[r] < 0) continue;
    const uint8_t* p = &V.tok_bytes[V.tok_off[r]];
    int len = (int)(V.tok_off[r + 1] - V.tok_off[r]);
    uint64_t h = trc_hash_bytes([&](int i) { return p[i]; }, len);
    uint32_t slot = (uint32_t)h & (tsz - 1);
    while (V.tokhash[slot] != ~0ULL) slot = (slot + 1) & (tsz - 1);
    V.tokhash[slot] = ((h >> 32) << 32) | r;
  }
  // pair table: every split of every token into two tokens
  std::vector<uint64_t> pairs;
  pairs.reserve((size_t)n * 3);
  for (uint32_t r = 0; r < ntok; ++r) {
    if (src[r] < 0) continue;
    const char* p = (const char*)&V.tok_bytes[V.tok_off[r]];
    uint32_t len = V.tok_off[r + 1] - V.tok_off[r];
    for (uint32_t s = 1; s < len; ++s) {
      auto l = ids.find(std::string(p, s));
      if (l == ids.end()) continue;
      auto rr = ids.find(std::string(p + s, len - s));
      if (rr == ids.end()) continue;
      pairs.push_back((TRC_PAIR_KEY(l->second, rr->second) << 21) | r);
    }
  }
  V.n_pairs = pairs.size();
  uint32_t psz = trc_pow2_at_least((uint64_t)pairs.size() * 2);
  V.pairhash.assign(psz, ~0ULL);
  for (uint64_t e : pairs) {
    uint64_t key = e >> 21;
    uint32_t slot = (uint32_t)trc_mix64(key) & (psz - 1);
    while (V.pairhash[slot] != ~0ULL) slot = (slot + 1) & (psz - 1);
    V.pairhash[slot] = e;
  }
  // exact-key sets for tokens of <= 8 and 9..16 bytes without NUL bytes; hot8 = direct-mapped image of the
  // lowest-rank short tokens (rank order ~ merge frequency), kept by whoever claims a slot first
  {
    std::vector<uint64_t> k8; std::vector<std::pair<uint64_t, uint64_t>> k16;
    for (uint32_t r = 0; r < ntok; ++r) {
      if (src[r] < 0) continue;
      const uint8_t* p = &V.tok_bytes[V.tok_off[r]];
      uint32_t len = V.tok_off[r + 1] - V.tok_off[r];
      if (len > 16) continue;
      uint64_t lo = 0, hi = 0; bool nul = false;
      for (uint32_t i = 0; i < len; ++i) { uint64_t b = p[i]; nul |= b == 0; if (i < 8) lo |= b << (8 * i); else hi |= b << (8 * (i - 8)); }
      if (nul) continue;
      if (len <= 8) k8.push_back(lo); else k16.push_back({lo, hi});
    }
    uint32_t s8 = trc_pow2_at_least(k8.size() * 2), s16 = trc_pow2_at_least(k16.size() * 2);
    V.tok8.assign(s8, 0); V.tok16.assign(2 * (size_t)s16, 0);
    const uint32_t HOT = TRC_HOT8;
    V.hot8.assign(HOT, 0);
    for (uint64_t k : k8) {          // k8 is in rank order
      uint32_t slot = trc_key8_slot(k, s8 - 1);
      while (V.tok8[slot] != 0) slot = (slot + 1) & (s8 - 1);
      V.tok8[slot] = k;
      // hot set: word-shaped keys only (space or letter first), lowest rank first, two candidate slots
      const uint8_t f0 = (uint8_t)k;
      if (f0 == ' ' || ((f0 | 0x20) >= 'a' && (f0 | 0x20) <= 'z')) {
        uint32_t hs = trc_key8_slot(k, HOT - 1);
        if (V.hot8[hs] == 0) V.hot8[hs] = k; else if (V.hot8[hs ^ 1] == 0) V.hot8[hs ^ 1] = k;
      }
    }
    for (auto& k : k16) {
      uint32_t slot = trc_key8_slot(k.first ^ (k.second * 0x9E3779B97F4A7C15ULL), s16 - 1);
      while (V.tok16[2 * (size_t)slot] != 0 || V.tok16[2 * (size_t)slot + 1] != 0) slot = (slot + 1) & (s16 - 1);
      V.tok16[2 * (size_t)slot] = k.first; V.tok16[2 * (size_t)slot + 1] = k.second;
    }
  }
  if (n_special > TRC_MAX_SPECIALS) return "too many special tokens";
  V.n_sp = (uint32_t)n_special;
  V.sp_off.assign(n_special + 1, 0);
  for (int32_t k = 0; k < n_special; ++k) V.sp_off[k + 1] = V.sp_off[k] + (uint32_t)(special_off[k + 1] - special_off[k]);
  V.sp_bytes.assign(V.sp_off[n_special] + 1, 0);
  if (n_special) memcpy(V.sp_bytes.data(), special_bytes + special_off[0], V.sp_off[n_special]);
  for (int32_t k = 0; k < n_special; ++k)
    if (V.sp_off[k + 1] > V.sp_off[k]) { uint8_t b = V.sp_bytes[V.sp_off[k]]; V.sp_first[b >> 6] |= 1ULL << (b & 63); }
  return "";
}

// The 1110 digit-group strings whose token counts feed trc_ts_tokens, followed by " [".
static inline void trc_digit_probe_strings(std::vector<uint8_t>& bytes, std::vector<int64_t>& off) {
  bytes.clear(); off.assign(1, 0);
  char buf[8];
  for (int len = 1; len <= 3; ++len) {
    int lim = len == 1 ? 10 : (len == 2 ? 100 : 1000);
    for (int v = 0; v < lim; ++v) {
      snprintf(buf, sizeof buf, len == 1 ? "%d" : (len == 2 ? "%02d" : "%03d"), v);
      bytes.insert(bytes.end(), buf, buf + len);
      off.push_back((int64_t)bytes.size());
    }
  }
  bytes.push_back(' '); bytes.push_back('[');
  off.push_back((int64_t)bytes.size());
}
