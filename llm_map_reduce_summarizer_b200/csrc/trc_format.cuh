// Timestamp and number formatting, bit-exact with CPython.
//   trc_fmt_ts      /root/reference/preprocessor.py:91-107 == big_chunkeroosky.py:234-242
//   trc_ts_tokens   token count of "[ts]" pieces under the cl100k pre-tokeniser (digits split 3 by 3)
//   trc_fmt_pct1    f"{x:.1f}" of big_chunkeroosky.py:221 (correctly rounded, ties-to-even on the exact
//                   binary value)
#pragma once
#include "trc_common.cuh"

#define TRC_TS_MAX 28  // "-9223372036854775808" can not occur (hours only printed when > 0): 16+6 digits max

// int(x): truncation toward zero, then Python floor-divmod.
TRC_HD void trc_ts_split(double x, int64_t* h, int* m, int* s) {
  int64_t t;
  if (!(x == x)) t = 0;                        // NaN: Python raises; host rejects NaN before packing
  else if (x >= 9.2e18) t = INT64_MAX;
  else if (x <= -9.2e18) t = INT64_MIN + 1;
  else t = (int64_t)x;
  if (t >= 0 && t < 2147483648LL) {             // the usual case in 32-bit arithmetic (divisions by constants: multiply + shift)
    const uint32_t u = (uint32_t)t, hq = u / 3600u, r32 = u - hq * 3600u, mq = r32 / 60u;
    *h = (int64_t)hq; *m = (int)mq; *s = (int)(r32 - mq * 60u);
    return;
  }
  int64_t hh = t / 3600, r = t % 3600;
  if (r < 0) { r += 3600; hh -= 1; }
  *h = hh; *m = (int)(r / 60); *s = (int)(r % 60);
}

TRC_HD int trc_udigits(uint64_t v, char* tmp) {  // writes reversed digits, returns count
  int n = 0;
  if (v < 4294967296ULL) { uint32_t u = (uint32_t)v; do { const uint32_t q = u / 10u; tmp[n++] = (char)('0' + (u - q * 10u)); u = q; } while (u); return n; }
  do { tmp[n++] = (char)('0' + (v % 10)); v /= 10; } while (v);
  return n;
}

// Writes the formatted timestamp (no terminator) and returns its length.
TRC_HD int trc_fmt_ts(double x, uint8_t* out) {
  int64_t h; int m, s;
  trc_ts_split(x, &h, &m, &s);
  int o = 0;
  if (h > 0) {
    char tmp[24];
    int nd = trc_udigits((uint64_t)h, tmp);
    if (nd < 2) out[o++] = '0';
    for (int k = nd - 1; k >= 0; --k) out[o++] = (uint8_t)tmp[k];
    out[o++] = ':';
  }
  out[o++] = (uint8_t)('0' + m / 10); out[o++] = (uint8_t)('0' + m % 10); out[o++] = ':';
  out[o++] = (uint8_t)('0' + s / 10); out[o++] = (uint8_t)('0' + s % 10);
  return o;
}

TRC_HD int trc_ts_len(double x) {
  int64_t h; int m, s;
  trc_ts_split(x, &h, &m, &s);
  if (h <= 0) return 5;
  char tmp[24];
  int nd = trc_udigits((uint64_t)h, tmp);
  return (nd < 2 ? 2 : nd) + 6;
}

// Number of BPE tokens of the timestamp text alone (digit groups by \p{N}{1,3} from the left, ':' = 1).
TRC_HD int trc_ts_tokens(const TrcTables& T, double x) {
  int64_t h; int m, s;
  trc_ts_split(x, &h, &m, &s);
  int c = T.dcount[10 + m] + 1 + T.dcount[10 + s];
  if (h > 0) {
    char tmp[24];
    int nd = trc_udigits((uint64_t)h, tmp);
    char d[24];
    int n = 0;
    if (nd < 2) d[n++] = '0';
    for (int k = nd - 1; k >= 0; --k) d[n++] = tmp[k];
    for (int p = 0; p < n; p += 3) {
      int g = n - p < 3 ? n - p : 3;
      int v = 0;
      for (int k = 0; k < g; ++k) v = v * 10 + (d[p + k] - '0');
      c += T.dcount[v + (g == 1 ? 0 : (g == 2 ? 10 : 110))];
    }
    c += 1;  // ':'
  }
  return c;
}

// f"{v:.1f}".  Returns length, or -1 when |v| >= 2^59 (caller raises; a percentage never gets there).
TRC_HD int trc_fmt_pct1(double v, uint8_t* out) {
  int o = 0;
  if (!(v == v)) { out[0] = 'n'; out[1] = 'a'; out[2] = 'n'; return 3; }
  union { double d; uint64_t u; } cv; cv.d = v;
  bool neg = (cv.u >> 63) != 0;
  double a = neg ? -v : v;
  if (neg) out[o++] = '-';
  if (a > 1.7976931348623157e308) { out[o++] = 'i'; out[o++] = 'n'; out[o++] = 'f'; return o; }
  if (a >= 576460752303423488.0) return -1;
  uint64_t ip = (uint64_t)a;
  double f = a - (double)ip;          // exact
  uint64_t d = 0; int cmp = -1;       // cmp: remainder vs half  (-1 below, 0 tie, +1 above)
  if (f > 0.0) {
    union { double d; uint64_t u; } cf; cf.d = f;
    int e = (int)((cf.u >> 52) & 0x7FF);
    uint64_t mant = cf.u & 0xFFFFFFFFFFFFFULL;
    int k;  // f = mant * 2^-k
    if (e == 0) { k = 1074; } else { mant |= 1ULL << 52; k = 1075 - e; }
    if (k <= 63) {
      uint64_t t10 = mant * 10ULL;  // < 2^57
      d = t10 >> k;
      uint64_t rem = t10 & ((1ULL << k) - 1), half = 1ULL << (k - 1);
      cmp = rem > half ? 1 : (rem == half ? 0 : -1);
    } else {
      d = 0; cmp = -1;  // mant*10 < 2^57 <= 2^(k-1)/2...: strictly below half
      if (k == 64) { uint64_t t10 = mant * 10ULL; cmp = t10 > (1ULL << 63) ? 1 : (t10 == (1ULL << 63) ? 0 : -1); }
    }
  }
  uint64_t q = ip * 10ULL + d;
  if (cmp > 0 || (cmp == 0 && (q & 1ULL))) q += 1;
  char tmp[24];
  int nd = trc_udigits(q / 10ULL, tmp);
  for (int k2 = nd - 1; k2 >= 0; --k2) out[o++] = (uint8_t)tmp[k2];
  out[o++] = '.';
  out[o++] = (uint8_t)('0' + (q % 10ULL));
  return o;
}

TRC_HD int trc_fmt_uint(uint64_t v, uint8_t* out) {
  char tmp[24];
  int nd = trc_udigits(v, tmp);
  for (int k = 0; k < nd; ++k) out[k] = (uint8_t)tmp[nd - 1 - k];
  return nd;
}
