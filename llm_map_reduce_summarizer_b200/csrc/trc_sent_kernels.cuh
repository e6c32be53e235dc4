// Sentence path: kernels + host staging (included by trchunk.cu after the context definition).
// big_chunkeroosky.py:331-433 (sentence greedy), :437-542 (clause greedy).

// ---- records ---------------------------------------------------------------------------------------
__global__ void k_rl_list(TrcDev D, const int32_t* __restrict__ nsub_g, const int64_t* __restrict__ rl_scan, int32_t* rl_groups,
                          int64_t* rl_len) {
  GSTRIDE(g, D.m) if (nsub_g[g] < 0) {
    int64_t q = rl_scan[g];
    rl_groups[q] = (int32_t)g;
    int32_t j = D.gstart[g];
    rl_len[q] = D.coff[j + 1] - D.coff[j];
  }
}
__global__ void k_rl_flag(TrcDev D, const int32_t* __restrict__ nsub_g, int32_t* flag) { GSTRIDE(g, D.m) flag[g] = nsub_g[g] < 0; }

TRC_HD int64_t rl_find(const int64_t* boff, int64_t n, int64_t pos) {   // last q with boff[q] <= pos
  int64_t lo = 0, hi = n;
  while (hi - lo > 1) { int64_t mid = (lo + hi) >> 1; if (boff[mid] <= pos) lo = mid; else hi = mid; }
  return lo;
}

// The same for the threads of a warp that ask for ascending positions (consecutive cells or records): one of them
// searches, the others step forward from its answer - ten dependent loads per thread became one or two.
TRC_HD int64_t rl_find_near(const int64_t* boff, int64_t n, int64_t pos) {
#if defined(__CUDA_ARCH__)
  const unsigned m = __activemask();
  const int leader = __ffs(m) - 1;
  int64_t q = 0;
  if ((int)(threadIdx.x & 31) == leader) q = rl_find(boff, n, pos);
  q = __shfl_sync(m, q, leader);
  if (boff[q] > pos) return rl_find(boff, n, pos);
  while (q + 1 < n && boff[q + 1] <= pos) ++q;
  return q;
#else
  return rl_find(boff, n, pos);
#endif
}
// Breaks before byte x of the concatenated texts: the scan runs over 8-byte cells (cellscan), the rest is counted.
TRC_HD int64_t pk_breaks_before(const uint8_t* brk, const int64_t* cellscan, int64_t x) {
  int64_t v = cellscan[x >> 3];
  for (int64_t p = x & ~(int64_t)7; p < x; ++p) v += brk[p] != 0;
  return v;
}

// Punkt break decision for every text byte of the over-budget raw segments, a thread per 8 bytes.  Ordinary
// cells (inside one segment, ASCII, no control bytes) are handled on bit masks: only the '.', '?' and '!' bytes
// go through the context test; the domain check of the sentence path (single U+0020 between words, no leading /
// trailing whitespace - it only ever fails for text that did not come through clean_text) is two mask tests.
TRC_HD void pk_break_byte(const TrcDev& D, const TrcTables& T, const uint8_t* s, int64_t n, int64_t i, int32_t j, uint8_t* out) {
  *out = pk_sentend(s[i]) && pk_is_break(T, s, n, i);
  uint8_t c = s[i];
  bool bad = false;
  if (c == ' ') bad = (i == 0 || i + 1 == n || s[i + 1] == ' ');
  else if (c < 0x80) bad = (T.props[c] & TRC_P_SPACE) != 0;
  else if (c >= 0xC0) { int l; uint32_t cp = trc_decode(s, i, n, &l); bad = (T.props[cp] & TRC_P_SPACE) != 0; }
  if (bad) D.doc_err[D.sdoc[j]] = -1;
}
// `cand(q, i, p)` decides the '.', '?' or '!' at byte i of over-budget segment q (byte p of the concatenated texts): the test
// build evaluates it on the spot; the CUDA kernel queues it, because the context test is ~2 000 instructions and ran with
// five of 32 lanes when every thread evaluated the candidates of its own cell.
template <class Cand>
TRC_HD void pk_break_cell(const TrcDev& D, const TrcTables& T, const TrcSentState& S, int64_t total, int64_t cell, uint8_t* brk, Cand&& cand) {
  const int64_t p0 = cell * 8;
  int64_t q = rl_find_near(S.rl_boff, S.n_rl, p0);
  if (p0 + 16 <= S.rl_boff[q + 1]) {                        // (two aligned words are read: both inside the segment)
    const int32_t j = D.gstart[S.rl_groups[q]];
    const int64_t c0 = D.coff[j], n = D.coff[j + 1] - c0, i = p0 - S.rl_boff[q];
    const PkBuf B = pk_buf(D.C);
    const int64_t c = c0 + i + B.mis;
    uint64_t lo, hi;
#if defined(__CUDA_ARCH__)
    lo = *reinterpret_cast<const unsigned long long*>(B.sb + (c & ~7LL)); hi = *reinterpret_cast<const unsigned long long*>(B.sb + (c & ~7LL) + 8);
#else
    memcpy(&lo, B.sb + (c & ~7LL), 8); memcpy(&hi, B.sb + (c & ~7LL) + 8, 8);
#endif
    const unsigned sh = (unsigned)(c & 7) * 8u;
    const uint64_t x = sh ? (lo >> sh) | (hi << (64 - sh)) : lo;
    const uint8_t nextb = i + 8 < n ? (uint8_t)((sh ? hi >> sh : hi) & 0xff) : (uint8_t)'x';
    const uint64_t sp = pk_eq8(x, ' ');
    const uint64_t low = ~(((x & ~SW_H) | SW_H) - SW_L * 0x21) & SW_H;       // ASCII bytes below 0x21
    if (!((x & SW_H) | (low & ~sp))) {
      uint64_t bad = sp & ((sp >> 8) | (nextb == ' ' ? (0x80ULL << 56) : 0ULL));   // two spaces in a row
      if (i == 0) bad |= sp & 0x80ULL;                                              // leading space
      if (n - i <= 8) bad |= sp & (0x80ULL << (8 * (n - 1 - i)));                   // trailing space
      if (bad) D.doc_err[D.sdoc[j]] = -1;
      uint32_t se = sw_pack(pk_eq8(x, '.') | pk_eq8(x, '?') | pk_eq8(x, '!'));
      uint64_t flags = 0;
      while (se) {
        const int k = (int)trc_ctz32(se);
        se &= se - 1;
        if (cand(q, i + k, p0 + k)) flags |= 1ULL << (8 * k);
      }
#if defined(__CUDA_ARCH__)
      *reinterpret_cast<unsigned long long*>(brk + p0) = flags;
#else
      memcpy(brk + p0, &flags, 8);
#endif
      return;
    }
  }
  for (int64_t p = p0; p < p0 + 8 && p < total; ++p) {      // cell across a segment edge, or with unusual bytes
    while (p >= S.rl_boff[q + 1]) ++q;
    const int32_t j = D.gstart[S.rl_groups[q]];
    pk_break_byte(D, T, D.C + D.coff[j], D.coff[j + 1] - D.coff[j], p - S.rl_boff[q], j, brk + p);
  }
}
#ifdef TRC_HOSTSIM
__global__ void k_punkt_breaks(TrcDev D, TrcTables T, TrcSentState S, int64_t total, uint8_t* brk) {
  GSTRIDE(cell, (total + 7) / 8) pk_break_cell(D, T, S, total, cell, brk, [&](int64_t q, int64_t i, int64_t) {
    const int32_t j = D.gstart[S.rl_groups[q]];
    return pk_is_break(T, D.C + D.coff[j], D.coff[j + 1] - D.coff[j], i);
  });
}
#else
#define PK_T 128
#define PK_CELLS 4                             // cells per thread and tile: ~120 candidates per tile, a candidate for every thread
#define PK_LIST (PK_T * 8)
__global__ void __launch_bounds__(PK_T) k_punkt_breaks(TrcDev D, TrcTables T, TrcSentState S, int64_t total, uint8_t* brk) {
  __shared__ int64_t s_p[PK_LIST];
  __shared__ int64_t s_i[PK_LIST];
  __shared__ int32_t s_q[PK_LIST];
  __shared__ int s_n;
  const int64_t ncell = (total + 7) / 8;
  for (int64_t base = (int64_t)blockIdx.x * (PK_T * PK_CELLS); base < ncell; base += (int64_t)gridDim.x * (PK_T * PK_CELLS)) {
    if (threadIdx.x == 0) s_n = 0;
    __syncthreads();
    for (int c = 0; c < PK_CELLS; ++c) {
      const int64_t cell = base + c * PK_T + threadIdx.x;
      if (cell < ncell) pk_break_cell(D, T, S, total, cell, brk, [&](int64_t q, int64_t i, int64_t p) {
        const int slot = atomicAdd(&s_n, 1);
        if (slot >= PK_LIST) {                  // (text that is mostly '.', '?' and '!': decided on the spot)
          const int32_t j = D.gstart[S.rl_groups[q]];
          return pk_is_break(T, D.C + D.coff[j], D.coff[j + 1] - D.coff[j], i);
        }
        s_p[slot] = p; s_i[slot] = i; s_q[slot] = (int32_t)q;
        return false;
      });
    }
    __syncthreads();                            // the cells' flag words are written (zero where a candidate is queued)
    const int nc = s_n < PK_LIST ? s_n : PK_LIST;
    for (int k = threadIdx.x; k < nc; k += PK_T) {
      const int32_t j = D.gstart[S.rl_groups[s_q[k]]];
      if (pk_is_break(T, D.C + D.coff[j], D.coff[j + 1] - D.coff[j], s_i[k])) brk[s_p[k]] = 1;
    }
    __syncthreads();
  }
}
#endif
__global__ void k_cell_counts(const uint8_t* __restrict__ brk, int64_t total, uint8_t* cellcnt) {
  GSTRIDE(cell, (total + 7) / 8) {
    const int64_t p0 = cell * 8, p1 = p0 + 8 < total ? p0 + 8 : total;
    int c = 0;
    for (int64_t p = p0; p < p1; ++p) c += brk[p] != 0;
    cellcnt[cell] = (uint8_t)c;
  }
}
__global__ void k_brk_pos(const uint8_t* __restrict__ brk, const int64_t* __restrict__ cellscan, int64_t total, int64_t* pos) {
  GSTRIDE(cell, (total + 7) / 8) {
    const int64_t p0 = cell * 8, p1 = p0 + 8 < total ? p0 + 8 : total;
    int64_t o = cellscan[cell];
    if (cellscan[cell + 1] == o) continue;
    for (int64_t p = p0; p < p1; ++p) if (brk[p]) pos[o++] = p;
  }
}
__global__ void k_sent_counts(TrcSentState S, const uint8_t* __restrict__ brk, const int64_t* __restrict__ cellscan, int32_t* ns) {
  GSTRIDE(q, S.n_rl) ns[q] = (int32_t)(pk_breaks_before(brk, cellscan, S.rl_boff[q + 1]) - pk_breaks_before(brk, cellscan, S.rl_boff[q])) + 1;
}

// one thread per sentence slice: slice bounds, realignment with its neighbours, strip, code points
__global__ void k_sent_build(TrcDev D, TrcSentState S, const uint8_t* __restrict__ brk, const int64_t* __restrict__ cellscan,
                             const int64_t* __restrict__ brk_pos, int64_t* chars64) {
  GSTRIDE(r, S.n_sent) {
    int64_t q = rl_find_near(S.sent_off, S.n_rl, r);
    int32_t g = S.rl_groups[q];
    int32_t j = D.gstart[g];
    const uint8_t* s = D.C + D.coff[j];
    int64_t n = D.coff[j + 1] - D.coff[j];
    int64_t i = r - S.sent_off[q];
    int64_t nb = S.sent_off[q + 1] - S.sent_off[q] - 1;     // breaks of this segment
    int64_t b0 = pk_breaks_before(brk, cellscan, S.rl_boff[q]);
    auto slice = [&](int64_t k, int64_t* a, int64_t* b) {   // _slices_from_text
      if (k == 0) *a = 0;
      else { int64_t c = brk_pos[b0 + k - 1] - S.rl_boff[q]; *a = pk_nonword(s[c + 1]) ? c + 1 : c + 2; }
      *b = k < nb ? (brk_pos[b0 + k] - S.rl_boff[q]) + 1 : n;
    };
    int64_t a, b, k0 = 0, e0 = 0, k1 = 0, e1 = 0;
    slice(i, &a, &b);
    if (i > 0) pk_realign(s, a, b, &k0, &e0);               // _realign_boundaries
    int64_t sa = a + e0, sb = b;
    if (i < nb) { int64_t a2, b2; slice(i + 1, &a2, &b2); pk_realign(s, a2, b2, &k1, &e1); if (k1) sb = a2 + k1; }
    while (sa < sb && s[sa] == ' ') ++sa;                   // sentence.strip()  (:357)
    while (sb > sa && s[sb - 1] == ' ') --sb;
    if (sb < sa) sb = sa;
    TrcSent& R = S.recs[r];
    R.group = g; R.kind = 1; R.off = D.coff[j] + sa; R.len = sb - sa; R.chars = pk_count_chars(s, sa, sb);
    R.tok = 0; R.sub = -1; R.last_in_sub = 0; R.off_in_sub = 0; R.parent = -1; R.ncl = 0; R.cl0 = 0; R.start = 0; R.end = 0;
    chars64[r] = R.chars;
  }
}

// sentence times (:351-366) and the unit range for token counting
__global__ void k_sent_times(TrcDev D, TrcSentState S, const int64_t* __restrict__ cscan, int64_t* ustart, int64_t* uend) {
  GSTRIDE(r, S.n_sent) {
    TrcSent& R = S.recs[r];
    int64_t q = rl_find_near(S.sent_off, S.n_rl, r);
    int32_t j = D.gstart[R.group];
    int32_t raw = D.k2r[j];
    double st = D.start[raw], en = D.end[raw];
    double span = en - st;
    int32_t chars = D.cchars[raw];
    double tpc = chars > 0 ? span / (double)chars : 0.0;
    int64_t done = cscan[r] - cscan[S.sent_off[q]];
    double m1 = tpc * (double)done;
    R.start = st + m1;
    double m2 = tpc * (double)R.chars;
    R.end = R.start + m2;
    ustart[r] = R.off; uend[r] = R.off + R.len;
  }
}

// Token count of " " + record text, split three ways.  For a record of an oversized segment the words that lie
// completely inside it are a difference of the per-word running sum of stage P (TrcWordSums) -> wsum.  A record
// edge inside a word (a closing quote followed by "--", a clause cut at the dot of "3.5") leaves a partial word:
// the head [a, first space) is counted as its own unit with the virtual leading space, the tail [last space, b)
// as a unit that starts with its real space.  Records of segments that were not pre-tokenised are one head unit.
// Empty unit ranges are skipped by the generic counter; their counts are ignored.
__global__ void k_sent_wordsum(TrcDev D, TrcSentState S, TrcWordSums W, int64_t rec0, int64_t nrec, int64_t* hs, int64_t* he,
                               int64_t* ts, int64_t* te, int64_t* wsum) {
  GSTRIDE(r, nrec) {                           // records rec0 .. rec0+nrec (sentences, or the clause records behind them)
    const TrcSent& R = S.recs[rec0 + r];
    const int64_t a = R.off, b = R.off + R.len;
    int64_t h0 = a, h1 = b, t0 = b, t1 = b, v = 0;
    if (W.of_seg && R.len > 0) {
      const int32_t j = D.gstart[R.group];
      const int32_t q = W.of_seg[D.k2r[j]];
      if (q >= 0) {
        const int64_t c0 = D.coff[j], c1 = D.coff[j + 1], gshift = W.toff[q] - c0;   // position x of C is x + gshift of the giants' buffer
        const bool left = a == c0 || D.C[a - 1] == ' ', right = b == c1 || D.C[b] == ' ';
        int64_t m0 = a, m1 = b;
        bool whole = false;                      // no usable word edge inside: the record is one generic unit
        if (!left) { while (m0 < b && D.C[m0] != ' ') ++m0; if (m0 >= b) whole = true; }
        if (!whole && !right) { m1 = b - 1; while (m1 >= m0 && D.C[m1] != ' ') --m1; if (m1 < m0) whole = true; }
        if (!whole) {
          h1 = left ? a : m0;                    // head: [a, first space) when the left edge is inside a word
          if (!right) t0 = m1;                   // tail: [last space, b) when the right edge is inside a word
          const int64_t w0 = W.wscan[(left ? (a == c0 ? a : a - 1) : m0) + gshift], w1 = W.wscan[m1 + gshift];
          v = W.wpre[w1] - W.wpre[w0];
        }
      }
    }
    else if (W.wtok && R.len > 0) {
      // cut mode: the fused kernel left the token count of every word at the word's first byte (its space, or the first
      // byte of the segment); zero elsewhere, 255 = not known there.  Same head / middle / tail split as above.
      const int32_t j = D.gstart[R.group];
      const int64_t c0 = D.coff[j], c1 = D.coff[j + 1];
      const bool left = a == c0 || D.C[a - 1] == ' ', right = b == c1 || D.C[b] == ' ';
      int64_t m0 = a, m1 = b;
      bool whole = false;
      if (!left) { while (m0 < b && D.C[m0] != ' ') ++m0; if (m0 >= b) whole = true; }
      if (!whole && !right) { m1 = b - 1; while (m1 >= m0 && D.C[m1] != ' ') --m1; if (m1 < m0) whole = true; }
      if (!whole) {
        const int64_t x0 = left ? (a == c0 ? a : a - 1) : m0, x1 = m1;      // words that start in [x0, x1)
        int64_t sum = 0;
        bool unknown = x1 - x0 > 512;        // (a long record is counted directly, a thread per 32 bytes: k_count_units_coop)
        int64_t x = unknown ? x1 : x0;
        for (; x < x1 && (x & 7); ++x) { const uint8_t t = W.wtok[x]; sum += t; unknown |= t == 255; }
        for (; x + 8 <= x1; x += 8) {
          uint64_t q;
#if defined(__CUDA_ARCH__)
          q = *reinterpret_cast<const unsigned long long*>(W.wtok + x);
#else
          memcpy(&q, W.wtok + x, 8);
#endif
          if (!q) continue;
          unknown |= pk_eq8(q, 0xFF) != 0;
          const uint64_t pair = (q & 0x00FF00FF00FF00FFULL) + ((q >> 8) & 0x00FF00FF00FF00FFULL);
          sum += (int64_t)((pair * 0x0001000100010001ULL) >> 48);
        }
        for (; x < x1; ++x) { const uint8_t t = W.wtok[x]; sum += t; unknown |= t == 255; }
        if (!unknown) {
          h1 = left ? a : m0;
          if (!right) t0 = m1;
          v = sum;
        }
      }
    }
    hs[r] = h0; he[r] = h1; ts[r] = t0; te[r] = t1; wsum[r] = v;
  }
}
TRC_HD int32_t sent_unit_tokens(int64_t r, const int64_t* hs, const int64_t* he, const int64_t* ts, const int64_t* te, const int64_t* wsum,
                                const uint32_t* hc, const uint32_t* tc) {
  return (int32_t)wsum[r] + (he[r] > hs[r] ? (int32_t)hc[r] : 0) + (te[r] > ts[r] ? (int32_t)tc[r] : 0);
}

__global__ void k_sent_tok(TrcDev D, TrcTables T, TrcSentState S, const int64_t* hs, const int64_t* he, const int64_t* ts, const int64_t* te,
                           const int64_t* wsum, const uint32_t* hc, const uint32_t* tc, int32_t* ncl,
                           int64_t* ob_list, unsigned long long* ob_cnt) {
  GSTRIDE(r, S.n_sent) {
    TrcSent& R = S.recs[r];
    int32_t n = 0;
    if (R.len > 0) {
      R.tok = 2 + trc_ts_tokens(T, R.start) + sent_unit_tokens(r, hs, he, ts, te, wsum, hc, tc);
      int64_t E = D.doc_E[D.sdoc[D.gstart[R.group]]];
      if (R.tok > E) {
        if (ob_list) ob_list[atomicAdd(ob_cnt, 1ULL)] = r;       // over budget: its clauses are counted by a warp (k_clause_count_w)
        else n = pk_clause_count(D.C, R.off, R.off + R.len);
      }
    }
    R.ncl = n; ncl[r] = n;
  }
}

// one thread per over-budget sentence: clause records with times (:474-506)
__global__ void k_clause_build(TrcDev D, TrcSentState S, const int64_t* __restrict__ clscan, int64_t* ustart, int64_t* uend) {
  GSTRIDE(r, S.n_sent) {
    TrcSent& P = S.recs[r];
    if (P.ncl == 0) continue;
    P.cl0 = S.n_sent + clscan[r];
    double span = P.end - P.start;
    double tpc = P.chars > 0 ? span / (double)P.chars : 0.0;
    int64_t done = 0;
    const uint8_t* s = D.C;
    pk_clauses(s, P.off, P.off + P.len, [&](int32_t k, int64_t ca, int64_t cb) {
      while (ca < cb && s[ca] == ' ') ++ca;
      while (cb > ca && s[cb - 1] == ' ') --cb;
      TrcSent& R = S.recs[P.cl0 + k];
      R.group = P.group; R.kind = 2; R.off = ca; R.len = cb - ca; R.chars = pk_count_chars(s, ca, cb);
      R.tok = 0; R.sub = -1; R.last_in_sub = 0; R.off_in_sub = 0; R.parent = (int32_t)r; R.ncl = 0; R.cl0 = 0;
      double m1 = tpc * (double)done;
      R.start = P.start + m1;
      double m2 = tpc * (double)R.chars;
      R.end = R.start + m2;
      done += R.chars;
      int64_t u = clscan[r] + k;
      ustart[u] = ca; uend[u] = cb;
    });
  }
}
#ifndef TRC_HOSTSIM
// ---- the same clause splitting (pk_clauses_t) by a warp per over-budget sentence ---------------------------------
// An over-budget sentence can be tens of KB; lanes take consecutive 8-byte words, neighbours exchange the edge
// bits with shuffles, clause / word indices are warp prefix sums.  Counting first (the records need their number),
// then the raw clause ranges, then one thread per clause for strip / code points / times.
struct PkWarpWord { PkWord w; uint32_t nxp, nxsp, prevsp; };
__device__ __forceinline__ PkWarpWord pk_warp_word(const PkBuf& B, int64_t base, int64_t W1, int64_t a2, int64_t b2, int lane, uint32_t carry_sp) {
  const int64_t W = base + lane;
  PkWarpWord o;
  o.w.p = 0; o.w.sp = 0xffu; o.w.cont = 0;
  if (W <= W1) o.w = pk_word(B.sb, W, a2, b2);
  o.nxp = __shfl_down_sync(0xffffffffu, o.w.p, 1);
  o.nxsp = __shfl_down_sync(0xffffffffu, o.w.sp, 1);
  if (lane == 31) { o.nxp = 0; o.nxsp = 0xffu; if (W + 1 <= W1) { const PkWord x = pk_word(B.sb, W + 1, a2, b2); o.nxp = x.p; o.nxsp = x.sp; } }
  o.prevsp = __shfl_up_sync(0xffffffffu, o.w.sp, 1);
  if (lane == 0) o.prevsp = carry_sp;
  return o;
}
__device__ __forceinline__ int32_t pk_warp_excl(int32_t v, int lane, int32_t* total) {
  int32_t x = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) { const int32_t y = __shfl_up_sync(0xffffffffu, x, o); if (lane >= o) x += y; }
  *total = __shfl_sync(0xffffffffu, x, 31);
  return x - v;
}
__global__ void k_clause_count_w(TrcDev D, TrcSentState S, const int64_t* __restrict__ ob, int64_t nob, int32_t* ncl, int32_t* ob_words) {
  WSTRIDE(i, nob) {
    const int lane = TRC_LANE;
    const int64_t r = ob[i];
    TrcSent& R = S.recs[r];
    const PkBuf B = pk_buf(D.C);
    const int64_t a2 = R.off + B.mis, b2 = R.off + R.len + B.mis, W0 = a2 >> 3, W1 = (b2 - 1) >> 3;
    int64_t ends = 0, words = 0;
    uint32_t carry = 0xffu;
    for (int64_t base = W0; base <= W1; base += 32) {
      const PkWarpWord x = pk_warp_word(B, base, W1, a2, b2, lane, carry);
      ends += __popc(x.w.p & ~((x.w.p >> 1) | ((x.nxp & 1u) << 7)));
      words += __popc(~x.w.sp & 0xffu & ((x.w.sp << 1) | ((x.prevsp >> 7) & 1u)));
      carry = __shfl_sync(0xffffffffu, x.w.sp, 31);
    }
    ends = trc_lanes_sum(ends); words = trc_lanes_sum(words);
    const int32_t nc = (int32_t)ends - (pk_clause_punct(D.C[R.off]) ? 1 : 0);
    const int32_t n = nc > 0 ? nc : (int32_t)((words + 19) / 20);
    if (lane == 0) { R.ncl = n; ncl[r] = n; ob_words[i] = nc > 0 ? -1 : (int32_t)words; }
  }
}
// raw clause ranges [rs[u], re[u]) of every clause record u (before strip)
__global__ void k_clause_ranges_w(TrcDev D, TrcSentState S, const int64_t* __restrict__ ob, int64_t nob, const int32_t* __restrict__ ob_words,
                                  const int64_t* __restrict__ clscan, int64_t* rs, int64_t* re) {
  WSTRIDE(i, nob) {
    const int lane = TRC_LANE;
    const int64_t r = ob[i];
    TrcSent& R = S.recs[r];
    if (R.ncl == 0) continue;
    const int64_t u0 = clscan[r];
    const int32_t n = R.ncl, words = ob_words[i];
    if (lane == 0) R.cl0 = S.n_sent + u0;
    const PkBuf B = pk_buf(D.C);
    const int64_t a = R.off, a2 = a + B.mis, b2 = a + R.len + B.mis, W0 = a2 >> 3, W1 = (b2 - 1) >> 3;
    uint32_t carry = 0xffu;
    int32_t kbase = 0;
    if (words < 0) {                                       // punctuation runs end the clauses
      const int32_t lead = pk_clause_punct(D.C[a]) ? 1 : 0;
      if (!lead && lane == 0) rs[u0] = a;
      for (int64_t base = W0; base <= W1; base += 32) {
        const PkWarpWord x = pk_warp_word(B, base, W1, a2, b2, lane, carry);
        uint32_t e = x.w.p & ~((x.w.p >> 1) | ((x.nxp & 1u) << 7));
        int32_t tot;
        int32_t k = kbase + pk_warp_excl(__popc(e), lane, &tot) - lead;
        while (e) {
          const int bit = __ffs(e) - 1;
          e &= e - 1;
          const int64_t ce = 8 * (base + lane) + bit + 1 - B.mis;
          if (k >= 0 && k < n) re[u0 + k] = ce;
          if (k + 1 < n) rs[u0 + k + 1] = ce;               // also the leading run's end (k == -1): start of clause 0
          ++k;
        }
        kbase += tot;
        carry = __shfl_sync(0xffffffffu, x.w.sp, 31);
      }
    } else {                                               // groups of 20 words
      for (int64_t base = W0; base <= W1; base += 32) {
        const PkWarpWord x = pk_warp_word(B, base, W1, a2, b2, lane, carry);
        const uint32_t ws = ~x.w.sp & 0xffu & ((x.w.sp << 1) | ((x.prevsp >> 7) & 1u));
        const uint32_t we = ~x.w.sp & 0xffu & ((x.w.sp >> 1) | ((x.nxsp & 1u) << 7));
        int32_t tot;
        const int32_t w0 = kbase + pk_warp_excl(__popc(ws), lane, &tot);   // words that start before this lane's word
        uint32_t t = ws;
        while (t) {
          const int bit = __ffs(t) - 1;
          t &= t - 1;
          const int32_t idx = w0 + __popc(ws & ((1u << bit) - 1u));
          if (idx % 20 == 0) rs[u0 + idx / 20] = 8 * (base + lane) + bit - B.mis;
        }
        t = we;
        while (t) {
          const int bit = __ffs(t) - 1;
          t &= t - 1;
          const int32_t idx = w0 + __popc(ws & ((2u << bit) - 1u)) - 1;   // the word that ends here
          if ((idx + 1) % 20 == 0 || idx + 1 == words) re[u0 + idx / 20] = 8 * (base + lane) + bit + 1 - B.mis;
        }
        kbase += tot;
        carry = __shfl_sync(0xffffffffu, x.w.sp, 31);
      }
    }
  }
}
__global__ void k_clause_recs(TrcDev D, TrcSentState S, const int64_t* __restrict__ clscan, int64_t n_cl, const int64_t* __restrict__ rs,
                              const int64_t* __restrict__ re, int64_t* chars64) {
  GSTRIDE(u, n_cl) {
    const int64_t r = rl_find(clscan, S.n_sent, u);
    const TrcSent& P = S.recs[r];
    const uint8_t* s = D.C;
    int64_t ca = rs[u], cb = re[u];
    while (ca < cb && s[ca] == ' ') ++ca;
    while (cb > ca && s[cb - 1] == ' ') --cb;
    TrcSent& R = S.recs[S.n_sent + u];
    R.group = P.group; R.kind = 2; R.off = ca; R.len = cb - ca; R.chars = pk_count_chars(s, ca, cb);
    R.tok = 0; R.sub = -1; R.last_in_sub = 0; R.off_in_sub = 0; R.parent = (int32_t)r; R.ncl = 0; R.cl0 = 0; R.start = 0; R.end = 0;
    chars64[u] = R.chars;
  }
}
__global__ void k_clause_times(TrcSentState S, const int64_t* __restrict__ clscan, int64_t n_cl, const int64_t* __restrict__ ccs,
                               int64_t* ustart, int64_t* uend) {
  GSTRIDE(u, n_cl) {
    TrcSent& R = S.recs[S.n_sent + u];
    const TrcSent& P = S.recs[R.parent];
    const double span = P.end - P.start;
    const double tpc = P.chars > 0 ? span / (double)P.chars : 0.0;
    const int64_t done = ccs[u] - ccs[clscan[R.parent]];
    const double m1 = tpc * (double)done;
    R.start = P.start + m1;
    const double m2 = tpc * (double)R.chars;
    R.end = R.start + m2;
    ustart[u] = R.off; uend[u] = R.off + R.len;
  }
}
#endif
__global__ void k_clause_tok(TrcTables T, TrcSentState S, const int64_t* hs, const int64_t* he, const int64_t* ts, const int64_t* te,
                             const int64_t* wsum, const uint32_t* hc, const uint32_t* tc, int64_t n_cl) {
  GSTRIDE(u, n_cl) {
    TrcSent& R = S.recs[S.n_sent + u];
    R.tok = R.len > 0 ? 2 + trc_ts_tokens(T, R.start) + sent_unit_tokens(u, hs, he, ts, te, wsum, hc, tc) : 0;
  }
}

// ---- nested greedy, one thread per over-budget raw segment -----------------------------------------------
struct SubAcc {
  int64_t tok, text_len; int32_t r0, r1, n; double start, end; int32_t start_src;
};

// The greedy below is sequential per segment; it is the form the test build runs (a thread per segment).  The CUDA build
// runs the same decisions in sg_warp (further down): a warp per segment, 32 records per load.
struct SentLite { int64_t len, cl0, piece; double start, end; int32_t tok, ncl; };
TRC_HD SentLite sent_lite(const TrcSent& R) {
  SentLite o; o.len = R.len; o.cl0 = R.cl0; o.start = R.start; o.end = R.end; o.tok = R.tok; o.ncl = R.ncl;
  o.piece = 1 + trc_ts_len(R.start) + 2 + R.len;
  return o;
}
// A record as the greedy sees it.
struct SentViewDirect {
  SentLite v;
  TRC_HD int32_t tok() const { return v.tok; }
  TRC_HD int64_t piece() const { return v.piece; }
  TRC_HD double start() const { return v.start; }
  TRC_HD double end() const { return v.end; }
  TRC_HD int64_t cl0() const { return v.cl0; }
  TRC_HD int32_t ncl() const { return v.ncl; }
};
struct SentFetchDirect {
  const TrcSent* recs;
  TRC_HD SentViewDirect operator()(int64_t r) { return SentViewDirect{sent_lite(recs[r])}; }
};

template <bool WRITE>
TRC_HD void sg_flush(TrcSentState& S, TrcSub* out, int32_t& ns, int32_t g, int32_t kind, SubAcc& a, bool wr) {
  if (WRITE && wr) {
    TrcSub& s = out[ns];
    s.group = g; s.kind = kind; s.r0 = a.r0; s.r1 = a.r1; s.start_src = a.start_src; s.end_src = -1; s.tok = (int32_t)a.tok; s.pad = 0;
    s.start = a.start; s.end = a.end; s.text_len = a.text_len;
    S.recs[a.r1].last_in_sub = 1;
  }
  ++ns;
}
// Appends record r (tok tokens) to the sub-chunk under construction, which will be flushed as number ns (nothing else is
// flushed between a sub-chunk's first append and its own flush).  Only the writing pass needs the record's text length
// and end time.
template <bool WRITE, class View>
TRC_HD void sg_append(TrcSentState& S, SubAcc& a, int32_t r, const View& R, int32_t tok, int64_t sub_index, bool wr) {
  if (a.n == 0) { a.r0 = r; a.text_len = 0; }
  if (WRITE) {
    const int64_t piece = R.piece();
    if (wr) { S.recs[r].sub = (int32_t)sub_index; S.recs[r].off_in_sub = a.n ? a.text_len + 1 : 0; }
    a.text_len += piece + (a.n ? 1 : 0);
    a.end = R.end();
  }
  a.tok += tok; a.r1 = r; a.n += 1;
}

// (a record is empty exactly when its token count is 0: k_sent_tok / k_clause_tok give every non-empty record 2 + ...)
template <bool WRITE, class Fetch>
TRC_HD int32_t trc_split_sentences(const TrcDev& D, TrcSentState S, int64_t q, TrcSub* out, int64_t sub_base, Fetch&& fetch, bool wr) {
  int32_t g = S.rl_groups[q];
  int32_t j = D.gstart[g];
  int64_t E = D.doc_E[D.sdoc[j]];
  int32_t ns = 0;
  SubAcc cur; cur.tok = 0; cur.text_len = 0; cur.n = 0; cur.r0 = cur.r1 = -1; cur.start = D.start[D.k2r[j]]; cur.end = 0; cur.start_src = j;
  const int64_t r_end = S.sent_off[q + 1];
  for (int64_t r = S.sent_off[q]; r < r_end; ++r) {
    const auto R = fetch(r);
    const int32_t tok = R.tok();
    if (tok == 0) continue;
    if (tok > E) {
      // clause sub-chunks of this sentence (:437-542)
      const double r_start = R.start(), r_endt = R.end();
      const int64_t cl0 = R.cl0(), ncl = R.ncl();
      SubAcc cl; cl.tok = 0; cl.text_len = 0; cl.n = 0; cl.r0 = cl.r1 = -1; cl.start = r_start; cl.end = 0; cl.start_src = -1;
      // an existing sub-chunk is closed first (:378-381), but it is appended BEFORE the clause chunks
      if (cur.tok > 0) { cur.end = r_start; sg_flush<WRITE>(S, out, ns, g, 1, cur, wr); }
      for (int64_t c = cl0; c < cl0 + ncl; ++c) {
        const SentViewDirect C{sent_lite(S.recs[c])};
        if (C.v.len == 0) continue;
        if (cl.tok + C.v.tok > E && cl.tok > 0) {
          sg_flush<WRITE>(S, out, ns, g, 2, cl, wr);
          cl.tok = 0; cl.text_len = 0; cl.n = 0; cl.start = C.v.start; cl.start_src = -1;
        }
        sg_append<WRITE>(S, cl, (int32_t)c, C, C.v.tok, sub_base + ns, wr);
      }
      if (cl.tok > 0) sg_flush<WRITE>(S, out, ns, g, 2, cl, wr);
      cur.tok = 0; cur.text_len = 0; cur.n = 0; cur.start = r_endt; cur.start_src = -1;    // :387-399
    } else if (cur.tok + tok > E && cur.tok > 0) {
      if (WRITE) cur.end = R.start();                                                      // :404
      sg_flush<WRITE>(S, out, ns, g, 1, cur, wr);
      cur.tok = 0; cur.text_len = 0; cur.n = 0; cur.start_src = -1;                        // :409-421
      if (WRITE) cur.start = cur.end;
      sg_append<WRITE>(S, cur, (int32_t)r, R, tok, sub_base + ns, wr);
    } else {
      sg_append<WRITE>(S, cur, (int32_t)r, R, tok, sub_base + ns, wr);                     // :423-429
    }
  }
  if (cur.tok > 0) sg_flush<WRITE>(S, out, ns, g, 1, cur, wr);
  return ns;
}

#ifdef TRC_HOSTSIM
__global__ void k_sent_greedy_count(TrcDev D, TrcSentState S, int32_t* nsub_g) {
  GSTRIDE(q, S.n_rl) nsub_g[S.rl_groups[q]] = trc_split_sentences<false>(D, S, q, nullptr, 0, SentFetchDirect{S.recs}, true);
}
__global__ void k_sent_greedy_write(TrcDev D, TrcSentState S) {
  GSTRIDE(q, S.n_rl) { int64_t base = D.sub_off[S.rl_groups[q]]; trc_split_sentences<true>(D, S, q, D.subs + base, base, SentFetchDirect{S.recs}, true); }
}
#else
// The same greedy for a warp, written for the loop it is: ~1 400 dependent iterations per segment in a warp that has the
// SM quadrant almost to itself, so every instruction of an iteration is paid in full (profiles: 73 instructions per
// sentence through the generic code, most of them 64-bit index arithmetic and struct traffic).  Lanes load 32 records'
// token counts (and, when writing, text lengths and start times) at once; the uniform inner loop moves one or two values
// per record between lanes and keeps 32-bit indices; what a record gets (sub-chunk, offset) is parked in the lane that
// loaded it and stored 32 records at a time.  Over-budget sentences (rare) go through the generic clause code.
template <bool WRITE>
__device__ int32_t sg_warp(const TrcDev& D, TrcSentState S, int64_t q, TrcSub* out, int64_t sub_base, int lane) {
  const int32_t g = S.rl_groups[q];
  const int32_t j = D.gstart[g];
  const int64_t E = D.doc_E[D.sdoc[j]];
  const int64_t rbase = S.sent_off[q];
  const int32_t n = (int32_t)(S.sent_off[q + 1] - rbase);
  TrcSent* const recs = S.recs + rbase;
  const bool wr = lane == 0;
  int32_t ns = 0;
  SubAcc cur; cur.tok = 0; cur.text_len = 0; cur.n = 0; cur.r0 = cur.r1 = -1; cur.start = D.start[D.k2r[j]]; cur.end = 0; cur.start_src = j;
  // (the next window's records are requested before the current window is walked: their latency hides under the walk)
  int32_t ntok = 0; double nstart = 0.0; int64_t nlen = 0;
  if (lane < n) { ntok = recs[lane].tok; if (WRITE) { nstart = recs[lane].start; nlen = recs[lane].len; } }
  for (int32_t base = 0; base < n; base += 32) {
    const int32_t k = base + lane;
    const int32_t mytok = ntok; const double mystart = nstart;
    int32_t mypiece = 0;
    if (WRITE && k < n) mypiece = (int32_t)(1 + trc_ts_len(mystart) + 2 + nlen);
    ntok = 0;
    if (k + 32 < n) { ntok = recs[k + 32].tok; if (WRITE) { nstart = recs[k + 32].start; nlen = recs[k + 32].len; } }
    int32_t mysub = -1; int64_t myoff = 0;
    const int cnt = n - base < 32 ? n - base : 32;
    for (int t = 0; t < cnt; ++t) {
      const int32_t tok = __shfl_sync(0xffffffffu, mytok, t);
      if (tok == 0) continue;                                   // an empty record (see trc_split_sentences)
      const int32_t r = (int32_t)(rbase + base + t);            // record index as the sub-chunks store it
      if (tok > E) {
        const TrcSent& R = S.recs[r];
        const double r_start = R.start, r_endt = R.end;
        const int64_t cl0 = R.cl0, ncl = R.ncl;
        SubAcc cl; cl.tok = 0; cl.text_len = 0; cl.n = 0; cl.r0 = cl.r1 = -1; cl.start = r_start; cl.end = 0; cl.start_src = -1;
        if (cur.tok > 0) { cur.end = r_start; sg_flush<WRITE>(S, out, ns, g, 1, cur, wr); }
        for (int64_t c = cl0; c < cl0 + ncl; ++c) {
          const SentViewDirect C{sent_lite(S.recs[c])};
          if (C.v.len == 0) continue;
          if (cl.tok + C.v.tok > E && cl.tok > 0) {
            sg_flush<WRITE>(S, out, ns, g, 2, cl, wr);
            cl.tok = 0; cl.text_len = 0; cl.n = 0; cl.start = C.v.start; cl.start_src = -1;
          }
          sg_append<WRITE>(S, cl, (int32_t)c, C, C.v.tok, sub_base + ns, wr);
        }
        if (cl.tok > 0) sg_flush<WRITE>(S, out, ns, g, 2, cl, wr);
        cur.tok = 0; cur.text_len = 0; cur.n = 0; cur.start = r_endt; cur.start_src = -1;
        continue;
      }
      if (cur.tok + tok > E && cur.tok > 0) {
        if (WRITE) cur.end = __shfl_sync(0xffffffffu, mystart, t);
        sg_flush<WRITE>(S, out, ns, g, 1, cur, wr);
        cur.tok = 0; cur.text_len = 0; cur.n = 0; cur.start_src = -1;
        if (WRITE) cur.start = cur.end;
      }
      if (cur.n == 0) { cur.r0 = r; cur.text_len = 0; }
      if (WRITE) {
        const int32_t piece = __shfl_sync(0xffffffffu, mypiece, t);
        if (lane == t) { mysub = (int32_t)(sub_base + ns); myoff = cur.n ? cur.text_len + 1 : 0; }
        cur.text_len += piece + (cur.n ? 1 : 0);
      }
      cur.tok += tok; cur.r1 = r; cur.n += 1;
    }
    if (WRITE && mysub >= 0) { recs[k].sub = mysub; recs[k].off_in_sub = myoff; }
  }
  if (cur.tok > 0) {
    if (WRITE) cur.end = S.recs[cur.r1].end;                    // the end of the last record appended
    sg_flush<WRITE>(S, out, ns, g, 1, cur, wr);
  }
  return ns;
}
__global__ void k_sent_greedy_count(TrcDev D, TrcSentState S, int32_t* nsub_g) {
  WSTRIDE(q, S.n_rl) {
    const int32_t n = sg_warp<false>(D, S, q, nullptr, 0, TRC_LANE);
    if (TRC_LANE == 0) nsub_g[S.rl_groups[q]] = n;
  }
}
__global__ void k_sent_greedy_write(TrcDev D, TrcSentState S) {
  WSTRIDE(q, S.n_rl) {
    const int64_t base = D.sub_off[S.rl_groups[q]];
    sg_warp<true>(D, S, q, D.subs + base, base, TRC_LANE);
  }
}
#endif

// K6 for sentence/clause records: "[ts] text" (+ joining space) into the item text; warp per record (test build), eight
// lanes per record (CUDA build: a sentence is ~35 bytes; a warp each left three quarters of the lanes without a byte)
#ifndef TRC_HOSTSIM
// Where the text of every sentence / clause sub-chunk goes (one lookup chain per sub-chunk instead of one per record), and
// the "\n\n" in front of it when it is not the first item of its chunk.
__global__ void k_sub_dst(TrcDev D, int64_t* sub_dst) {
  GSTRIDE(sub, D.n_subs) {
    if (D.subs[sub].kind == 0) { sub_dst[sub] = -1; continue; }
    const int32_t grp = D.subs[sub].group;
    const int64_t t = D.item_off[grp] + (sub - D.sub_off[grp]);
    const int64_t d = D.it_dst[t];
    sub_dst[sub] = d;
    if (t > D.chunks[D.cscanf[t + 1] - 1].item0) { D.twc[d - 2] = '\n'; D.twc[d - 1] = '\n'; }
  }
}
__global__ void k_emit_sent_recs(TrcDev D, TrcSentState S, const int64_t* __restrict__ sub_dst) {
  for (int64_t q = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; (q >> 3) < S.n_recs; q += (int64_t)gridDim.x * blockDim.x) {
    const int64_t r = q >> 3;
    const int lane = (int)(q & 7);
    const TrcSent* Rp = S.recs + r;
    const int32_t sub = Rp->sub;
    if (sub < 0) continue;
    const int64_t off = Rp->off, len = Rp->len, ois = Rp->off_in_sub;
    const double st = Rp->start;
    const int32_t last = Rp->last_in_sub;
    uint8_t* dst = D.twc + sub_dst[sub] + ois;
    int64_t h; int m, sec;
    trc_ts_split(st, &h, &m, &sec);
    int o;
    if (h <= 0) {                          // "[MM:SS] ": one byte per lane
      const unsigned long long pre = (unsigned long long)'[' | ((unsigned long long)('0' + m / 10) << 8) | ((unsigned long long)('0' + m % 10) << 16) |
                                     ((unsigned long long)':' << 24) | ((unsigned long long)('0' + sec / 10) << 32) |
                                     ((unsigned long long)('0' + sec % 10) << 40) | ((unsigned long long)']' << 48) | ((unsigned long long)' ' << 56);
      dst[lane] = (uint8_t)(pre >> (8 * lane));
      o = 8;
    } else {
      uint8_t buf[TRC_TS_MAX + 4];
      const int tl = trc_fmt_ts(st, buf + 1);
      buf[0] = '['; buf[tl + 1] = ']'; buf[tl + 2] = ' ';
      for (int k = lane; k < tl + 3; k += 8) dst[k] = buf[k];
      o = tl + 3;
    }
    const uint8_t* src = D.C + off;
    uint8_t* db = dst + o;
    if (!last && lane == 7) db[len] = ' ';
    if (len < 24) { for (int64_t k = lane; k < len; k += 8) db[k] = src[k]; continue; }
    // 4-byte words aligned to the destination (two aligned source words + a funnel shift each), as in k_emit_pieces_flat
    const int head = (int)((4 - ((uintptr_t)db & 3)) & 3);
    if (lane < head) db[lane] = src[lane];
    const int64_t nw = (len - head) >> 2;
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
#else
__global__ void k_emit_sent_recs(TrcDev D, TrcSentState S) {
  WSTRIDE(r, S.n_recs) {
    const TrcSent R = S.recs[r];
    if (R.sub < 0) continue;
    const TrcSub& sb = D.subs[R.sub];
    int64_t t = D.item_off[sb.group] + (R.sub - D.sub_off[sb.group]);
    uint8_t* base = D.twc + D.it_dst[t];
    int lane = TRC_LANE, nl = TRC_NL;
    if (R.off_in_sub == 0 && lane == 0 && t > D.chunks[D.cscanf[t + 1] - 1].item0) { base[-2] = '\n'; base[-1] = '\n'; }
    uint8_t* dst = base + R.off_in_sub;
    uint8_t buf[TRC_TS_MAX + 4];
    int tl = trc_fmt_ts(R.start, buf + 1);
    buf[0] = '['; buf[tl + 1] = ']'; buf[tl + 2] = ' ';
    for (int k = lane; k < tl + 3; k += nl) dst[k] = buf[k];
    const uint8_t* src = D.C + R.off;
    for (int64_t k = lane; k < R.len; k += nl) dst[tl + 3 + k] = src[k];
    if (!R.last_in_sub && lane == 0) dst[tl + 3 + R.len] = ' ';
  }
}
#endif

// ---- host staging -----------------------------------------------------------------------------------------
// (count_units is defined in trchunk.cu before this file is included)

static int sentence_stage_count(trc_ctx* ctx, cudaStream_t st, int32_t* nsub_g, int32_t nraw) {
  TrcDev& D = ctx->D;
  TrcSentState& S = ctx->SS;
  int rc;
  const int64_t m = D.m;
  // list of over-budget raw segments and their byte prefix
  int32_t* flag; int64_t* rl_scan; int32_t* rl_groups; int64_t* rl_len; int64_t* rl_boff;
  ALLOC(flag, int32_t, m); ALLOC(rl_scan, int64_t, m + 1); ALLOC(rl_groups, int32_t, nraw); ALLOC(rl_len, int64_t, nraw);
  ALLOC(rl_boff, int64_t, nraw + 1);
  LAUNCH(k_rl_flag, m, 256, D, nsub_g, flag);
  if ((rc = scan_excl(ctx, st, flag, rl_scan, m))) return rc;
  LAUNCH(k_rl_list, m, 256, D, nsub_g, rl_scan, rl_groups, rl_len);
  if ((rc = scan_excl(ctx, st, rl_len, rl_boff, (int64_t)nraw))) return rc;
  int64_t total = 0;
  if ((rc = read_back(ctx, st, rl_boff + nraw, 8, &total))) return rc;
  S.rl_groups = rl_groups; S.n_rl = nraw; S.rl_boff = rl_boff;
  // Punkt breaks
  // (the break flags are scanned per 8-byte cell: an eighth of the scan over bytes, positions inside a cell are counted)
  uint8_t* brk; uint8_t* cellcnt; int64_t* cellscan;
  const int64_t ncell = (total + 7) / 8;
  ALLOC(brk, uint8_t, total + 8); ALLOC(cellcnt, uint8_t, ncell); ALLOC(cellscan, int64_t, ncell + 1);
#ifdef TRC_HOSTSIM
  LAUNCH(k_punkt_breaks, ncell, 128, D, ctx->T, S, total, brk);
#else
  LAUNCH(k_punkt_breaks, (ncell + PK_CELLS - 1) / PK_CELLS, PK_T, D, ctx->T, S, total, brk);
#endif
  LAUNCH(k_cell_counts, ncell, 256, brk, total, cellcnt);
  if ((rc = scan_excl(ctx, st, cellcnt, cellscan, ncell))) return rc;
  int64_t nbrk = 0;
  if ((rc = read_back(ctx, st, cellscan + ncell, 8, &nbrk))) return rc;
  int64_t* brk_pos; int32_t* ns; int64_t* sent_off;
  ALLOC(brk_pos, int64_t, nbrk); ALLOC(ns, int32_t, nraw); ALLOC(sent_off, int64_t, nraw + 1);
  LAUNCH(k_brk_pos, ncell, 256, brk, cellscan, total, brk_pos);
  LAUNCH(k_sent_counts, (int64_t)nraw, 128, S, brk, cellscan, ns);
  if ((rc = scan_excl(ctx, st, ns, sent_off, (int64_t)nraw))) return rc;
  S.sent_off = sent_off;
  S.n_sent = nbrk + nraw;
  // sentence records (the clause records are appended once their number is known)
  TrcSent* recs0; int64_t* chars64; int64_t* cscan; int64_t* ustart; int64_t* uend; uint32_t* cnt; int32_t* ncl; int64_t* clscan;
  ALLOC(recs0, TrcSent, S.n_sent); ALLOC(chars64, int64_t, S.n_sent); ALLOC(cscan, int64_t, S.n_sent + 1);
  ALLOC(ustart, int64_t, S.n_sent); ALLOC(uend, int64_t, S.n_sent); ALLOC(cnt, uint32_t, S.n_sent); ALLOC(ncl, int32_t, S.n_sent);
  ALLOC(clscan, int64_t, S.n_sent + 1);
  S.recs = recs0; S.n_recs = S.n_sent;
  LAUNCH(k_sent_build, S.n_sent, 128, D, S, brk, cellscan, brk_pos, chars64);
  if ((rc = scan_excl(ctx, st, chars64, cscan, S.n_sent))) return rc;
  LAUNCH(k_sent_times, S.n_sent, 128, D, S, cscan, ustart, uend);
  int64_t* wsum; int64_t* tst; int64_t* ten; uint32_t* tcnt;
  ALLOC(wsum, int64_t, S.n_sent); ALLOC(tst, int64_t, S.n_sent); ALLOC(ten, int64_t, S.n_sent); ALLOC(tcnt, uint32_t, S.n_sent);
  LAUNCH(k_sent_wordsum, S.n_sent, 128, D, S, ctx->WS, (int64_t)0, S.n_sent, ustart, uend, tst, ten, wsum);
  if ((rc = count_units(ctx, st, D.C, ustart, uend, S.n_sent, total, 1, cnt, nullptr, false, 128))) return rc;
  if (ctx->WS.of_seg || ctx->WS.wtok) { if ((rc = count_units(ctx, st, D.C, tst, ten, S.n_sent, total, 0, tcnt, nullptr, false, 128))) return rc; }
#ifdef TRC_HOSTSIM
  LAUNCH(k_sent_tok, S.n_sent, 128, D, ctx->T, S, ustart, uend, tst, ten, wsum, cnt, tcnt, ncl, (int64_t*)nullptr, (unsigned long long*)nullptr);
#else
  int64_t* ob; unsigned long long* ob_cnt; int32_t* ob_words = nullptr;
  ALLOC(ob, int64_t, S.n_sent); ALLOC(ob_cnt, unsigned long long, 1);
  CK(cudaMemsetAsync(ob_cnt, 0, 8, st));
  LAUNCH(k_sent_tok, S.n_sent, 128, D, ctx->T, S, ustart, uend, tst, ten, wsum, cnt, tcnt, ncl, ob, ob_cnt);
  int64_t nob = 0;
  if ((rc = read_back(ctx, st, ob_cnt, 8, &nob))) return rc;
  if (nob > 0) {
    ALLOC(ob_words, int32_t, nob);
    LAUNCH_W(k_clause_count_w, nob, D, S, ob, nob, ncl, ob_words);
  }
#endif
  if ((rc = scan_excl(ctx, st, ncl, clscan, S.n_sent))) return rc;
  int64_t ncls = 0;
  if ((rc = read_back(ctx, st, clscan + S.n_sent, 8, &ncls))) return rc;
  if (ncls > 0) {
    TrcSent* recs;
    ALLOC(recs, TrcSent, S.n_sent + ncls);
    CK(cudaMemcpyAsync(recs, recs0, sizeof(TrcSent) * (size_t)S.n_sent, cudaMemcpyDeviceToDevice, st));
    S.recs = recs; S.n_recs = S.n_sent + ncls;
    int64_t* cus; int64_t* cue; uint32_t* ccnt;
    ALLOC(cus, int64_t, ncls); ALLOC(cue, int64_t, ncls); ALLOC(ccnt, uint32_t, ncls);
#ifdef TRC_HOSTSIM
    LAUNCH(k_clause_build, S.n_sent, 64, D, S, clscan, cus, cue);
#else
    {
      int64_t* rs; int64_t* re; int64_t* cch; int64_t* ccs;
      ALLOC(rs, int64_t, ncls); ALLOC(re, int64_t, ncls); ALLOC(cch, int64_t, ncls); ALLOC(ccs, int64_t, ncls + 1);
      LAUNCH_W(k_clause_ranges_w, nob, D, S, ob, nob, ob_words, clscan, rs, re);
      LAUNCH(k_clause_recs, ncls, 128, D, S, clscan, ncls, rs, re, cch);
      if ((rc = scan_excl(ctx, st, cch, ccs, ncls))) return rc;
      LAUNCH(k_clause_times, ncls, 128, S, clscan, ncls, ccs, cus, cue);
    }
#endif
    int64_t* cwsum; int64_t* cts; int64_t* cte; uint32_t* ctc;
    ALLOC(cwsum, int64_t, ncls); ALLOC(cts, int64_t, ncls); ALLOC(cte, int64_t, ncls); ALLOC(ctc, uint32_t, ncls);
    LAUNCH(k_sent_wordsum, ncls, 128, D, S, ctx->WS, S.n_sent, ncls, cus, cue, cts, cte, cwsum);
    if ((rc = count_units(ctx, st, D.C, cus, cue, ncls, total, 1, ccnt, nullptr, false, 128))) return rc;
    if (ctx->WS.of_seg || ctx->WS.wtok) { if ((rc = count_units(ctx, st, D.C, cts, cte, ncls, total, 0, ctc, nullptr, false, 128))) return rc; }
    LAUNCH(k_clause_tok, ncls, 128, ctx->T, S, cus, cue, cts, cte, cwsum, ccnt, ctc, ncls);
  }
  LAUNCH_W(k_sent_greedy_count, (int64_t)nraw, D, S, nsub_g);
  return TRC_OK;
}

static int sentence_stage_write(trc_ctx* ctx, cudaStream_t st, const int32_t*) {
  LAUNCH_W(k_sent_greedy_write, ctx->SS.n_rl, ctx->D, ctx->SS);
  return TRC_OK;
}
static int sentence_stage_emit(trc_ctx* ctx, cudaStream_t st) {
#ifdef TRC_HOSTSIM
  LAUNCH_W(k_emit_sent_recs, ctx->SS.n_recs, ctx->D, ctx->SS);
#else
  int64_t* sub_dst;
  ALLOC(sub_dst, int64_t, ctx->D.n_subs);
  LAUNCH(k_sub_dst, ctx->D.n_subs, 128, ctx->D, sub_dst);
  LAUNCH(k_emit_sent_recs, ctx->SS.n_recs * 8, 256, ctx->D, ctx->SS, sub_dst);
#endif
  return TRC_OK;
}
