// Fixed time intervals over the processed segments: /root/reference/preprocessor.py:217-324
// (aggregate_by_time_interval), SURVEY.md §8(f)1.  Included by trchunk.cu after the launch helpers.
//
// The reference tests every (interval, segment) pair (:247-257) and, for combined segments, every
// (interval, component) pair (:268-275).  Here every merged segment finds its own window of intervals with
// four binary searches over the interval bounds (they are monotone in the interval index, fp rounding
// included), evaluates the reference's predicate exactly inside the window, and the (interval, segment)
// incidences are transposed with a counting pass; the reference's two stable sorts by 'start' (:283,:302) are
// ranks computed inside each (small) list.  Interval bounds use the reference's own expression order
// (start_time + i*interval, min(lo + interval, end_time), :241-242) in fp64.
#pragma once
#include <math.h>
#include "trc_stages.cuh"

struct TrcAggDoc { double t0, t1; int64_t n; };

struct TrcAggRec {            // == trc_agg (include/trchunk.h)
  int32_t doc, interval_index;
  double start, end;
  int64_t mem0, mem1;         // members [mem0, mem1) in TRC_A_AGG_MEM, sorted by start (stable)
  int64_t text_off, text_len; // the aggregated 'text' in TRC_A_AGG_TEXT
  uint8_t ts[48];             // start_formatted / end_formatted, 24-byte slots [len][chars]
};
struct TrcAggMemRec {         // == trc_agg_mem
  int32_t group, n_pieces;    // merged segment; number of surviving components (0: not a combined segment)
  int64_t piece0;             // its components [piece0, piece0 + n_pieces) in TRC_A_AGG_PIECE, sorted by start (stable)
  int64_t text_off, text_len; // the member's own text (after its "[ts speaker] " prefix) in TRC_A_AGG_TEXT
};

struct TrcAggArgs {
  double dt;
  TrcAggDoc* docs; const int64_t* int_off;      // per document; first interval id of every document (n_doc+1)
  int32_t* gcnt; int32_t* gfcnt;                // per merged segment: member entries, surviving components over them
  const int64_t* eoff; const int64_t* foff;     // their exclusive scans
  int32_t* icnt; const int64_t* ioff; int32_t* icur;   // per interval: members, scan, fill cursor
  int32_t* ent_group; int64_t* ent_int; int32_t* ent_f; int64_t* ent_tlen; int64_t* ent_foff;
  int32_t* ent_piece;
  int64_t* ilist; int64_t* sorted; int64_t* mlen; const int64_t* moff;
  int32_t* aflag; const int64_t* aoff;
  TrcAggRec* rec; TrcAggMemRec* mem; uint8_t* text; int64_t* adoc_off;
  int64_t n_int, n_ent;
};

TRC_HD int32_t trc_agg_bump(int32_t* p) {
#if defined(__CUDA_ARCH__)
  return atomicAdd(p, 1);
#else
  int32_t o = *p; *p = o + 1; return o;
#endif
}
TRC_HD double trc_agg_lo(const TrcAggDoc& a, double dt, int64_t i) { return a.t0 + (double)i * dt; }          // :241
TRC_HD double trc_agg_hi(const TrcAggDoc& a, double dt, double lo) { double h = lo + dt; return a.t1 < h ? a.t1 : h; }   // :242
// :253-256 and :270-273
TRC_HD bool trc_agg_in(double lo, double hi, double s, double e) { return (lo <= s && s < hi) || (s <= lo && e > lo); }

// math.ceil((end_time - start_time) / interval_seconds), :236; 0 when the range would be empty.
TRC_HD void trc_agg_doc(const TrcDev& D, double dt, int64_t x, TrcAggDoc* out, int32_t* nint) {
  TrcAggDoc a; a.t0 = a.t1 = 0.0; a.n = 0;
  int64_t g0 = D.gdoc_off[x], g1 = D.gdoc_off[x + 1];
  if (g1 > g0) {
    a.t0 = D.start[D.k2r[D.gstart[g0]]];
    a.t1 = D.end[D.k2r[D.gstart[g1] - 1]];
    double q = (a.t1 - a.t0) / dt;
    if (q > 0.0) {
      double c = ceil(q);
      a.n = c >= 2147483646.0 ? 2147483646 : (int64_t)c;
    }
  }
  *out = a; *nint = (int32_t)a.n;
}

// Window [i0, i1) of interval indices outside which the predicate is false for the span [s, e].
TRC_HD void trc_agg_window(const TrcAggDoc& a, double dt, double s, double e, int64_t* i0, int64_t* i1) {
  int64_t lo = 0, hi = a.n;                     // first i with  hi(i) > s  or  lo(i) >= s
  while (lo < hi) {
    int64_t mid = lo + ((hi - lo) >> 1);
    double l = trc_agg_lo(a, dt, mid);
    if (trc_agg_hi(a, dt, l) > s || l >= s) hi = mid; else lo = mid + 1;
  }
  *i0 = lo;
  hi = a.n;                                     // first i with  lo(i) > s  and  lo(i) >= e
  while (lo < hi) {
    int64_t mid = lo + ((hi - lo) >> 1);
    double l = trc_agg_lo(a, dt, mid);
    if (l > s && l >= e) hi = mid; else lo = mid + 1;
  }
  *i1 = lo;
}

TRC_HD int64_t trc_agg_piece_text_len(const TrcDev& D, int32_t j) {   // "[ts] text", :284-285
  return 3 + trc_ts_len(D.start[D.k2r[j]]) + (D.coff[j + 1] - D.coff[j]);
}

// One merged segment: its member entries (interval-major inside the segment).  WRITE=false counts.
template <bool WRITE>
TRC_HD void trc_agg_group(const TrcDev& D, const TrcAggArgs& A, int64_t g) {
  const int32_t a = D.gstart[g], b = D.gstart[g + 1];
  const int32_t x = D.sdoc[a];
  const TrcAggDoc doc = A.docs[x];
  const double s = D.start[D.k2r[a]], e = D.end[D.k2r[b - 1]];
  const bool comb = b - a > 1;                  // 'segment_timestamps' in segment, :262
  int64_t i0, i1;
  trc_agg_window(doc, A.dt, s, e, &i0, &i1);
  int32_t cnt = 0; int64_t fcnt = 0;
  bool sorted_pieces = true;
  if (WRITE && comb) for (int32_t j = a + 1; j < b; ++j) if (D.start[D.k2r[j]] < D.start[D.k2r[j - 1]]) { sorted_pieces = false; break; }
  for (int64_t i = i0; i < i1; ++i) {
    const double lo = trc_agg_lo(doc, A.dt, i), hi = trc_agg_hi(doc, A.dt, lo);
    if (!trc_agg_in(lo, hi, s, e)) continue;
    int32_t f = 0; int64_t tlen = 0;
    if (comb) {
      for (int32_t j = a; j < b; ++j) {
        const int32_t r = D.k2r[j];
        if (trc_agg_in(lo, hi, D.start[r], D.end[r])) { ++f; if (WRITE) tlen += trc_agg_piece_text_len(D, j) + 1; }
      }
      if (f == 0) continue;                     // :277 no surviving component: the segment is left out
      tlen -= 1;                                // " ".join
    } else if (WRITE) tlen = D.coff[a + 1] - D.coff[a];
    if (WRITE) {
      const int64_t ent = A.eoff[g] + cnt, I = A.int_off[x] + i, fo = A.foff[g] + fcnt;
      A.ent_group[ent] = (int32_t)g; A.ent_int[ent] = I; A.ent_f[ent] = f; A.ent_tlen[ent] = tlen; A.ent_foff[ent] = fo;
      if (comb) {                               // surviving components, sorted by start (stable), :283
        int32_t w = 0;
        for (int32_t j = a; j < b; ++j) {
          const int32_t r = D.k2r[j];
          if (!trc_agg_in(lo, hi, D.start[r], D.end[r])) continue;
          int32_t pos = w;
          if (!sorted_pieces) {
            pos = 0;
            const double sj = D.start[r];
            for (int32_t k = a; k < b; ++k) {
              const int32_t rk = D.k2r[k];
              if (k != j && trc_agg_in(lo, hi, D.start[rk], D.end[rk]) && (D.start[rk] < sj || (D.start[rk] == sj && k < j))) ++pos;
            }
          }
          A.ent_piece[fo + pos] = j;
          ++w;
        }
      }
      const int32_t slot = trc_agg_bump(&A.icur[I]);
      A.ilist[A.ioff[I] + slot] = ent;
    } else {
      trc_agg_bump(&A.icnt[A.int_off[x] + i]);
    }
    ++cnt; fcnt += f;
  }
  if (!WRITE) { A.gcnt[g] = cnt; A.gfcnt[g] = (int32_t)(fcnt > 2147483647 ? 2147483647 : fcnt); }
}

// Position of one member inside its interval after the stable sort by start (:302), and the bytes it
// contributes to the aggregated text: ("\n\n")? "[" ts " " speaker "] " text, :303-307.
TRC_HD void trc_agg_rank(const TrcDev& D, const TrcAggArgs& A, int64_t p) {
  const int64_t ent = A.ilist[p], I = A.ent_int[ent];
  const int64_t a = A.ioff[I], b = A.ioff[I + 1];
  const int32_t g = A.ent_group[ent];
  const int32_t r = D.k2r[D.gstart[g]];
  const double s = D.start[r];
  int64_t rank = 0;
  for (int64_t q = a; q < b; ++q) {
    const int32_t g2 = A.ent_group[A.ilist[q]];
    const double s2 = D.start[D.k2r[D.gstart[g2]]];
    if (s2 < s || (s2 == s && g2 < g)) ++rank;
  }
  const int32_t sp = D.spk[r];
  A.sorted[a + rank] = ent;
  A.mlen[a + rank] = (rank ? 2 : 0) + 1 + trc_ts_len(s) + 1 + (D.spk_off[sp + 1] - D.spk_off[sp]) + 2 + A.ent_tlen[ent];
}

// Writes one member (lanes cooperate) and its record.
TRC_HD void trc_agg_write(const TrcDev& D, const TrcAggArgs& A, int64_t q, int lane, int nl) {
  const int64_t ent = A.sorted[q], I = A.ent_int[ent];
  const int32_t g = A.ent_group[ent];
  const int32_t pa = D.gstart[g];
  const int32_t r = D.k2r[pa];
  const int32_t sp = D.spk[r];
  uint8_t* dst = A.text + A.moff[q];
  int64_t o = 0;
  uint8_t buf[TRC_TS_MAX + 8];
  int n = 0;
  if (q > A.ioff[I]) { buf[n++] = '\n'; buf[n++] = '\n'; }
  buf[n++] = '[';
  n += trc_fmt_ts(D.start[r], buf + n);
  buf[n++] = ' ';
  for (int k = lane; k < n; k += nl) dst[k] = buf[k];
  o = n;
  const int64_t sl = D.spk_off[sp + 1] - D.spk_off[sp];
  for (int64_t k = lane; k < sl; k += nl) dst[o + k] = D.spk_bytes[D.spk_off[sp] + k];
  o += sl;
  if (lane == 0) { dst[o] = ']'; dst[o + 1] = ' '; }
  o += 2;
  const int32_t f = A.ent_f[ent];
  if (f == 0) {
    const uint8_t* src = D.C + D.coff[pa];
    const int64_t len = D.coff[pa + 1] - D.coff[pa];
    for (int64_t k = lane; k < len; k += nl) dst[o + k] = src[k];
  } else {
    int64_t w = o;
    for (int32_t t = 0; t < f; ++t) {
      const int32_t j = A.ent_piece[A.ent_foff[ent] + t];
      trc_emit_piece(D, j, dst + w, true, t + 1 < f, lane, nl);
      w += trc_agg_piece_text_len(D, j) + 1;
    }
  }
  if (lane == 0) {
    TrcAggMemRec m;
    m.group = g; m.n_pieces = f; m.piece0 = A.ent_foff[ent]; m.text_off = A.moff[q] + o; m.text_len = A.ent_tlen[ent];
    A.mem[q] = m;
  }
}

// One record per non-empty interval, :294-319.
TRC_HD void trc_agg_record(const TrcDev& D, const TrcAggArgs& A, int64_t I) {
  if (!A.aflag[I]) return;
  int64_t lo = 0, hi = D.n_doc;                 // document of interval I: last x with int_off[x] <= I
  while (hi - lo > 1) { int64_t mid = (lo + hi) >> 1; if (A.int_off[mid] <= I) lo = mid; else hi = mid; }
  const TrcAggDoc doc = A.docs[lo];
  const int64_t i = I - A.int_off[lo];
  TrcAggRec R;
  R.doc = (int32_t)lo; R.interval_index = (int32_t)i;
  R.start = trc_agg_lo(doc, A.dt, i); R.end = trc_agg_hi(doc, A.dt, R.start);
  R.mem0 = A.ioff[I]; R.mem1 = A.ioff[I + 1];
  R.text_off = A.moff[R.mem0]; R.text_len = A.moff[R.mem1] - R.text_off;
  for (int k = 0; k < 48; ++k) R.ts[k] = 0;
  R.ts[0] = (uint8_t)trc_fmt_ts(R.start, R.ts + 1);
  R.ts[24] = (uint8_t)trc_fmt_ts(R.end, R.ts + 25);
  A.rec[A.aoff[I]] = R;
}

__global__ void k_agg_doc(TrcDev D, TrcAggArgs A, int32_t* nint) { GSTRIDE(x, D.n_doc) trc_agg_doc(D, A.dt, x, A.docs + x, nint + x); }
__global__ void k_agg_count(TrcDev D, TrcAggArgs A) { GSTRIDE(g, D.m) trc_agg_group<false>(D, A, g); }
__global__ void k_agg_fill(TrcDev D, TrcAggArgs A) { GSTRIDE(g, D.m) trc_agg_group<true>(D, A, g); }
__global__ void k_agg_flag(TrcAggArgs A) { GSTRIDE(I, A.n_int) A.aflag[I] = A.icnt[I] > 0; }
__global__ void k_agg_rank(TrcDev D, TrcAggArgs A) { GSTRIDE(p, A.n_ent) trc_agg_rank(D, A, p); }
__global__ void k_agg_write(TrcDev D, TrcAggArgs A) { WSTRIDE(q, A.n_ent) trc_agg_write(D, A, q, TRC_LANE, TRC_NL); }
__global__ void k_agg_record(TrcDev D, TrcAggArgs A) { GSTRIDE(I, A.n_int) trc_agg_record(D, A, I); }
__global__ void k_agg_doc_off(TrcDev D, TrcAggArgs A) { GSTRIDE(x, D.n_doc + 1) A.adoc_off[x] = A.aoff[A.int_off[x]]; }

#define TRC_AGG_MAX_INTERVALS ((int64_t)1 << 27)
#define TRC_AGG_MAX_ENTRIES ((int64_t)1 << 28)

static int aggregate_stage(trc_ctx* ctx, cudaStream_t st, double dt) {
  TrcDev& D = ctx->D;
  int rc;
  const int64_t nd = D.n_doc, m = D.m;
  TrcAggArgs A{};
  A.dt = dt;
  int32_t* nint; int64_t* int_off;
  ALLOC(A.docs, TrcAggDoc, nd); ALLOC(nint, int32_t, nd); ALLOC(int_off, int64_t, nd + 1);
  LAUNCH(k_agg_doc, nd, 128, D, A, nint);
  if ((rc = scan_excl(ctx, st, nint, int_off, nd))) return rc;
  A.int_off = int_off;
  int64_t ni = 0;
  if ((rc = read_back(ctx, st, int_off + nd, 8, &ni))) return rc;
  if (ni > TRC_AGG_MAX_INTERVALS) { ctx->err = "aggregate: more than 2^27 intervals in the batch"; return TRC_EINVAL; }
  A.n_int = ni;
  int64_t *eoff, *foff, *ioff;
  ALLOC(A.gcnt, int32_t, m); ALLOC(A.gfcnt, int32_t, m); ALLOC(eoff, int64_t, m + 1); ALLOC(foff, int64_t, m + 1);
  ALLOC(A.icnt, int32_t, ni); ALLOC(A.icur, int32_t, ni); ALLOC(ioff, int64_t, ni + 1); ALLOC(A.aflag, int32_t, ni);
  CK(cudaMemsetAsync(A.icnt, 0, 4 * (size_t)(ni > 0 ? ni : 1), st));
  CK(cudaMemsetAsync(A.icur, 0, 4 * (size_t)(ni > 0 ? ni : 1), st));
  LAUNCH(k_agg_count, m, 128, D, A);
  if ((rc = scan_excl(ctx, st, A.gcnt, eoff, m))) return rc;
  if ((rc = scan_excl(ctx, st, A.gfcnt, foff, m))) return rc;
  if ((rc = scan_excl(ctx, st, A.icnt, ioff, ni))) return rc;
  A.eoff = eoff; A.foff = foff; A.ioff = ioff;
  int64_t h2[2] = {0, 0};
  if ((rc = read_back(ctx, st, eoff + m, 8, &h2[0]))) return rc;
  if ((rc = read_back(ctx, st, foff + m, 8, &h2[1]))) return rc;
  const int64_t ne = h2[0], nf = h2[1];
  if (ne > TRC_AGG_MAX_ENTRIES || nf > TRC_AGG_MAX_ENTRIES) { ctx->err = "aggregate: more than 2^28 (interval, segment) incidences"; return TRC_EINVAL; }
  A.n_ent = ne;
  ALLOC(A.ent_group, int32_t, ne); ALLOC(A.ent_int, int64_t, ne); ALLOC(A.ent_f, int32_t, ne); ALLOC(A.ent_tlen, int64_t, ne);
  ALLOC(A.ent_foff, int64_t, ne); ALLOC(A.ent_piece, int32_t, nf); ALLOC(A.ilist, int64_t, ne); ALLOC(A.sorted, int64_t, ne);
  int64_t *moff, *aoff;
  ALLOC(A.mlen, int64_t, ne); ALLOC(moff, int64_t, ne + 1); ALLOC(aoff, int64_t, ni + 1);
  LAUNCH(k_agg_fill, m, 128, D, A);
  LAUNCH(k_agg_flag, ni, 256, A);
  LAUNCH(k_agg_rank, ne, 128, D, A);
  if ((rc = scan_excl(ctx, st, A.mlen, moff, ne))) return rc;
  if ((rc = scan_excl(ctx, st, A.aflag, aoff, ni))) return rc;
  A.moff = moff; A.aoff = aoff;
  if ((rc = read_back(ctx, st, moff + ne, 8, &h2[0]))) return rc;
  if ((rc = read_back(ctx, st, aoff + ni, 8, &h2[1]))) return rc;
  const int64_t bt = h2[0], na = h2[1];
  ALLOC(A.rec, TrcAggRec, na); ALLOC(A.mem, TrcAggMemRec, ne); ALLOC(A.text, uint8_t, bt); ALLOC(A.adoc_off, int64_t, nd + 1);
  LAUNCH_W(k_agg_write, ne, D, A);
  LAUNCH(k_agg_record, ni, 128, D, A);
  LAUNCH(k_agg_doc_off, nd + 1, 128, D, A);
  ctx->counts[TRC_N_AGG] = na; ctx->counts[TRC_N_AGG_MEM] = ne; ctx->counts[TRC_B_AGG] = bt;
  set_arr(ctx, TRC_A_AGG, A.rec, (int64_t)sizeof(TrcAggRec) * na);
  set_arr(ctx, TRC_A_AGG_MEM, A.mem, (int64_t)sizeof(TrcAggMemRec) * ne);
  set_arr(ctx, TRC_A_AGG_PIECE, A.ent_piece, 4 * nf);
  set_arr(ctx, TRC_A_AGG_TEXT, A.text, bt);
  set_arr(ctx, TRC_A_AGG_DOC_OFF, A.adoc_off, 8 * (nd + 1));
  return TRC_OK;
}
