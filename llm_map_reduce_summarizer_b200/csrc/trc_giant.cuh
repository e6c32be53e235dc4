// Segments too large for a fused tile ("giants": a monologue of tens of KB in one segment, config 4).
// They are cleaned and tokenised BEFORE the fused kernel by byte-parallel grid kernels over global memory
// (the per-byte local rules of trc_clean_par.cuh, preprocessor.py:38,69-89), so that the fused kernel's
// slow lane only has to copy their cleaned text and take over their counts instead of walking tens of KB
// in one thread.  Tokens: every space of the cleaned text starts a word; words are counted as units
// (big_chunkeroosky.py:83 via DESIGN.md §3) and summed per giant.
#pragma once

#define TRC_GIANT_MIN 4096          // raw bytes; larger segments are pre-processed here

struct GiantState {
  const int32_t* list;              // raw segment index of every giant (ascending)
  const int64_t* boff;              // prefix of their raw lengths, ng+1
  int64_t ng, gb;                   // giants, raw bytes
  uint8_t* cls; uint8_t* code;      // per raw byte
  int32_t* fns;                     // first non-space byte per giant
  int64_t* pk;                      // per raw byte: output bytes | code points << 32   (scanned in place -> opk)
  int64_t* opk;                     // exclusive scan of pk, gb+1
  uint8_t* text;                    // cleaned text of all giants back to back
  int64_t* toff;                    // cleaned offset of every giant, ng+1
  int32_t* chars;                   // cleaned code points
  uint32_t* tok;                    // tokens of " " + cleaned text
  long long* spw;                   // first special-token hit: (word << 4 | k), or LLONG_MAX
  const int32_t* of_seg;            // per raw segment: giant index or -1 (n entries)
};

// cnt[0]: oversized segments.  cnt[1 + q]: tiles of cand.v[q] segments whose raw bytes would not fit a
// tile of the fused kernel (`cap` bytes from the 16-byte aligned start) - the host picks the tile size from these.
struct TileCand { int v[4]; int all8; };   // all8: every candidate is a multiple of 8 (only every eighth segment can start a tile)
__global__ void k_giant_any(TrcDev D, unsigned long long* cnt, TileCand cand, int64_t cap) {
  int found = 0;
  GSTRIDE(i, D.n) {
    const int64_t a = D.seg_off[i];
    found += (D.seg_off[i + 1] - a) > TRC_GIANT_MIN;
    if (cand.all8 && (i & 7)) continue;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int ts = cand.v[q];
      if (ts > 0 && (uint32_t)i % (uint32_t)ts == 0u) {            // (segment indices fit 31 bits: k2r is int32)
        const int64_t e = i + ts < D.n ? i + ts : D.n;
        if (D.seg_off[e] - (a & ~(int64_t)15) > cap) atomicAdd(cnt + 1 + q, 1ULL);
      }
    }
  }
  if (found) atomicAdd(cnt, (unsigned long long)found);       // (rare: most batches have none)
}
// Tiles of varying size for batches whose segments do not fit tiles of a fixed segment count (oversized segments, heavy
// tails): a tile is a run of segments that start inside the same window of `win` bytes, at most `max_segs` of them, and an
// segment of more than `iso` bytes is a tile of its own.  The raw bytes of any other tile span less than win + iso + 16.
__global__ void k_tile_flags(TrcDev D, int32_t* flag, int64_t win, int max_segs, int64_t iso) {
  GSTRIDE(i, D.n) {
    int f = (uint32_t)i % (uint32_t)max_segs == 0u;
    if (!f) {
      const int64_t a = D.seg_off[i], p = D.seg_off[i - 1], b = D.seg_off[i + 1];
      f = (a / win != p / win) || (b - a > iso) || (a - p > iso);
    }
    flag[i] = f;
  }
}
__global__ void k_giant_flag(TrcDev D, int32_t* flag) { GSTRIDE(i, D.n) flag[i] = (D.seg_off[i + 1] - D.seg_off[i]) > TRC_GIANT_MIN; }
__global__ void k_giant_list(TrcDev D, const int32_t* __restrict__ flag, const int64_t* __restrict__ scan, int32_t* list, int64_t* len,
                             int32_t* of_seg) {
  GSTRIDE(i, D.n) {
    if (flag[i]) { int64_t q = scan[i]; list[q] = (int32_t)i; len[q] = D.seg_off[i + 1] - D.seg_off[i]; of_seg[i] = (int32_t)q; }
    else of_seg[i] = -1;
  }
}
// ---- cutting oversized segments into parts (generation 5) ----------------------------------------------------------------
// An oversized segment is handed to the fused kernel as PARTS of (normally) at most TRC_CUT_STEP + TRC_CUT_SEARCH bytes.  A cut sits
// in front of a whitespace run that follows an ASCII non-whitespace byte, so the part before it ends in text and the
// part after it starts with the run; the kernel lets a continuing part keep that run (one space), and the cleaned parts
// back to back ARE the cleaned segment, provided no rule of clean_text looks across the cut:
//  * whitespace collapse and strip (preprocessor.py:72): the run lies inside one part;
//  * duplicate words (:76): only fires if the word before the run equals the word after it - a cut is taken only where
//    they differ in their last byte, where one of them is no \w-run at all, or where nothing follows the run;
//  * space after . ! ? (:80): needs a letter right after the mark, and a whitespace byte follows the part's last byte.
// One thread per cut target (every TRC_CUT_STEP bytes) looks for the first such place within TRC_CUT_SEARCH bytes.  A
// target without one is skipped; when that leaves a part too long for a tile (kilobytes of one repeated word, non-ASCII
// text without ASCII whitespace) *bad is set and the batch takes the byte-parallel pre-pass below instead.
#define TRC_CUT_STEP 1024
#define TRC_CUT_SEARCH 512
#define TRC_CUT_ISO 2048             // parts are at most STEP + SEARCH bytes unless a target had no cut: longer ones get a tile each
__global__ void k_cut_counts(const int64_t* __restrict__ len, int64_t ng, int32_t* tcnt) { GSTRIDE(q, ng) tcnt[q] = (int32_t)((len[q] - 1) / TRC_CUT_STEP); }
__global__ void k_cut_find(TrcDev D, TrcTables T, const int32_t* __restrict__ list, const int64_t* __restrict__ len, const int64_t* __restrict__ tpre,
                           int64_t ng, int64_t nt, int64_t* cutpos) {
  GSTRIDE(t, nt) {
    int64_t lo = 0, hi = ng;
    while (hi - lo > 1) { const int64_t mid = (lo + hi) >> 1; if (tpre[mid] <= t) lo = mid; else hi = mid; }
    const int64_t c = t - tpre[lo] + 1;                       // the c-th target of this segment
    const int64_t s = D.seg_off[list[lo]], L = len[lo];
    const uint8_t* p = D.text + s;
    int64_t k = c * TRC_CUT_STEP;
    const int64_t lim = k + TRC_CUT_SEARCH < L ? k + TRC_CUT_SEARCH : L;
    int64_t found = -1;
    while (k < lim) {
      const uint8_t b = p[k], a = p[k - 1];
      if (!(b < 0x80 && (T.props[b] & TRC_P_SPACE) && a < 0x80 && !(T.props[a] & TRC_P_SPACE))) { ++k; continue; }
      int64_t r = k + 1;                                      // the whitespace run [k, r)
      bool ascii = true;
      while (r < L) { const uint8_t x = p[r]; if (x >= 0x80) { ascii = false; break; } if (!(T.props[x] & TRC_P_SPACE)) break; ++r; }
      if (!ascii) { k = r + 1; continue; }                    // the run may go on with a multi-byte character: not decided here
      if (r == L) { found = k; break; }                       // nothing but whitespace follows
      int64_t e = r;                                          // the \w-run [r, e) after it
      while (e < L) { const uint8_t x = p[e]; if (x >= 0x80) { ascii = false; break; } if (!(T.props[x] & TRC_P_WORD)) break; ++e; }
      if (!ascii) { k = e + 1; continue; }
      if (e == r || !(T.props[a] & TRC_P_WORD) || p[e - 1] != a) { found = k; break; }
      k = e;                                                  // same last byte: the words may be equal
    }
    cutpos[t] = found < 0 ? -1 : s + found;                   // -1: no cut at this target (k_vseg_build makes the part before it longer)
  }
}
// The segment list the fused kernel sees: the caller's segments with every oversized one replaced by its parts.
// A target without a cut leaves an EMPTY part behind (its start is the next part's start), so the part in front of it grows
// by a step; it still has to fit a tile of the fused kernel on its own (max_part), else *bad is set.
__global__ void k_vseg_build(TrcDev D, const int32_t* __restrict__ gflag, const int64_t* __restrict__ gscan, const int64_t* __restrict__ tpre,
                             const int64_t* __restrict__ cutpos, int64_t nt, int64_t* vseg_off, int32_t* vorig, uint8_t* vcont,
                             int64_t max_part, unsigned long long* bad) {
  GSTRIDE(i, D.n) {
    const int64_t q = gscan[i];                               // oversized segments before i
    const int64_t v = i + tpre[q];
    vseg_off[v] = D.seg_off[i]; vorig[v] = (int32_t)i; vcont[v] = 0;
    if (gflag[i]) {
      const int64_t t0 = tpre[q], t1 = tpre[q + 1];
      int64_t nxt = D.seg_off[i + 1];                         // start of the part after the one being laid down
      bool too_long = false;
      for (int64_t t = t1 - 1; t >= t0; --t) {
        const int64_t u = v + 1 + (t - t0);
        const int64_t cpos = cutpos[t] < 0 ? nxt : cutpos[t];
        too_long |= nxt - cpos > max_part;
        vseg_off[u] = cpos; vorig[u] = (int32_t)i; vcont[u] = 1;
        nxt = cpos;
      }
      too_long |= nxt - D.seg_off[i] > max_part;
      if (too_long) atomicAdd(bad, 1ULL);
    }
    if (i == D.n - 1) vseg_off[D.n + nt] = D.seg_off[D.n];
  }
}
__global__ void k_special_from_pos(const unsigned long long* __restrict__ pos, int64_t nk, int32_t* seg_special) {
  GSTRIDE(j, nk) seg_special[j] = pos[j] == ~0ULL ? -1 : (int32_t)(pos[j] & 15ULL);
}
struct CutView { int64_t nv; const int64_t* vseg_off; const int32_t* vorig; const uint8_t* vcont; unsigned long long* bad; };

__device__ __forceinline__ int64_t giant_of_byte(const GiantState& G, int64_t p) {
  int64_t lo = 0, hi = G.ng;
  while (hi - lo > 1) { int64_t mid = (lo + hi) >> 1; if (G.boff[mid] <= p) lo = mid; else hi = mid; }
  return lo;
}
__global__ void k_giant_classify(TrcDev D, TrcTables T, GiantState G) {
  GSTRIDE(p, G.gb) {
    const int64_t q = giant_of_byte(G, p);
    const int64_t s = D.seg_off[G.list[q]];
    const int32_t n = (int32_t)(G.boff[q + 1] - G.boff[q]);
    cp_classify(T, D.text + s, G.cls + G.boff[q], G.code + G.boff[q], (int32_t)(p - G.boff[q]), n);
  }
}
__global__ void k_giant_fns(GiantState G) {
  GSTRIDE(q, G.ng) G.fns[q] = cp_first_nonspace(G.cls + G.boff[q], 0, (int32_t)(G.boff[q + 1] - G.boff[q]));
}
__global__ void k_giant_markdup(TrcDev D, GiantState G) {
  GSTRIDE(p, G.gb) {
    const int64_t q = giant_of_byte(G, p);
    const int32_t k = (int32_t)(p - G.boff[q]), n = (int32_t)(G.boff[q + 1] - G.boff[q]);
    const uint8_t* cls = G.cls + G.boff[q];
    if (cp_dup_candidate(cls, k, 0)) cp_mark_dup(D.text + D.seg_off[G.list[q]], cls, G.code + G.boff[q], k, 0, n);
  }
}
__global__ void k_giant_code(TrcDev D, GiantState G) {
  GSTRIDE(p, G.gb) {
    const int64_t q = giant_of_byte(G, p);
    const int32_t k = (int32_t)(p - G.boff[q]), n = (int32_t)(G.boff[q + 1] - G.boff[q]);
    const uint8_t* cls = G.cls + G.boff[q];
    const uint32_t r = cp_code(D.text + D.seg_off[G.list[q]], cls, G.code + G.boff[q], k, 0, n, G.fns[q]);
    const int64_t ch = (r == 3) ? 2 : ((r == 2 || (r == 1 && !(cls[k] & CP_CONT))) ? 1 : 0);
    G.pk[p] = (int64_t)cp_outlen(r) | (ch << 32);
  }
}
__global__ void k_giant_offsets(GiantState G) {
  GSTRIDE(q, G.ng + 1) {
    const int64_t v = G.opk[G.boff[q]];
    G.toff[q] = v & 0xffffffffLL;
    if (q > 0) G.chars[q - 1] = (int32_t)((v >> 32) - (G.opk[G.boff[q - 1]] >> 32));
    if (q < G.ng) { G.tok[q] = 0; G.spw[q] = 0x7fffffffffffffffLL; }
  }
}
__global__ void k_giant_scatter(TrcDev D, GiantState G) {
  GSTRIDE(p, G.gb) {
    const uint8_t r = G.code[p];
    if (!r) continue;
    const int64_t q = giant_of_byte(G, p);
    const int64_t o = G.opk[p] & 0xffffffffLL;
    const uint8_t b = D.text[D.seg_off[G.list[q]] + (p - G.boff[q])];
    if (r == 1) G.text[o] = b; else if (r == 2) G.text[o] = ' '; else { G.text[o] = ' '; G.text[o + 1] = b; }
  }
}
// words of the cleaned giants: a word starts at every space and at every giant's first byte
__global__ void k_giant_wflag(GiantState G, int64_t gc, int32_t* wflag, const int64_t* __restrict__ toff) {
  GSTRIDE(x, gc) wflag[x] = G.text[x] == ' ';
}
__global__ void k_giant_first(GiantState G, int32_t* wflag) { GSTRIDE(q, G.ng) if (G.toff[q + 1] > G.toff[q]) wflag[G.toff[q]] = 1; }
__global__ void k_giant_words(GiantState G, int64_t gc, const int32_t* __restrict__ wflag, const int64_t* __restrict__ wscan, int64_t* ustart) {
  GSTRIDE(x, gc) if (wflag[x]) ustart[wscan[x]] = x;
}
__global__ void k_giant_wends(GiantState G, int64_t nw, int64_t gc, const int64_t* __restrict__ ustart, int64_t* uend, int32_t* wgiant, uint8_t* wfirst) {
  GSTRIDE(w, nw) {
    const int64_t x = ustart[w];
    int64_t lo = 0, hi = G.ng;                       // giant of cleaned byte x
    while (hi - lo > 1) { int64_t mid = (lo + hi) >> 1; if (G.toff[mid] <= x) lo = mid; else hi = mid; }
    while (lo + 1 < G.ng && G.toff[lo + 1] <= x) ++lo;   // skip blank giants that share the offset
    const int64_t gend = G.toff[lo + 1];
    const int64_t nx = w + 1 < nw ? ustart[w + 1] : gc;
    uend[w] = nx < gend ? nx : gend;
    wgiant[w] = (int32_t)lo;
    wfirst[w] = x == G.toff[lo] && G.text[x] != ' ';
  }
}
// first word of every giant (it gets the virtual leading space) as its own unit list
__global__ void k_giant_firstunits(GiantState G, const int64_t* __restrict__ wscan, const int64_t* __restrict__ uend, int64_t* fs, int64_t* fe) {
  GSTRIDE(q, G.ng) {
    const int64_t x = G.toff[q];
    fs[q] = x;
    fe[q] = G.toff[q + 1] > x ? uend[wscan[x]] : x;
  }
}
// final count of every word (the first word of a giant takes the count made with the virtual leading space);
// special-token hits are rare and go through an atomic
__global__ void k_giant_wcount(GiantState G, int64_t nw, const int32_t* __restrict__ wgiant, const uint32_t* __restrict__ c0,
                               const uint32_t* __restrict__ c1, const uint8_t* __restrict__ wfirst, const int32_t* __restrict__ sp, int32_t* wc) {
  GSTRIDE(w, nw) {
    const int32_t q = wgiant[w];
    wc[w] = (int32_t)(wfirst[w] ? c1[q] : c0[w]);
    if (sp[w] >= 0) atomicMin(&G.spw[q], (long long)((w << 4) | sp[w]));
  }
}
__global__ void k_giant_totals(GiantState G, const int64_t* __restrict__ wscan, const int64_t* __restrict__ wpre) {
  GSTRIDE(q, G.ng) G.tok[q] = (uint32_t)(wpre[wscan[G.toff[q + 1]]] - wpre[wscan[G.toff[q]]]);
}
