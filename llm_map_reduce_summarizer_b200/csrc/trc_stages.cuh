// Per-unit routines of the preprocess and chunk stages (everything except cleaning and tokenising).
// Each routine is the body of one kernel in trchunk.cu; citations give the reference lines it replaces.
#pragma once
#include "trc_common.cuh"
#include "trc_format.cuh"

// ----------------------------------------------------------------------------------------------------
// Device-resident state of one batch (all pointers are device pointers; host mirror in trchunk.cu).
// Index spaces:  raw segment i in [0,n) ; kept segment ("piece") j in [0,n_kept) ; merged segment
// ("group") g in [0,m) ; item t in [0,n_items) ; chunk c in [0,n_chunks) ; document x in [0,n_doc).
// ----------------------------------------------------------------------------------------------------
struct TrcSub {          // one sub-chunk of an over-budget segment (big_chunkeroosky.py:267-542)
  int32_t group;         // parent merged segment
  int32_t kind;          // 0 combined-piece path (:280-329), 1 sentence path (:331-433), 2 clause (:437-542)
  int32_t r0, r1;        // inclusive piece range (kind 0) or sentence/clause record range (kind 1/2)
  int32_t start_src;     // piece whose ORIGINAL start object is the value, or -1 (computed float)
  int32_t end_src;       // piece whose ORIGINAL end object is the value, or -1
  int32_t tok;
  int32_t pad;
  double start, end;
  int64_t text_len;
};

struct TrcChunk {
  int32_t item0, item1;      // item range [item0, item1)
  int32_t doc;
  int32_t chunk_index;       // within the document
  int64_t token_count;
  double start_time, end_time;
  double position_percentage;
  int32_t pct_is_int_zero;   // the reference's `else 0` branch yields the int 0 (big_chunkeroosky.py:184)
  int32_t start_kind;        // 0: start_time is group `start_ref`'s original start ; 1: sub `start_ref`'s start
  int32_t start_ref;
  int32_t n_speakers;        // ids at spk_list[item0 .. item0+n_speakers)
  int32_t hdr_len;           // bytes of the context header (0 when add_context is false)
  int32_t err;               // 1: percentage outside the formatter's exact range
  int64_t text_len;
  int64_t twc_off;           // offset of text_with_context in the output stream
};

struct TrcDev {
  // ---- inputs
  const uint8_t* text; const int64_t* seg_off; const double* start; const double* end; const int32_t* spk;
  const int64_t* doc_off; const int32_t* doc_E;
  const uint8_t* spk_bytes; const int64_t* spk_off; const uint32_t* spk_tok;  // spk_tok[id] = count(" "+spk+":")
  int64_t n, n_doc; int32_t n_spk;
  int32_t merge, has_cap, preserve, add_context; double cap;
  // ---- preprocess
  int64_t* clen; int32_t* cchars;          // per raw segment
  int64_t* kscan;                          // exclusive scan of keep flags, n+1
  int64_t* cscan;                          // exclusive scan of clen, n+1
  int64_t n_kept;
  int32_t* k2r; int64_t* coff; uint8_t* C; int32_t* sdoc; int64_t* kdoc_off;
  uint32_t* ctok;                          // count(" " + cleaned text)
  int32_t* kspk; double* kdur;             // per piece: speaker id and end - start (null unless the capped merge runs)
  int32_t* mnxt; int32_t* gflag; int64_t* gscan; int64_t m; int32_t* gstart; int64_t* gdoc_off;
  int32_t* plen; int64_t* poff; uint8_t* M;
  int64_t* poffs;                          // piece layout WITH timestamps (== poff when preserve), for sub-chunk text
  uint8_t* gts;                            // m * 2 * 24: [len][chars] start_formatted / end_formatted
  int64_t* gtok;                           // token count of "[ts] speaker: text" per group
  int64_t* glen;                           // byte length of that formatted string
  // ---- chunk
  int32_t* nitems_g; int64_t* item_off;    // per group, m+1
  int64_t n_items;
  int32_t* it_group; int32_t* it_sub; int32_t* it_tok; int64_t* it_len; uint8_t* it_forced;
  int64_t* W; int64_t* L;                  // n_items+1 prefix sums of it_tok and it_len+2
  int64_t* idoc_off; int32_t* pnxt; int32_t* cflag; int64_t* cscanf; int64_t n_chunks; int32_t* cstart;
  int64_t* cdoc_off; TrcChunk* chunks; int32_t* spk_list; int64_t* twc_len; int64_t* twc_scan; uint8_t* twc; int64_t* it_dst;
  TrcSub* subs; int64_t n_subs; int64_t* sub_off; int32_t* pc_item;
  int64_t* pdst; uint8_t* pflag;           // per piece: destination of its text in TWC (-1: none), bit0 "[ts] " prefix, bit1 joining space
  int32_t* doc_err;                        // per document: 0 ok, 1+k special token k found, -1 unsupported
  int64_t* doc_err_seg;
};

// ---- P3: same-speaker merge, preprocessor.py:132-158 ------------------------------------------------
// mnxt[j] = first kept segment that is NOT in the group that starts at j: exact left-to-right fp64
// running sum from j, so that the group starts of a document are the orbit of its
// first segment under mnxt.
TRC_HD void trc_merge_next(const TrcDev& D, int64_t j) {
  int32_t x = D.sdoc[j];
  int64_t hi = D.kdoc_off[x + 1];
  int64_t k = j + 1;
  if (D.merge && D.kspk) {                   // the same scan on the per-piece copies (contiguous loads, no indirection)
    const int32_t* const ps = D.kspk; const double* const pd = D.kdur;
    const double cap = D.cap; const bool has_cap = D.has_cap != 0;
    const int32_t sp = ps[j];
    double cur = pd[j];
    int32_t kk = (int32_t)k; const int32_t hh = (int32_t)hi;      // (piece indices fit 31 bits: k2r, mnxt are int32)
    while (kk < hh) {
      if (ps[kk] != sp) break;
      const double t = cur + pd[kk];           // the running sum the reference would have after taking this piece
      if (has_cap && t > cap) break;
      cur = t;
      ++kk;
    }
    k = kk;
  } else if (D.merge) {
    int32_t r = D.k2r[j];
    int32_t sp = D.spk[r];
    double cur = D.end[r] - D.start[r];
    while (k < hi) {
      int32_t rk = D.k2r[k];
      if (D.spk[rk] != sp) break;
      double d = D.end[rk] - D.start[rk];
      if (D.has_cap && cur + d > D.cap) break;
      cur += d;
      ++k;
    }
  }
  D.mnxt[j] = (int32_t)k;
}

// Without a duration cap (max_segment_duration=None, preprocessor.py:140-141) or without merging, a group starts
// exactly where the speaker changes (or at every segment): no forward scan and no orbit, one flag per piece.  A
// forward scan per piece would be quadratic in the length of a single-speaker run.
TRC_HD void trc_group_flag_nocap(const TrcDev& D, int64_t j) {
  const int32_t x = D.sdoc[j];
  const bool first = j == D.kdoc_off[x];
  D.gflag[j] = (!D.merge || first || D.spk[D.k2r[j]] != D.spk[D.k2r[j - 1]]) ? 1 : 0;
}

// Orbit marking (v1): one thread walks one document.  flag[] must be zeroed.
TRC_HD void trc_orbit_walk(const int64_t* range_off, const int32_t* nxt, int32_t* flag, int64_t x) {
  int64_t a = range_off[x], b = range_off[x + 1];
  while (a < b) { flag[a] = 1; a = nxt[a]; }
}

// ---- P4: per piece byte length inside the merged text, preprocessor.py:190-200 ---------------------------
TRC_HD bool trc_group_combined(const TrcDev& D, int64_t g) { return D.gstart[g + 1] - D.gstart[g] > 1; }

TRC_HD void trc_piece_len(const TrcDev& D, int64_t j) {
  int64_t g = D.gscan[j + 1] - 1;            // inclusive scan of gflag, minus one
  int32_t a = D.gstart[g], b = D.gstart[g + 1];
  int64_t len = D.coff[j + 1] - D.coff[j];
  if (b - a > 1) {
    if (D.preserve) len += 3 + trc_ts_len(D.start[D.k2r[j]]);
    if (j + 1 < b) len += 1;
  }
  D.plen[j] = (int32_t)len;
}

// Copy one piece ("[ts] " + cleaned text + optional joining space) to dst; lanes cooperate.
TRC_HD void trc_emit_piece(const TrcDev& D, int64_t j, uint8_t* dst, bool with_ts, bool trailing_space, int lane, int nl) {
  int64_t o = 0;
  if (with_ts) {
    uint8_t buf[TRC_TS_MAX + 4];
    int tl = trc_fmt_ts(D.start[D.k2r[j]], buf + 1);
    buf[0] = '['; buf[tl + 1] = ']'; buf[tl + 2] = ' ';
    for (int k = lane; k < tl + 3; k += nl) dst[k] = buf[k];
    o = tl + 3;
  }
  const uint8_t* src = D.C + D.coff[j];
  int64_t len = D.coff[j + 1] - D.coff[j];
  for (int64_t k = lane; k < len; k += nl) dst[o + k] = src[k];
  if (trailing_space && lane == 0) dst[o + len] = ' ';
}

TRC_HD void trc_emit_merged(const TrcDev& D, int64_t j, int lane, int nl) {
  int64_t g = D.gscan[j + 1] - 1;
  int32_t a = D.gstart[g], b = D.gstart[g + 1];
  bool comb = b - a > 1;
  trc_emit_piece(D, j, D.M + D.poff[j], comb && D.preserve, comb && j + 1 < b, lane, nl);
}

// ---- P5/C0: per group token total and formatted length -------------------------------------------------
// formatted = "[" ts(start) "] " speaker ": " text        (big_chunkeroosky.py:244-265)
// tokens    = count("[ts]") + count(" speaker:") + count(" " + text), exact because the pre-tokeniser
//             always has match boundaries after "]" and after ":" (DESIGN.md §3).
TRC_HD void trc_group_stats(const TrcDev& D, const TrcTables& T, int64_t g, int lane, int nl, uint32_t mask = 0xffffffffu) {   // lanes share the pieces
  int32_t a = D.gstart[g], b = D.gstart[g + 1];
  int32_t ra = D.k2r[a], rb = D.k2r[b - 1];
  double st = D.start[ra];
  int64_t part = 0;
  if (b - a == 1) { if (lane == 0) part = D.ctok[a]; }
  else {
    for (int32_t j = a + lane; j < b; j += nl) {
      part += D.ctok[j];
      if (D.preserve) part += T.cnt_space_lbracket + trc_ts_tokens(T, D.start[D.k2r[j]]) + 1;
    }
  }
  const int64_t tok = 2 + trc_ts_tokens(T, st) + D.spk_tok[D.spk[ra]] + (nl > 1 ? trc_lanes_sum_w(part, nl, mask) : part);
  if (lane != 0) return;
  D.gtok[g] = tok;
  int64_t mlen = D.poff[b] - D.poff[a];
  int32_t s = D.spk[ra];
  D.glen[g] = 1 + trc_ts_len(st) + 2 + (D.spk_off[s + 1] - D.spk_off[s]) + 2 + mlen;
  uint8_t* o = D.gts + g * 48;
  o[0] = (uint8_t)trc_fmt_ts(st, o + 1);
  o[24] = (uint8_t)trc_fmt_ts(D.end[rb], o + 25);
}

// ---- C1: sub-chunks of an over-budget COMBINED segment, big_chunkeroosky.py:280-329 ----------------------
// piece tokens = count("[ts_k] text_k") = 1 + ts + 1 + count(" " + text_k).  Greedy: a new sub-chunk
// starts at piece k when cur + p_k > E and cur > 0.  WRITE=false only counts.
TRC_HD int32_t trc_piece_tokens(const TrcDev& D, const TrcTables& T, int32_t j) {
  return 2 + trc_ts_tokens(T, D.start[D.k2r[j]]) + (int32_t)D.ctok[j];
}

// The greedy is sequential, but the piece token counts are not: `tokens(j)` may hand out counts that the lanes of a
// warp computed 32 at a time (PieceTokWarp); every lane runs the same loop and `wr` picks the lane that writes.
struct PieceTokDirect {
  const TrcDev* D; const TrcTables* T;
  TRC_HD int64_t operator()(int32_t j) { return trc_piece_tokens(*D, *T, j); }
};
#ifndef TRC_HOSTSIM
struct PieceTokWarp {
  const TrcDev* D; const TrcTables* T; int32_t hi; int lane; int32_t win0; int32_t mine;
  __device__ int64_t operator()(int32_t j) {
    if (j < win0 || j >= win0 + 32) {
      win0 = j;
      const int32_t x = j + lane;
      if (x < hi) mine = trc_piece_tokens(*D, *T, x);
    }
    return (int64_t)__shfl_sync(0xffffffffu, mine, j - win0);
  }
};
#endif
template <bool WRITE, class Tokens>
TRC_HD int32_t trc_split_combined(const TrcDev& D, int64_t g, TrcSub* out, Tokens&& tokens, bool wr) {
  int32_t a = D.gstart[g], b = D.gstart[g + 1];
  int64_t E = D.doc_E[D.sdoc[a]];
  int32_t ns = 0;
  int64_t cur = 0;
  int32_t first = a;
  for (int32_t j = a; j < b; ++j) {
    int64_t p = tokens(j);
    if (cur + p > E && cur > 0) {
      if (WRITE && wr) {
        TrcSub& s = out[ns];
        s.group = (int32_t)g; s.kind = 0; s.r0 = first; s.r1 = j - 1; s.start_src = first; s.end_src = j - 1;
        s.tok = (int32_t)cur; s.pad = 0;
        s.start = D.start[D.k2r[first]]; s.end = D.end[D.k2r[j - 1]];
        s.text_len = D.poffs[j] - D.poffs[first] - 1;
      }
      ++ns; cur = 0; first = j;
    }
    cur += p;
  }
  if (cur > 0) {
    if (WRITE && wr) {
      TrcSub& s = out[ns];
      s.group = (int32_t)g; s.kind = 0; s.r0 = first; s.r1 = b - 1; s.start_src = first; s.end_src = b - 1;
      s.tok = (int32_t)cur; s.pad = 0;
      s.start = D.start[D.k2r[first]]; s.end = D.end[D.k2r[b - 1]];
      s.text_len = D.poffs[b] - D.poffs[first];
    }
    ++ns;
  }
  return ns;
}

// ---- C2: items ------------------------------------------------------------------------------------------
TRC_HD void trc_build_items(const TrcDev& D, int64_t g) {
  int64_t t0 = D.item_off[g];
  int32_t cnt = D.nitems_g[g];
  int32_t a = D.gstart[g];
  bool large = D.gtok[g] > (int64_t)D.doc_E[D.sdoc[a]];
  if (!large) {
    D.it_group[t0] = (int32_t)g; D.it_sub[t0] = -1; D.it_tok[t0] = (int32_t)D.gtok[g];
    D.it_len[t0] = D.glen[g]; D.it_forced[t0] = 0;
    return;
  }
  int64_t s0 = D.sub_off[g];
  for (int32_t k = 0; k < cnt; ++k) {
    const TrcSub& s = D.subs[s0 + k];
    D.it_group[t0 + k] = (int32_t)g; D.it_sub[t0 + k] = (int32_t)(s0 + k); D.it_tok[t0 + k] = s.tok;
    D.it_len[t0 + k] = s.text_len; D.it_forced[t0 + k] = (k == 0);
  }
}

// ---- C3: greedy packing, big_chunkeroosky.py:86-137 ------------------------------------
// pnxt[a] = first item t > a with W[t+1] - W[a] > E (binary search over the integer prefix sums), cut at
// the next forced break (first sub-item of an over-budget segment) and at the document end.
TRC_HD void trc_pack_next(const TrcDev& D, const int64_t* forced_pos, const int64_t* n_forced_ptr, int64_t a) {
  const int64_t n_forced = *n_forced_ptr;            // counted on the device: no host round trip for it
  int32_t x = D.sdoc[D.gstart[D.it_group[a]]];
  int64_t hi = D.idoc_off[x + 1];
  int64_t E = D.doc_E[x];
  int64_t base = D.W[a];
  // smallest t in (a, hi] such that t == hi or W[t+1] - base > E
  int64_t lo = a + 1, up = hi;
  while (lo < up) {
    int64_t mid = (lo + up) >> 1;
    if (D.W[mid + 1] - base > E) up = mid; else lo = mid + 1;
  }
  // next forced item after a
  int64_t fl = 0, fu = n_forced;
  while (fl < fu) { int64_t mid = (fl + fu) >> 1; if (forced_pos[mid] > a) fu = mid; else fl = mid + 1; }
  if (fl < n_forced && forced_pos[fl] < lo) lo = forced_pos[fl];
  D.pnxt[a] = (int32_t)lo;
}

// ---- C4: chunk records, big_chunkeroosky.py:147-232 ------------------------------------------------------
TRC_HD double trc_item_start(const TrcDev& D, int64_t t) {
  int32_t s = D.it_sub[t];
  return s >= 0 ? D.subs[s].start : D.start[D.k2r[D.gstart[D.it_group[t]]]];
}
TRC_HD double trc_item_end(const TrcDev& D, int64_t t) {
  int32_t s = D.it_sub[t];
  if (s >= 0) return D.subs[s].end;
  int32_t g = D.it_group[t];
  return D.end[D.k2r[D.gstart[g + 1] - 1]];
}
TRC_HD int32_t trc_item_speaker(const TrcDev& D, int64_t t) {
  int32_t s = D.it_sub[t];
  if (s >= 0 && D.subs[s].kind == 2) return 0;     // clause sub-chunks keep speaker "" (id 0), :485,521
  return D.spk[D.k2r[D.gstart[D.it_group[t]]]];
}

#define TRC_HDR_A "--- TRANSCRIPT CHUNK INFORMATION ---\nTime Range: "
#define TRC_HDR_B "\nSpeakers: "
#define TRC_HDR_C "\nPosition: Chunk "
#define TRC_HDR_D " (approximately "
#define TRC_HDR_E "% through the transcript)\n--- TRANSCRIPT CHUNK CONTENT ---"

template <bool WRITE>
TRC_HD int64_t trc_put(uint8_t* dst, int64_t o, const char* s) {
  int64_t k = 0;
  while (s[k]) { if (WRITE) dst[o + k] = (uint8_t)s[k]; ++k; }
  return o + k;
}

// Header text of chunk c; WRITE=false returns the length only.
template <bool WRITE>
TRC_HD int64_t trc_header(const TrcDev& D, const TrcChunk& c, uint8_t* dst) {
  uint8_t buf[40];
  int64_t o = trc_put<WRITE>(dst, 0, TRC_HDR_A);
  int n = trc_fmt_ts(c.start_time, buf);
  if (WRITE) for (int k = 0; k < n; ++k) dst[o + k] = buf[k];
  o += n;
  o = trc_put<WRITE>(dst, o, " - ");
  n = trc_fmt_ts(c.end_time, buf);
  if (WRITE) for (int k = 0; k < n; ++k) dst[o + k] = buf[k];
  o += n;
  o = trc_put<WRITE>(dst, o, TRC_HDR_B);
  for (int32_t k = 0; k < c.n_speakers; ++k) {
    if (k) o = trc_put<WRITE>(dst, o, ", ");
    int32_t s = D.spk_list[c.item0 + k];
    int64_t a = D.spk_off[s], b = D.spk_off[s + 1];
    if (WRITE) for (int64_t q = a; q < b; ++q) dst[o + (q - a)] = D.spk_bytes[q];
    o += b - a;
  }
  o = trc_put<WRITE>(dst, o, TRC_HDR_C);
  n = trc_fmt_uint((uint64_t)c.chunk_index + 1, buf);
  if (WRITE) for (int k = 0; k < n; ++k) dst[o + k] = buf[k];
  o += n;
  o = trc_put<WRITE>(dst, o, TRC_HDR_D);
  n = c.pct_is_int_zero ? trc_fmt_pct1(0.0, buf) : trc_fmt_pct1(c.position_percentage, buf);
  if (n < 0) n = 0;
  if (WRITE) for (int k = 0; k < n; ++k) dst[o + k] = buf[k];
  o += n;
  o = trc_put<WRITE>(dst, o, TRC_HDR_E);
  return o;
}

// The record of chunk c in three parts, so that the CUDA build can give the middle one (the speaker set) to a whole warp
// (k_chunk_record_w in trchunk.cu): everything that is one value per chunk / the sorted distinct speakers / the lengths.
TRC_HD void trc_chunk_record_head(const TrcDev& D, int64_t c) {
  TrcChunk& ch = D.chunks[c];
  int64_t a = D.cstart[c], b = D.cstart[c + 1];
  int32_t x = D.sdoc[D.gstart[D.it_group[a]]];
  ch.item0 = (int32_t)a; ch.item1 = (int32_t)b; ch.doc = x;
  ch.chunk_index = (int32_t)(c - D.cdoc_off[x]);
  ch.token_count = D.W[b] - D.W[a];
  // start_time: :74,:95 the (parent) segment's start, or :116 the sub-chunk's start when a sub-chunk
  // that is not the first of its segment opened the chunk
  int32_t s = D.it_sub[a];
  if (s >= 0 && !D.it_forced[a]) { ch.start_kind = 1; ch.start_ref = s; ch.start_time = D.subs[s].start; }
  else { ch.start_kind = 0; ch.start_ref = D.it_group[a]; ch.start_time = D.start[D.k2r[D.gstart[D.it_group[a]]]]; }
  ch.end_time = trc_item_end(D, b - 1);
  double first = trc_item_start(D, a), last = ch.end_time;
  double t0 = trc_item_start(D, D.cstart[D.cdoc_off[x]]);
  ch.err = 0;
  if (last > t0) {
    double num = first - t0, den = last - t0;
    double q = num / den;
    ch.position_percentage = q * 100.0;
    ch.pct_is_int_zero = 0;
    double av = ch.position_percentage < 0 ? -ch.position_percentage : ch.position_percentage;
    if (av >= 576460752303423488.0 && av <= 1.7976931348623157e308) ch.err = 1;
  } else { ch.position_percentage = 0.0; ch.pct_is_int_zero = 1; }
}
TRC_HD void trc_chunk_record_tail(const TrcDev& D, int64_t c, int32_t ns) {
  TrcChunk& ch = D.chunks[c];
  const int64_t a = ch.item0, b = ch.item1;
  ch.n_speakers = ns;
  ch.text_len = (D.L[b] - D.L[a]) - 2;
  ch.hdr_len = D.add_context ? (int32_t)trc_header<false>(D, ch, nullptr) : 0;
  D.twc_len[c] = ch.text_len + (D.add_context ? ch.hdr_len + 2 : 0);
}
TRC_HD void trc_chunk_record(const TrcDev& D, int64_t c) {
  trc_chunk_record_head(D, c);
  const int64_t a = D.chunks[c].item0, b = D.chunks[c].item1;
  // sorted distinct speaker ids (ids are assigned in code-point order of the strings, so sorted(set) == ascending ids)
  int32_t ns = 0;
  int32_t* lst = D.spk_list + a;
  for (int64_t t = a; t < b; ++t) {
    int32_t v = trc_item_speaker(D, t);
    int32_t k = 0;
    while (k < ns && lst[k] < v) ++k;
    if (k < ns && lst[k] == v) continue;
    for (int32_t q = ns; q > k; --q) lst[q] = lst[q - 1];
    lst[k] = v; ++ns;
  }
  trc_chunk_record_tail(D, c, ns);
}

TRC_HD void trc_chunk_header_write(const TrcDev& D, int64_t c) {
  TrcChunk& ch = D.chunks[c];
  ch.twc_off = D.twc_scan[c];
  int64_t body = ch.twc_off;
  if (D.add_context) {
    uint8_t* dst = D.twc + ch.twc_off;
    int64_t o = trc_header<true>(D, ch, dst);
    dst[o] = '\n'; dst[o + 1] = '\n';
    body += o + 2;
  }
  // destination of every item text of this chunk
  for (int64_t t = ch.item0; t < ch.item1; ++t) D.it_dst[t] = body + (D.L[t] - D.L[ch.item0]);
}

// ---- C5: item text emission (K6) -----------------------------------------------------------------------
// One call per piece j: writes the piece into its item's text, and the first piece of an item also
// writes the item prefix ("[ts] speaker: " for whole segments) and the "\n\n" that precedes the item.
TRC_HD void trc_emit_chunk_piece(const TrcDev& D, int64_t j, int lane, int nl) {
  int64_t t = D.pc_item[j];
  if (t < 0) return;                      // piece belongs to a sentence-path item (emitted elsewhere)
  int32_t g = D.it_group[t];
  int32_t s = D.it_sub[t];
  int32_t ga = D.gstart[g], gb = D.gstart[g + 1];
  int32_t p0 = s >= 0 ? D.subs[s].r0 : ga;
  int32_t p1 = s >= 0 ? D.subs[s].r1 : gb - 1;
  bool comb = gb - ga > 1;
  uint8_t* base = D.twc + D.it_dst[t];
  int64_t pre = 0;
  if (s < 0) {
    int32_t r = D.k2r[ga];
    int32_t sp = D.spk[r];
    int64_t sl = D.spk_off[sp + 1] - D.spk_off[sp];
    pre = 1 + trc_ts_len(D.start[r]) + 2 + sl + 2;
    if (j == p0) {
      uint8_t buf[TRC_TS_MAX + 4];
      int tl = trc_fmt_ts(D.start[r], buf + 1);
      buf[0] = '['; buf[tl + 1] = ']'; buf[tl + 2] = ' ';
      for (int k = lane; k < tl + 3; k += nl) base[k] = buf[k];
      for (int64_t k = lane; k < sl; k += nl) base[tl + 3 + k] = D.spk_bytes[D.spk_off[sp] + k];
      if (lane == 0) { base[tl + 3 + sl] = ':'; base[tl + 4 + sl] = ' '; }
    }
  }
  if (j == p0 && lane == 0 && t > D.chunks[D.cscanf[t + 1] - 1].item0) { base[-2] = '\n'; base[-1] = '\n'; }
  const int64_t* po = s >= 0 ? D.poffs : D.poff;
  trc_emit_piece(D, j, base + pre + (po[j] - po[p0]), comb && (D.preserve || s >= 0), comb && j < p1, lane, nl);
}


// ---- C5 (v2): item layout once per item, then a flat copy per piece ---------------------------------------
// One call per item t: writes the "\n\n" that precedes the item and, for whole segments, the
// "[ts] speaker: " prefix; records for every piece of the item where its text goes.
TRC_HD void trc_item_layout(const TrcDev& D, int64_t t) {
  int32_t g = D.it_group[t];
  int32_t s = D.it_sub[t];
  if (s >= 0 && D.subs[s].kind != 0) return;           // sentence / clause items are laid out by k_emit_sent_recs
  int32_t ga = D.gstart[g], gb = D.gstart[g + 1];
  int32_t p0 = s >= 0 ? D.subs[s].r0 : ga;
  int32_t p1 = s >= 0 ? D.subs[s].r1 : gb - 1;
  bool comb = gb - ga > 1;
  uint8_t* base = D.twc + D.it_dst[t];
  int64_t pre = 0;
  if (s < 0) {
    int32_t r = D.k2r[ga];
    int32_t sp = D.spk[r];
    int64_t sl = D.spk_off[sp + 1] - D.spk_off[sp];
    uint8_t buf[TRC_TS_MAX + 4];
    int tl = trc_fmt_ts(D.start[r], buf + 1);
    buf[0] = '['; buf[tl + 1] = ']'; buf[tl + 2] = ' ';
    for (int k = 0; k < tl + 3; ++k) base[k] = buf[k];
    for (int64_t k = 0; k < sl; ++k) base[tl + 3 + k] = D.spk_bytes[D.spk_off[sp] + k];
    base[tl + 3 + sl] = ':'; base[tl + 4 + sl] = ' ';
    pre = tl + 5 + sl;
  }
  if (t > D.chunks[D.cscanf[t + 1] - 1].item0) { base[-2] = '\n'; base[-1] = '\n'; }
  const int64_t* po = s >= 0 ? D.poffs : D.poff;
  bool with_ts = comb && (D.preserve || s >= 0);
  for (int32_t j = p0; j <= p1; ++j) {
    D.pdst[j] = D.it_dst[t] + pre + (po[j] - po[p0]);
    D.pflag[j] = (uint8_t)((with_ts ? 1 : 0) | ((comb && j < p1) ? 2 : 0));
  }
}

// One call per (piece, lane of an 8-lane group): "[ts] " (lane 0) + the cleaned text + the joining space.
TRC_HD void trc_emit_piece_flat(const TrcDev& D, int64_t j, int lane, int nl) {
  int64_t d = D.pdst[j];
  if (d < 0) return;
  uint8_t f = D.pflag[j];
  uint8_t* dst = D.twc + d;
  int64_t o = 0;
  if (f & 1) {
    double st = D.start[D.k2r[j]];
    o = trc_ts_len(st) + 3;
    if (lane == 0) {
      uint8_t buf[TRC_TS_MAX + 4];
      int tl = trc_fmt_ts(st, buf + 1);
      buf[0] = '['; buf[tl + 1] = ']'; buf[tl + 2] = ' ';
      for (int k = 0; k < tl + 3; ++k) dst[k] = buf[k];
    }
  }
  int64_t c0 = D.coff[j];
  int64_t len = D.coff[j + 1] - c0;
  const uint8_t* src = D.C + c0;
  for (int64_t k = lane; k < len; k += nl) dst[o + k] = src[k];
  if ((f & 2) && lane == nl - 1) dst[o + len] = ' ';
}
