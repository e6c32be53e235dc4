// Sentence path of over-budget UN-MERGED segments: big_chunkeroosky.py:331-433 (sentences) and :437-542
// (clauses), including the untrained Punkt tokenizer behind `self.sentence_tokenizer.tokenize(text)`
// (:44, :334; nltk==3.8.1 punkt.py restated for empty parameters - see oracle/punkt.py, parity
// unpinned for that one function because nltk is not installable offline).
//
// Domain: the text is a CLEANED segment (output of trc_clean_segment), so its only whitespace is single
// U+0020 between words and it has no leading/trailing whitespace.  All characters the Punkt regexes
// name are ASCII, so the classifier works on bytes; multi-byte UTF-8 sequences are ordinary word bytes.
#pragma once
#include "trc_common.cuh"
#include "trc_format.cuh"
#include "trc_swar.cuh"

struct TrcSent {          // one sentence (kind 1) or clause (kind 2) of an over-budget raw segment
  int32_t group;
  int32_t kind;
  int64_t off, len;       // byte range inside the cleaned text buffer (after strip); len 0 = skipped
  int32_t chars;          // code points
  int32_t tok;            // tokens of "[ts] text"
  double start, end;      // :364-365 / :504-505
  int32_t sub;            // sub-chunk that holds this record, or -1 (empty / replaced by its clauses)
  int32_t last_in_sub;
  int64_t off_in_sub;     // byte offset of "[ts] text" inside the sub-chunk text
  int32_t parent;         // clauses: index of the sentence record
  int32_t ncl;            // sentences: number of clause records (0 unless over budget); cl0 = first
  int64_t cl0;
};

struct TrcSentState {
  TrcSent* recs;          // sentences [0, n_sent) followed by clauses [n_sent, n_recs)
  int64_t n_sent, n_recs;
  const int32_t* rl_groups;   // over-budget raw groups
  int64_t n_rl;
  const int64_t* rl_boff;     // byte prefix of their texts, n_rl+1
  const int64_t* sent_off;    // first sentence record of each, n_rl+1
};

// Per-word token counts of the oversized segments (trc_giant.cuh), kept from stage P: every space of a cleaned
// text starts a word and words are tokenised independently, so the tokens of " " + sentence are a
// difference of the running sum over its words whenever the sentence begins and ends at word edges.
struct TrcWordSums {
  const int32_t* of_seg;      // per raw segment: giant index or -1; nullptr: none
  const int64_t* toff;        // cleaned offset of every giant inside the giants' text buffer
  const int64_t* wscan;       // words that start before byte x of that buffer
  const int64_t* wpre;        // tokens of the words before word w
  const uint8_t* wtok;        // cut mode instead of the four above: per byte of the cleaned text, the token count of the word that
                              // starts there (0: no word starts there, 255: not known); nullptr: none
};

TRC_HD bool pk_nonword(uint8_t c) {
  return c == ')' || c == '"' || c == ';' || c == '}' || c == ']' || c == '*' || c == ':' || c == '@' || c == '\'' ||
         c == '(' || c == '{' || c == '[' || c == '?' || c == '!';
}
TRC_HD bool pk_wordstart(uint8_t c) {
  return !(c == '(' || c == '"' || c == '`' || c == '{' || c == '[' || c == ':' || c == ';' || c == '&' || c == '#' ||
           c == '*' || c == '@' || c == ')' || c == '}' || c == ']' || c == '-' || c == ',');
}
TRC_HD bool pk_sentend(uint8_t c) { return c == '.' || c == '?' || c == '!'; }
// _re_multi_char_punct at i: "--+" or "..+"  (the ". . ." form needs two whitespace and never fits a context)
TRC_HD int64_t pk_multi(const uint8_t* s, int64_t i, int64_t n) {
  if (i + 1 < n && ((s[i] == '-' && s[i + 1] == '-') || (s[i] == '.' && s[i + 1] == '.'))) {
    int64_t j = i + 2;
    while (j < n && s[j] == s[i]) ++j;
    return j - i;
  }
  return 0;
}
// period_context_re: [.?!] followed by a NonWord char, or by whitespace and another token
TRC_HD bool pk_candidate(const uint8_t* s, int64_t i, int64_t n) {
  if (!pk_sentend(s[i]) || i + 1 >= n) return false;
  uint8_t c = s[i + 1];
  return pk_nonword(c) || (c == ' ' && i + 2 < n);
}
// next token of word_tokenize at i (s[i] != ' '): returns its end
TRC_HD int64_t pk_token_end(const uint8_t* s, int64_t i, int64_t n) {
  int64_t m = pk_multi(s, i, n);
  if (m) return i + m;
  if (pk_wordstart(s[i])) {
    int64_t j = i + 1;
    for (; j < n; ++j) {
      uint8_t c = s[j];
      if (c == ' ' || pk_nonword(c) || pk_multi(s, j, n)) break;
      if (c == ',' && (j + 1 == n || s[j + 1] == ' ' || pk_nonword(s[j + 1]) || pk_multi(s, j + 1, n))) break;
    }
    return j;
  }
  return i + 1;
}

struct PkTok { int64_t a, b; bool brk, pfinal, initial, number; };

TRC_HD void pk_classify(const TrcTables& T, const uint8_t* s, PkTok& t) {
  int64_t L = t.b - t.a;
  const uint8_t* p = s + t.a;
  t.pfinal = p[L - 1] == '.';
  t.brk = false; t.initial = false; t.number = false;
  if (L == 1 && pk_sentend(p[0])) { t.brk = true; }
  else {
    bool alldots = true;
    for (int64_t k = 0; k < L; ++k) if (p[k] != '.') { alldots = false; break; }
    if (alldots && L >= 2) { /* ellipsis */ }
    else if (t.pfinal && !(L >= 2 && p[L - 2] == '.')) t.brk = true;
  }
  if (!t.pfinal) return;
  int l0;
  uint32_t c0 = trc_decode(p, 0, L, &l0);
  uint32_t pr = trc_prop(T, c0);
  t.initial = (pr & TRC_P_WORD) && !(pr & TRC_P_DIGIT) && L == l0 + 1;
  // ^-?[.,]?\d[\d,.-]*\.?$
  int64_t k = 0;
  if (k < L && p[k] == '-') ++k;
  if (k < L && (p[k] == '.' || p[k] == ',')) ++k;
  bool ok = false;
  if (k < L) {
    int l; uint32_t c = trc_decode(p, k, L, &l);
    if (trc_prop(T, c) & TRC_P_DIGIT) {
      k += l; ok = true;
      while (k < L) {
        uint32_t c2 = trc_decode(p, k, L, &l);
        if (!((trc_prop(T, c2) & TRC_P_DIGIT) || c2 == ',' || c2 == '.' || c2 == '-')) { ok = false; break; }
        k += l;
      }
    }
  }
  t.number = ok;
}

// text_contains_sentbreak(context) for context = s[a, b)
TRC_HD bool pk_context_has_break(const TrcTables& T, const uint8_t* s, int64_t a, int64_t b) {
  PkTok prev; bool have = false;
  int64_t i = a;
  while (i < b) {
    if (s[i] == ' ') { ++i; continue; }
    PkTok t; t.a = i; t.b = pk_token_end(s, i, b);
    pk_classify(T, s, t);
    if (have) {
      if (prev.pfinal && (prev.initial || prev.number)) {           // second pass
        int64_t L2 = t.b - t.a;
        uint8_t c = s[t.a];
        bool punct = L2 == 1 && (c == ';' || c == ':' || c == ',' || c == '.' || c == '!' || c == '?');
        int l; uint32_t c0 = trc_decode(s + t.a, 0, L2, &l);
        uint32_t pr = trc_prop(T, c0);
        if (punct || (pr & TRC_P_LOWER)) prev.brk = false;
        else if (prev.initial && (pr & TRC_P_UPPER)) prev.brk = false;
      }
      if (prev.brk) return true;
    }
    prev = t; have = true;
    i = t.b;
  }
  return false;   // the last token is ignored
}

// Is byte i of text s[0,n) a sentence break (the yielded match of its overlap chain says "break")?
TRC_HD bool pk_is_break(const TrcTables& T, const uint8_t* s, int64_t n, int64_t i) {
  if (!pk_candidate(s, i, n)) return false;
  // _match_potential_end_contexts: candidates with no whitespace between them form a chain that shares
  // one "previous word"; only the LAST candidate of a chain is yielded - plus the first one when it sits
  // exactly at the word start (its own previous-word slice is empty, so nothing overlaps it).
  bool last = true;
  for (int64_t j = i + 1; j < n && s[j] != ' '; ++j) if (pk_candidate(s, j, n)) { last = false; break; }
  int64_t first = i;
  for (int64_t j = i - 1; j >= 0 && s[j] != ' '; --j) if (pk_candidate(s, j, n)) first = j;
  int64_t ws = first;
  while (ws > 0 && s[ws - 1] != ' ') --ws;        // word start: after the last space before the chain
  if (!last) {
    if (!(first == i && ws == i)) return false;
    return pk_context_has_break(T, s, i, i + 2);  // context = "" + s[i] + the NonWord char that follows
  }
  int64_t ce;
  if (pk_nonword(s[i + 1])) ce = i + 2;
  else { ce = i + 2; while (ce < n && s[ce] != ' ') ++ce; }
  return pk_context_has_break(T, s, ws, ce);
}

// realign: ["')\]}]+?(?:\s+|(?=--)|$) matched at the start of slice [a,b): k = quote chars, e = m.end()
TRC_HD void pk_realign(const uint8_t* s, int64_t a, int64_t b, int64_t* k_out, int64_t* e_out) {
  *k_out = 0; *e_out = 0;
  for (int64_t k = 1; a + k <= b; ++k) {
    uint8_t c = s[a + k - 1];
    if (!(c == '"' || c == '\'' || c == ')' || c == ']' || c == '}')) return;
    int64_t p = a + k;
    if (p < b && s[p] == ' ') { int64_t q = p; while (q < b && s[q] == ' ') ++q; *k_out = k; *e_out = q - a; return; }
    if (p + 1 < b && s[p] == '-' && s[p + 1] == '-') { *k_out = k; *e_out = k; return; }
    if (p == b) { *k_out = k; *e_out = k; return; }
  }
}

TRC_HD bool pk_clause_punct(uint8_t c) { return c == ',' || c == '.' || c == ';' || c == ':' || c == '?' || c == '!'; }

// ---- eight bytes at a time -----------------------------------------------------------------------------------
// An over-budget sentence can be tens of KB (config 4); one thread walks it, so the walk reads aligned 64-bit
// words and works on 8-bit masks: clause punctuation, space, UTF-8 continuation bytes.
struct PkWord { uint32_t p, sp, cont; };
TRC_HD uint64_t pk_eq8(uint64_t x, uint8_t c) {                // 0x80 in every byte of x that equals c (exact)
  const uint64_t t = x ^ (SW_L * c);
  return ~(((t & ~SW_H) + ~SW_H) | t | ~SW_H);
}
// word W (bytes 8W .. 8W+7) of the 8-byte aligned buffer sb, restricted to [a, b); bytes outside count as spaces
TRC_HD PkWord pk_word(const uint8_t* sb, int64_t W, int64_t a, int64_t b) {
  uint64_t x;
#if defined(__CUDA_ARCH__)
  x = *reinterpret_cast<const unsigned long long*>(sb + 8 * W);
#else
  memcpy(&x, sb + 8 * W, 8);
#endif
  const int64_t k0 = 8 * W;
  uint32_t valid = 0xffu;
  if (k0 < a) valid &= (a - k0 >= 8) ? 0u : (0xffu << (a - k0));
  if (k0 + 8 > b) valid &= (b <= k0) ? 0u : (0xffu >> (k0 + 8 - b));
  valid &= 0xffu;
  PkWord o;
  o.p = sw_pack(pk_eq8(x, ',') | pk_eq8(x, '.') | pk_eq8(x, ';') | pk_eq8(x, ':') | pk_eq8(x, '?') | pk_eq8(x, '!')) & valid;
  o.sp = (sw_pack(pk_eq8(x, ' ')) | ~valid) & 0xffu;
  o.cont = sw_pack((x & SW_H) & ~((x << 1) & SW_H)) & valid;
  return o;
}
struct PkBuf { const uint8_t* sb; int64_t mis; };              // aligned view of s: position k of s is byte k + mis of sb
TRC_HD PkBuf pk_buf(const uint8_t* s) { PkBuf o; o.mis = (int64_t)((uintptr_t)s & 7); o.sb = s - o.mis; return o; }

TRC_HD int32_t pk_count_chars(const uint8_t* s, int64_t a, int64_t b) {
  if (b <= a) return 0;
  if (b - a < 16) { int32_t c = 0; for (int64_t i = a; i < b; ++i) c += (s[i] & 0xC0) != 0x80; return c; }
  // continuation bytes (10xxxxxx) of the aligned 8-byte words that cover [a, b); the first and the last word are masked
  const PkBuf B = pk_buf(s);
  const int64_t a2 = a + B.mis, b2 = b + B.mis, W0 = a2 >> 3, W1 = (b2 - 1) >> 3;
  auto cm = [&](int64_t W) -> uint64_t {
    uint64_t x;
#if defined(__CUDA_ARCH__)
    x = *reinterpret_cast<const unsigned long long*>(B.sb + 8 * W);
#else
    memcpy(&x, B.sb + 8 * W, 8);
#endif
    return (x & SW_H) & ~((x << 1) & SW_H);
  };
  auto pc = [](uint64_t m) -> int32_t {
#if defined(__CUDA_ARCH__)
    return (int32_t)__popcll(m);
#else
    return (int32_t)__builtin_popcountll(m);
#endif
  };
  const unsigned e = (unsigned)(b2 - 8 * W1);                            // bytes of the last word that belong to [a, b): 1..8
  const uint64_t head = ~0ULL << (8 * (unsigned)(a2 & 7)), tail = e < 8u ? (1ULL << (8 * e)) - 1ULL : ~0ULL;
  int32_t cont;
  if (W0 == W1) cont = pc(cm(W0) & head & tail);
  else {
    cont = pc(cm(W0) & head) + pc(cm(W1) & tail);
    for (int64_t W = W0 + 1; W < W1; ++W) cont += pc(cm(W));
  }
  return (int32_t)(b - a) - cont;
}

// Clause splitter for one over-budget sentence s[a,b) (:455-472): the matches of ([^,.;:?!]+[,.;:?!]+), i.e. one
// clause per punctuation run that has something in front of it (a leading run matches nothing, text after the last
// run is dropped, :455-456); when there is none, groups of 20 space-separated words (:459-472).
// emit(k, ca, cb) receives the raw clause ranges in order (before strip); returns the count.  COUNT only counts.
template <bool COUNT, class Emit>
TRC_HD int32_t pk_clauses_t(const uint8_t* s, int64_t a, int64_t b, Emit&& emit) {
  if (b <= a) return 0;
  const PkBuf B = pk_buf(s);
  const int64_t a2 = a + B.mis, b2 = b + B.mis, W0 = a2 >> 3, W1 = (b2 - 1) >> 3;
  const bool lead = pk_clause_punct(s[a]);
  int32_t ends = 0, words = 0;
  {
    PkWord cur = pk_word(B.sb, W0, a2, b2);
    uint32_t prev_sp = 1u;
    for (int64_t W = W0; W <= W1; ++W) {
      PkWord nx; nx.p = 0; nx.sp = 0xffu; nx.cont = 0;
      if (W < W1) nx = pk_word(B.sb, W + 1, a2, b2);
      ends += (int32_t)trc_popc32(cur.p & ~((cur.p >> 1) | ((nx.p & 1u) << 7)));
      words += (int32_t)trc_popc32(~cur.sp & 0xffu & ((cur.sp << 1) | prev_sp));
      prev_sp = (cur.sp >> 7) & 1u;
      cur = nx;
    }
  }
  const int32_t ncl = ends - (lead ? 1 : 0);
  const int32_t n = ncl > 0 ? ncl : (words + 19) / 20;
  if (COUNT) return n;
  int32_t k = 0;
  PkWord cur = pk_word(B.sb, W0, a2, b2);
  if (ncl > 0) {
    int64_t cs = a;
    bool skip = lead;                                    // the leading run's end only moves the start
    for (int64_t W = W0; W <= W1; ++W) {
      PkWord nx; nx.p = 0; nx.sp = 0xffu; nx.cont = 0;
      if (W < W1) nx = pk_word(B.sb, W + 1, a2, b2);
      uint32_t e = cur.p & ~((cur.p >> 1) | ((nx.p & 1u) << 7));
      while (e) {
        const int bit = (int)trc_ctz32(e);
        e &= e - 1;
        const int64_t ce = 8 * W + bit + 1 - B.mis;
        if (skip) skip = false; else { emit(k, cs, ce); ++k; }
        cs = ce;
      }
      cur = nx;
    }
    return k;
  }
  uint32_t prev_sp = 1u;
  int64_t gs = a; int32_t w = 0;                         // w: words seen so far
  for (int64_t W = W0; W <= W1; ++W) {
    PkWord nx; nx.p = 0; nx.sp = 0xffu; nx.cont = 0;
    if (W < W1) nx = pk_word(B.sb, W + 1, a2, b2);
    const uint32_t ws = ~cur.sp & 0xffu & ((cur.sp << 1) | prev_sp);                // first byte of a word
    const uint32_t we = ~cur.sp & 0xffu & ((cur.sp >> 1) | ((nx.sp & 1u) << 7));    // last byte of a word
    uint32_t ev = ws | we;
    while (ev) {
      const int bit = (int)trc_ctz32(ev);
      ev &= ev - 1;
      const int64_t pos = 8 * W + bit - B.mis;
      if ((ws >> bit) & 1u) { if (w % 20 == 0) gs = pos; ++w; }
      if (((we >> bit) & 1u) && (w % 20 == 0 || w == words)) { emit(k, gs, pos + 1); ++k; }
    }
    prev_sp = (cur.sp >> 7) & 1u;
    cur = nx;
  }
  return k;
}
template <class Emit>
TRC_HD int32_t pk_clauses(const uint8_t* s, int64_t a, int64_t b, Emit&& emit) { return pk_clauses_t<false>(s, a, b, emit); }
TRC_HD int32_t pk_clause_count(const uint8_t* s, int64_t a, int64_t b) { return pk_clauses_t<true>(s, a, b, [](int32_t, int64_t, int64_t) {}); }
