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

TRC_HD int32_t pk_count_chars(const uint8_t* s, int64_t a, int64_t b) {
  int32_t c = 0;
  for (int64_t i = a; i < b; ++i) c += (s[i] & 0xC0) != 0x80;
  return c;
}
TRC_HD bool pk_clause_punct(uint8_t c) { return c == ',' || c == '.' || c == ';' || c == ':' || c == '?' || c == '!'; }

// Clause splitter for one over-budget sentence s[a,b) (:455-472).  emit(k, ca, cb) receives the raw
// clause ranges in order (before strip); returns the count.
template <class Emit>
TRC_HD int32_t pk_clauses(const uint8_t* s, int64_t a, int64_t b, Emit&& emit) {
  int32_t n = 0;
  int64_t i = a;
  while (i < b && pk_clause_punct(s[i])) ++i;       // a leading punctuation run matches nothing
  int64_t cs = i;
  while (i < b) {
    while (i < b && !pk_clause_punct(s[i])) ++i;
    if (i >= b) break;                              // trailing text without punctuation is dropped (:455-456)
    while (i < b && pk_clause_punct(s[i])) ++i;
    emit(n, cs, i); ++n;
    cs = i;
  }
  if (n) return n;
  // no clause boundary: groups of 20 whitespace-separated words (:459-472)
  i = a;
  while (i < b) {
    while (i < b && s[i] == ' ') ++i;
    if (i >= b) break;
    int64_t gs = i, ge = i; int w = 0;
    while (i < b && w < 20) {
      while (i < b && s[i] != ' ') ++i;
      ge = i; ++w;
      while (i < b && s[i] == ' ') ++i;
    }
    emit(n, gs, ge); ++n;
  }
  return n;
}
