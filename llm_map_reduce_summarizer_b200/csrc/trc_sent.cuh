// Sentence path of over-budget RAW segments (big_chunkeroosky.py:331-433, 437-542): records.
#pragma once
#include "trc_common.cuh"

struct TrcSent {          // one sentence or clause of an over-budget raw segment
  int32_t group;
  int32_t kind;           // 1 sentence, 2 clause
  int64_t off, len;       // byte range inside the cleaned text buffer
  int32_t chars;          // code points
  int32_t tok;            // tokens of "[ts] text"
  double start, end;
};

struct TrcSentState {
  TrcSent* recs;
  int64_t n_recs;
};
