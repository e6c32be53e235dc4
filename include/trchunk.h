/* trchunk - C ABI of the B200 (sm_100a) preprocess+chunk library  (libtrchunk.so)
 *
 * The reference (consilience-dev/llm-map-reduce-summarizer) is pure Python and has no FFI; its boundary
 * for this path is the Python import surface used at /root/reference/main.py:27-28,139-149.  The entry
 * points below are what a ctypes binding of that surface needs: packed transcripts in, packed
 * processed-segments / chunks out.  The drop-in modules that wrap them with the reference's exact
 * Python signatures live in llm_map_reduce_summarizer_b200/{preprocessor,big_chunkeroosky}.py and the
 * binding a maintainer would add to the reference is shown in INTEGRATION.md.
 *
 * Conventions: plain pointers and sizes only; every function returns 0 or a negative TRC_E* code and
 * leaves a message for trc_last_error(); a context is bound to one CUDA device and must not be used
 * from two threads at once; `stream` is a cudaStream_t passed as void* (NULL = the context's own).
 * Input pointers are HOST pointers when `memory == TRC_MEM_HOST` (the library copies them to the
 * device inside the call) or DEVICE pointers when `memory == TRC_MEM_DEVICE` (used in place).
 */
#ifndef TRCHUNK_H
#define TRCHUNK_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define TRC_OK 0
#define TRC_ECUDA (-1)      /* CUDA runtime error, no device, or wrong architecture            */
#define TRC_EINVAL (-2)     /* bad argument / call order                                        */
#define TRC_ENOVOCAB (-3)   /* tables or vocabulary not loaded                                   */
#define TRC_EUNSUPPORTED (-4) /* input outside the supported domain (reported, never approximated) */

#define TRC_MEM_HOST 0
#define TRC_MEM_DEVICE 1

typedef struct trc_ctx trc_ctx;

/* Life cycle.  trc_create fails (TRC_ECUDA) when no sm_100 device is present: there is no CPU path. */
int trc_create(int device, trc_ctx** out);
void trc_destroy(trc_ctx* ctx);
const char* trc_last_error(trc_ctx* ctx);

/* Unicode property table, one byte per code point (0x110000 entries, layout in
 * tools/gen_unicode_tables.py).  Replaces the tables CPython's str.split()/re and the Rust regex crate
 * carry internally (preprocessor.py:38,80,84; tiktoken pattern classes). */
int trc_load_tables(trc_ctx* ctx, const uint8_t* props);

/* Rank table of the BPE vocabulary + special-token literals.  Replaces
 * tiktoken.get_encoding(tokenizer_name) at big_chunkeroosky.py:43 (rank == token id). */
int trc_load_vocab(trc_ctx* ctx, const uint8_t* token_bytes, const int64_t* token_off, const int32_t* ranks,
                   int32_t n_tokens, const uint8_t* special_bytes, const int64_t* special_off, int32_t n_special);

/* A batch of transcripts ("documents").  Segment i owns text[seg_off[i] .. seg_off[i+1]) (UTF-8);
 * document x owns segments [doc_off[x], doc_off[x+1]).  speaker[i] indexes the speaker string table,
 * whose ids MUST be assigned in code-point order of the strings and MUST contain "" as id 0
 * (big_chunkeroosky.py:166 sorts speakers; :485 uses ""). */
typedef struct {
  const uint8_t* text;
  const int64_t* seg_off;      /* n_seg + 1 */
  const double* start;         /* n_seg     */
  const double* end;           /* n_seg     */
  const int32_t* speaker;      /* n_seg     */
  int64_t n_seg;
  const int64_t* doc_off;      /* n_doc + 1 */
  int64_t n_doc;
  const uint8_t* spk_bytes;
  const int64_t* spk_off;      /* n_spk + 1 */
  int32_t n_spk;
  int32_t memory;              /* TRC_MEM_HOST / TRC_MEM_DEVICE */
  int32_t merge_same_speaker;  /* preprocessor.py:17 */
  int32_t has_max_duration;    /* 0 == max_segment_duration is None, preprocessor.py:19,141 */
  double max_duration;
  int32_t preserve_timestamps; /* preprocessor.py:20 */
  int32_t want_merged_text;    /* 1: also materialise the merged text buffer (TRC_A_MERGED) */
  /* --- already processed input: chunk_transcript called on a list that did not come from this library
   * (big_chunkeroosky.py:46 accepts any list of processed-segment dicts).  With already_processed = 1 the
   * "segments" above are the PIECES of the processed segments (segment_timestamps entries of a combined
   * segment, else the segment itself), their text is used verbatim (no clean_text) and the grouping is
   * given instead of computed.  The segments' own start/end/text are verified on the device against what
   * the pieces imply (preprocessor.py:186-213); a mismatch marks the document unsupported (-1).
   * With group_off == NULL the pieces are merged as usual (combine_same_speaker_segments,
   * preprocessor.py:109-165, on already cleaned segments) and nothing is verified. */
  int32_t already_processed;
  int32_t reserved_;
  const int64_t* group_off;    /* n_groups + 1 piece ranges */
  int64_t n_groups;
  const double* group_start;   /* n_groups */
  const double* group_end;     /* n_groups */
  const uint8_t* full_text;    /* the segments' own text, back to back */
  const int64_t* full_off;     /* n_groups + 1 */
  /* --- optional compact form of seg_off for callers that upload over PCIe: n_seg byte lengths (each < 2^31) and
   * the total text size; seg_off may then be NULL and is rebuilt on the device by a prefix sum. */
  const int32_t* seg_len32;    /* n_seg, or NULL */
  int64_t text_bytes;          /* bytes of `text`; read only when seg_len32 is given */
} trc_batch;

/* Stage P: preprocessor.preprocess_transcript (preprocessor.py:15-67) for every document of the batch:
 * blank filter, clean_text, timestamp formatting, same-speaker merge, token count of every cleaned
 * segment.  Results stay on the device inside the context until the next trc_preprocess. */
int trc_preprocess(trc_ctx* ctx, const trc_batch* batch, void* stream);

/* Stage C: BigChunkeroosky.chunk_transcript + postprocess_chunks (big_chunkeroosky.py:46-145,544-567)
 * on the result of the last trc_preprocess.  doc_max_tokens[x] is max_tokens_per_chunk of document x
 * (host pointer, n_doc entries); context_tokens as in big_chunkeroosky.py:28,42. */
int trc_chunk(trc_ctx* ctx, const int32_t* doc_max_tokens, int32_t context_tokens, int32_t add_context,
              void* stream);

/* Counters of the current batch. */
enum {
  TRC_N_SEG = 0, TRC_N_KEPT, TRC_N_GROUPS, TRC_N_ITEMS, TRC_N_SUBS, TRC_N_CHUNKS, TRC_B_RAW, TRC_B_CLEAN,
  TRC_B_MERGED, TRC_B_TWC, TRC_TOKENS, TRC_N_DOC, TRC_N_LONG_SPANS, TRC_N_ERR_DOCS, TRC_N_SENT, TRC_N_GIANTS,
  TRC_N_AGG, TRC_N_AGG_MEM, TRC_B_AGG,
  TRC_N_WORDS,        /* whitespace-delimited words of the cleaned text that went through the word-level token counter */
  TRC_N_WORD_FIRST,   /* of those: not found in the word cache / launch memo (first sightings -> packed-word matcher)  */
  TRC_N_WORD_SLOW,    /* of those: left to the generic matcher + byte-pair merge (k_count_slow_words)                 */
  TRC_N_COUNTERS = 24
};
int trc_get_counts(trc_ctx* ctx, int64_t* counts /* TRC_N_COUNTERS */);

/* Result arrays (device resident).  trc_get_array copies array `which` to dst (host or device memory
 * per `memory`); with dst == NULL it only reports the size in bytes. */
enum {
  TRC_A_K2R = 0,      /* int32 [n_kept]      raw index of every kept (non-blank) segment                  */
  TRC_A_COFF,         /* int64 [n_kept+1]    offsets into TRC_A_CLEAN                                      */
  TRC_A_CLEAN,        /* uint8 [B_clean]     cleaned segment text (preprocessor.py:69-89)                  */
  TRC_A_CTOK,         /* uint32[n_kept]      tokens of " " + cleaned text                                  */
  TRC_A_GSTART,       /* int32 [m+1]         first kept segment of every merged segment                    */
  TRC_A_POFF,         /* int64 [n_kept+1]    piece offsets into TRC_A_MERGED                               */
  TRC_A_MERGED,       /* uint8 [B_merged]    merged segment text, groups back to back                      */
  TRC_A_GTS,          /* uint8 [m*48]        start_formatted / end_formatted, 24-byte slots [len][chars]   */
  TRC_A_GTOK,         /* int64 [m]           tokens of "[ts] speaker: text" (big_chunkeroosky.py:83)        */
  TRC_A_ITEM_GROUP,   /* int32 [n_items]                                                                  */
  TRC_A_ITEM_SUB,     /* int32 [n_items]     index into TRC_A_SUBS or -1                                   */
  TRC_A_ITEM_TOK,     /* int32 [n_items]                                                                  */
  TRC_A_ITEM_LEN,     /* int64 [n_items]     bytes of the item text inside TRC_A_TWC                       */
  TRC_A_ITEM_DST,     /* int64 [n_items]     offset of the item text inside TRC_A_TWC                      */
  TRC_A_SUBS,         /* trc_sub [n_subs]                                                                 */
  TRC_A_CHUNKS,       /* trc_chunk_rec [n_chunks]                                                         */
  TRC_A_SPK_LIST,     /* int32 [n_items]     sorted speaker ids of chunk c at [item0, item0+n_speakers)    */
  TRC_A_TWC,          /* uint8 [B_twc]       text_with_context of all chunks back to back                  */
  TRC_A_DOC_ERR,      /* int32 [n_doc]       0 ok, 1+k: special token k in the text (tiktoken ValueError), */
                      /*                     -1 unsupported input                                          */
  TRC_A_GDOC_OFF,     /* int64 [n_doc+1]     merged-segment range of every document                        */
  TRC_A_CDOC_OFF,     /* int64 [n_doc+1]     chunk range of every document                                 */
  TRC_A_CCHARS,       /* int32 [n_seg]       code points of the cleaned text of every RAW segment          */
  TRC_A_SENT,         /* trc_sent [n_sent]   sentence / clause records of the sentence path                */
  TRC_A_PDST,         /* int64 [n_kept]      offset inside TRC_A_TWC where every piece starts: its "[ts] " marker  */
                      /*                     if it has one, then its cleaned text (-1: the piece                  */
                      /*                     belongs to a sentence-path item; use TRC_A_CLEAN for it)             */
  TRC_A_AGG,          /* trc_agg [n_agg]     time-interval segments of trc_aggregate, documents back to back      */
  TRC_A_AGG_MEM,      /* trc_agg_mem [n_agg_mem]  their member segments, every interval sorted by start (stable)  */
  TRC_A_AGG_PIECE,    /* int32 []            kept-segment indices: the components that survive in a member        */
  TRC_A_AGG_TEXT,     /* uint8 [B_agg]       the aggregated texts back to back                                    */
  TRC_A_AGG_DOC_OFF,  /* int64 [n_doc+1]     trc_agg range of every document                                      */
  TRC_A_PROMPTS,      /* uint8 []            prompts of trc_assemble_prompts, chunks back to back                 */
  TRC_A_PROMPT_OFF,   /* int64 [n_chunks+1]  their offsets                                                        */
  TRC_A_CLEN32,       /* int32 [n_kept]      byte length of every kept segment's cleaned text (== differences of       */
                      /*                     TRC_A_COFF; half the bytes for callers that download results over PCIe)  */
  TRC_A__COUNT
};
int trc_get_array(trc_ctx* ctx, int which, void* dst, int64_t dst_bytes, int32_t memory, int64_t* nbytes,
                  void* stream);

/* Several result arrays in one call, e.g. everything main.py:148-149 needs for its chunk dicts: array which[k] is
 * copied to dst + offsets[k] (offsets chosen by the library, 256-byte aligned, returned together with the sizes);
 * one copy per array and ONE stream synchronisation when dst is host memory.  dst == NULL: sizes only. */
int trc_get_arrays(trc_ctx* ctx, const int32_t* which, int32_t n_which, void* dst, int64_t dst_bytes, int32_t memory,
                   int64_t* offsets /* n_which */, int64_t* nbytes /* n_which */, int64_t* total_bytes, void* stream);

/* preprocessor.py:91-107 format_timestamp == big_chunkeroosky.py:234-242 _format_time for n values with one small
 * kernel: out24 receives n slots of 24 bytes, [length][characters].  The resident batch is not touched (the other
 * scripts of the reference call _format_time per chunk: llm_executor.py:503, result_aggregator.py:569). */
int trc_format_timestamps(trc_ctx* ctx, const double* seconds, int64_t n, uint8_t* out24, int32_t memory, void* stream);

/* Record layouts of TRC_A_SUBS / TRC_A_CHUNKS / TRC_A_SENT (mirrored by ctypes.Structure in _native.py). */
typedef struct {
  int32_t group, kind, r0, r1, start_src, end_src, tok, pad;
  double start, end;
  int64_t text_len;
} trc_sub;

typedef struct {
  int32_t item0, item1, doc, chunk_index;
  int64_t token_count;
  double start_time, end_time, position_percentage;
  int32_t pct_is_int_zero, start_kind, start_ref, n_speakers, hdr_len, err;
  int64_t text_len, twc_off;
} trc_chunk_rec;

typedef struct {            /* sentence (kind 1) / clause (kind 2) of an over-budget un-merged segment */
  int32_t group, kind;
  int64_t off, len;         /* byte range in TRC_A_CLEAN */
  int32_t chars, tok;
  double start, end;
  int32_t sub, last_in_sub;
  int64_t off_in_sub;
  int32_t parent, ncl;
  int64_t cl0;
} trc_sent;

/* Fixed time intervals over the merged segments of the current batch: replaces aggregate_by_time_interval
 * (preprocessor.py:217-324; preprocess_transcript calls it when time_interval_seconds is set, :62-65).
 * Runs after trc_preprocess on the resident batch.  Interval i of a document spans
 * [start_time + i*interval_seconds, min(that + interval_seconds, end_time)) with start_time / end_time the
 * first segment's start / the last one's end (:232-242); a segment belongs to every interval it starts in or
 * spans (:253-256); a combined segment keeps only the components that do (:262-287, its text is rebuilt as
 * "[ts] text" joined by spaces) and is left out when none does; an interval's text is its members sorted by
 * start, "[ts speaker] text" joined by blank lines (:302-307); empty intervals produce nothing (:292).
 * The 'speaker' field of an interval (:311) is a join over a Python set, whose order depends on the
 * interpreter's string hashing: the host builds it from the member records.
 * Results: TRC_A_AGG / TRC_A_AGG_MEM / TRC_A_AGG_PIECE / TRC_A_AGG_TEXT / TRC_A_AGG_DOC_OFF and the counters
 * TRC_N_AGG, TRC_N_AGG_MEM, TRC_B_AGG.  interval_seconds must be a non-zero number (Python raises
 * ZeroDivisionError, :236); a non-positive quotient yields no intervals, as range() does. */
int trc_aggregate(trc_ctx* ctx, double interval_seconds, void* stream);

typedef struct {
  int32_t doc, interval_index;
  double start, end;          /* interval_start, interval_end (:241-242) */
  int64_t mem0, mem1;         /* members [mem0, mem1) in TRC_A_AGG_MEM */
  int64_t text_off, text_len; /* 'text' in TRC_A_AGG_TEXT */
  uint8_t ts[48];             /* start_formatted / end_formatted, 24-byte slots [len][chars] */
} trc_agg;

typedef struct {
  int32_t group, n_pieces;    /* merged segment (index into the document-ordered processed list of the batch);
                                 surviving components, 0 for a segment without 'segment_timestamps' */
  int64_t piece0;             /* components [piece0, piece0+n_pieces) in TRC_A_AGG_PIECE, sorted by start (stable) */
  int64_t text_off, text_len; /* the member's text (what 'segment_timestamps'[k]['text'] holds, :318) in TRC_A_AGG_TEXT */
} trc_agg_mem;

/* Prompt assembly from the device buffers: prompt(c) = lit[0] + twc(c) + lit[1] + twc(c) + ... + lit[n_lit-1] for
 * every chunk c, where twc(c) is the chunk's text_with_context.  Replaces prompt_template.format(transcript=
 * chunk['text_with_context'], summary_type=...) (llm_executor.py:187-190): the caller formats the template once
 * with a placeholder for the transcript and passes the literal pieces around it (n_lit >= 1; lit_off has n_lit+1
 * entries; host pointers).  Runs after trc_chunk; results: TRC_A_PROMPTS / TRC_A_PROMPT_OFF. */
int trc_assemble_prompts(trc_ctx* ctx, const uint8_t* lit_bytes, const int64_t* lit_off, int32_t n_lit, void* stream);

/* Ingest (host side): n_doc transcripts as JSON text (document x = json[doc_off[x] .. doc_off[x+1])) -> the packed
 * batch layout above, without building Python objects.  Replaces json.load + transcript_data.get('segments', [])
 * (main.py:337-339, :133) and the per-segment segment.get(...) defaults (preprocessor.py:46-50): missing text /
 * speaker -> "", missing start / end -> the int 0, a repeated key keeps its last value, unknown members are
 * validated and skipped.  Numbers go through strtod / strtoll (correctly rounded, as Python's float() / int());
 * start_kind / end_kind say which Python type json.load would have produced (0 float, 1 int, 2 bool).  Invalid
 * JSON, invalid UTF-8, and inputs the reference itself fails on at first use (segment not an object, text or
 * speaker not a string, null / string / NaN / Infinity times, lone surrogates) or that do not fit the layout
 * (integer times beyond 2^53, which float64 cannot hold exactly) are errors: err receives a message with the document and byte offset.
 * Documents are parsed in parallel on n_threads host threads.  The result owns its arrays; trc_ingest_view fills a
 * trc_batch (memory = TRC_MEM_HOST, merge parameters untouched) whose pointers stay valid until trc_ingest_free. */
typedef struct trc_ingest trc_ingest;
int trc_ingest_json(const uint8_t* json, const int64_t* doc_off, int64_t n_doc, int32_t n_threads, trc_ingest** out,
                    char* err, int64_t err_cap);
int trc_ingest_view(const trc_ingest* r, trc_batch* batch, const uint8_t** start_kind, const uint8_t** end_kind);
void trc_ingest_free(trc_ingest* r);

/* Token counting service: counts[i] = len(encode([" "] + bytes[off[i] .. off[i+1]))) for n strings.
 * Replaces len(self.tokenizer.encode(x)) (big_chunkeroosky.py:83,298,370,510).  special[i] (may be
 * NULL) receives the index of a special-token literal found in string i, or -1. */
int trc_count_tokens(trc_ctx* ctx, const uint8_t* bytes, const int64_t* off, int64_t n, int32_t leading_space,
                     int32_t memory, uint32_t* counts, int32_t* special, void* stream);

/* Wall-clock free timing helper: device time in ms of the last trc_preprocess / trc_chunk pair, measured
 * with CUDA events on the stream they ran on (0 if events were not recorded). */
int trc_last_stage_ms(trc_ctx* ctx, float* preprocess_ms, float* chunk_ms);
int trc_kernel_launches(trc_ctx* ctx, int64_t* launches);

/* Per-kernel device timing for bench.py's roofline line: with profiling on, every kernel launch is
 * bracketed by CUDA events on its stream; trc_get_profile returns "name total_ms launches\n" lines
 * accumulated since profiling was switched on.  (No reference counterpart: the reference only has
 * time.time() around the whole run, main.py:110,239.) */
int trc_set_profiling(trc_ctx* ctx, int32_t on);
int trc_get_profile(trc_ctx* ctx, char* buf, int64_t buf_bytes);

/* Token count of ONE pre-tokenised span without allocating anything.  (No counterpart in the
 * reference: tiktoken allocates freely on the host.)  For embedders that must not allocate device
 * memory while running - a server whose own pool already owns all of HBM, a plugin host that bans
 * driver allocations after start-up, a loop that cannot take the device-wide stall of cudaMalloc/
 * cudaFree - this does what the long-span path of trc_count_tokens does (vocabulary hit => 1, else
 * byte-pair merge by minimum rank, leftmost first) with its work arrays in a fixed 1 MiB array inside
 * the library image, present on every GPU that loads the library until the process exits.
 *   d_span  : DEVICE pointer, n_bytes bytes forming one span (the caller has already split the text);
 *   d_count : DEVICE pointer to one uint32 that receives the count; the call is asynchronous on `stream`.
 * Limit: 1 <= n_bytes <= 65536, larger spans are rejected with TRC_EINVAL and a message.
 * The first call on a device looks up the array's device address once; make that call before starting
 * any stream capture.  All calls on a device share the array and it is not cleared between calls (the
 * kernel initialises what it reads), so calls that overlap on the GPU race: issue them on one stream
 * or order them yourself. */
int trc_bpe_span_count_static(trc_ctx* ctx, const uint8_t* d_span, int64_t n_bytes, uint32_t* d_count, void* stream);

#ifdef __cplusplus
}
#endif
#endif
