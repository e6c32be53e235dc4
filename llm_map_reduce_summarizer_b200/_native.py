"""ctypes binding of include/trchunk.h.  Fails loudly when the CUDA library is missing."""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
DEFAULT_LIB = os.environ.get("TRCHUNK_LIB") or os.path.join(_HERE, "libtrchunk.so")

TRC_MEM_HOST, TRC_MEM_DEVICE = 0, 1
COUNTERS = ["n_seg", "n_kept", "n_groups", "n_items", "n_subs", "n_chunks", "b_raw", "b_clean", "b_merged", "b_twc",
            "tokens", "n_doc", "n_long_spans", "n_err_docs", "n_sent", "n_giants", "n_agg", "n_agg_mem", "b_agg",
            "n_words", "n_word_first", "n_word_slow"]
N_COUNTERS = 24
ARRAYS = ["K2R", "COFF", "CLEAN", "CTOK", "GSTART", "POFF", "MERGED", "GTS", "GTOK", "ITEM_GROUP", "ITEM_SUB", "ITEM_TOK",
          "ITEM_LEN", "ITEM_DST", "SUBS", "CHUNKS", "SPK_LIST", "TWC", "DOC_ERR", "GDOC_OFF", "CDOC_OFF", "CCHARS", "SENT", "PDST",
          "AGG", "AGG_MEM", "AGG_PIECE", "AGG_TEXT", "AGG_DOC_OFF", "PROMPTS", "PROMPT_OFF", "CLEN32"]
A = {n: i for i, n in enumerate(ARRAYS)}
ERRORS = {-1: "TRC_ECUDA", -2: "TRC_EINVAL", -3: "TRC_ENOVOCAB", -4: "TRC_EUNSUPPORTED"}


class Batch(C.Structure):
    _fields_ = [("text", C.c_void_p), ("seg_off", C.c_void_p), ("start", C.c_void_p), ("end", C.c_void_p),
                ("speaker", C.c_void_p), ("n_seg", C.c_int64), ("doc_off", C.c_void_p), ("n_doc", C.c_int64),
                ("spk_bytes", C.c_void_p), ("spk_off", C.c_void_p), ("n_spk", C.c_int32), ("memory", C.c_int32),
                ("merge_same_speaker", C.c_int32), ("has_max_duration", C.c_int32), ("max_duration", C.c_double),
                ("preserve_timestamps", C.c_int32), ("want_merged_text", C.c_int32),
                ("already_processed", C.c_int32), ("reserved_", C.c_int32), ("group_off", C.c_void_p),
                ("n_groups", C.c_int64), ("group_start", C.c_void_p), ("group_end", C.c_void_p),
                ("full_text", C.c_void_p), ("full_off", C.c_void_p),
                ("seg_len32", C.c_void_p), ("text_bytes", C.c_int64)]


SUB_DTYPE = [("group", "<i4"), ("kind", "<i4"), ("r0", "<i4"), ("r1", "<i4"), ("start_src", "<i4"), ("end_src", "<i4"),
             ("tok", "<i4"), ("pad", "<i4"), ("start", "<f8"), ("end", "<f8"), ("text_len", "<i8")]
CHUNK_DTYPE = [("item0", "<i4"), ("item1", "<i4"), ("doc", "<i4"), ("chunk_index", "<i4"), ("token_count", "<i8"),
               ("start_time", "<f8"), ("end_time", "<f8"), ("position_percentage", "<f8"), ("pct_is_int_zero", "<i4"),
               ("start_kind", "<i4"), ("start_ref", "<i4"), ("n_speakers", "<i4"), ("hdr_len", "<i4"), ("err", "<i4"),
               ("text_len", "<i8"), ("twc_off", "<i8")]
SENT_DTYPE = [("group", "<i4"), ("kind", "<i4"), ("off", "<i8"), ("len", "<i8"), ("chars", "<i4"), ("tok", "<i4"),
              ("start", "<f8"), ("end", "<f8"), ("sub", "<i4"), ("last_in_sub", "<i4"), ("off_in_sub", "<i8"),
              ("parent", "<i4"), ("ncl", "<i4"), ("cl0", "<i8")]
AGG_DTYPE = [("doc", "<i4"), ("interval_index", "<i4"), ("start", "<f8"), ("end", "<f8"), ("mem0", "<i8"), ("mem1", "<i8"),
             ("text_off", "<i8"), ("text_len", "<i8"), ("ts", "u1", (48,))]
AGG_MEM_DTYPE = [("group", "<i4"), ("n_pieces", "<i4"), ("piece0", "<i8"), ("text_off", "<i8"), ("text_len", "<i8")]

EXPORTS = ["trc_create", "trc_destroy", "trc_last_error", "trc_load_tables", "trc_load_vocab", "trc_preprocess",
           "trc_chunk", "trc_aggregate", "trc_get_counts", "trc_get_array", "trc_count_tokens", "trc_last_stage_ms", "trc_kernel_launches", "trc_set_profiling",
           "trc_get_profile", "trc_get_arrays", "trc_format_timestamps", "trc_bpe_span_count_static", "trc_ingest_json", "trc_ingest_view", "trc_ingest_free", "trc_assemble_prompts"]


def load_library(path=None):
    path = path or DEFAULT_LIB
    if not os.path.exists(path):
        raise RuntimeError(
            f"{path} is missing: build it with `python -m llm_map_reduce_summarizer_b200.build` "
            "(nvcc, sm_100a).  There is no CPU fallback for this path.")
    lib = C.CDLL(path)
    vp, i32, i64 = C.c_void_p, C.c_int32, C.c_int64
    lib.trc_create.argtypes = [C.c_int, C.POINTER(vp)]
    lib.trc_destroy.argtypes = [vp]
    lib.trc_destroy.restype = None
    lib.trc_last_error.argtypes = [vp]
    lib.trc_last_error.restype = C.c_char_p
    lib.trc_load_tables.argtypes = [vp, vp]
    lib.trc_load_vocab.argtypes = [vp, vp, vp, vp, i32, vp, vp, i32]
    lib.trc_preprocess.argtypes = [vp, C.POINTER(Batch), vp]
    lib.trc_chunk.argtypes = [vp, vp, i32, i32, vp]
    if hasattr(lib, "trc_aggregate"):
        lib.trc_aggregate.argtypes = [vp, C.c_double, vp]
    lib.trc_get_counts.argtypes = [vp, vp]
    lib.trc_get_array.argtypes = [vp, C.c_int, vp, i64, i32, C.POINTER(i64), vp]
    if hasattr(lib, "trc_get_arrays"):
        lib.trc_get_arrays.argtypes = [vp, vp, i32, vp, i64, i32, vp, vp, C.POINTER(i64), vp]
        lib.trc_format_timestamps.argtypes = [vp, vp, i64, vp, i32, vp]
    lib.trc_count_tokens.argtypes = [vp, vp, vp, i64, i32, i32, vp, vp, vp]
    lib.trc_last_stage_ms.argtypes = [vp, C.POINTER(C.c_float), C.POINTER(C.c_float)]
    lib.trc_kernel_launches.argtypes = [vp, C.POINTER(i64)]
    lib.trc_set_profiling.argtypes = [vp, i32]
    lib.trc_get_profile.argtypes = [vp, C.c_char_p, i64]
    if hasattr(lib, "trc_assemble_prompts"):
        lib.trc_assemble_prompts.argtypes = [vp, vp, vp, i32, vp]
    if hasattr(lib, "trc_ingest_json"):
        lib.trc_ingest_json.argtypes = [vp, vp, i64, i32, C.POINTER(vp), C.c_char_p, i64]
        lib.trc_ingest_view.argtypes = [vp, C.POINTER(Batch), C.POINTER(vp), C.POINTER(vp)]
        lib.trc_ingest_free.argtypes = [vp]
        lib.trc_ingest_free.restype = None
    if hasattr(lib, "trc_bpe_span_count_static"):
        lib.trc_bpe_span_count_static.argtypes = [vp, vp, i64, vp, vp]
    return lib
