"""Drop-in for the reference's `preprocessor` module (/root/reference/preprocessor.py).

`preprocess_transcript`, `clean_text`, `format_timestamp`, `combine_same_speaker_segments` and
`create_combined_segment` keep the reference's names, arguments, printed lines and dict formats; the
work runs in the CUDA kernels of libtrchunk.so (stage P, include/trchunk.h: trc_preprocess).
`aggregate_by_time_interval` (preprocessor.py:217-324) runs on the device too (trc_aggregate).
`extract_speakers` / `get_transcript_duration` (:326-361) are trivial host accessors.
"""
import math
from typing import Any, Dict, List, Optional, Tuple

from . import engine as _eng


def _fingerprint(d):
    """Cheap content fingerprint of a processed dict: an in-place edit between preprocess_transcript and
    chunk_transcript must be honoured (the reference would), so the resident batch is only reused for untouched dicts."""
    return hash((d.get("text"), d.get("speaker"), repr(d.get("start")), repr(d.get("end")), len(d)))


def resident_handle(segments):
    """The _trc handle of `segments` when it is this package's own, unedited output; else None."""
    h = getattr(segments, "_trc", None)
    if h is None or len(segments) != len(h["ids"]):
        return None
    for d, i, f in zip(segments, h["ids"], h["fp"]):
        if id(d) != i or _fingerprint(d) != f:
            return None
    return h


class ProcessedSegments(list):
    """The list preprocess_transcript returns.  A plain list of dicts for every consumer; it also
    remembers the packed input so that BigChunkeroosky.chunk_transcript can chunk on the device without
    re-packing (the batch stays resident in the engine while its serial matches)."""
    _trc = None


def preprocess_transcript(segments: List[Dict[Any, Any]], merge_same_speaker: bool = True,
                          time_interval_seconds: Optional[int] = None, max_segment_duration: Optional[int] = 120,
                          preserve_timestamps: bool = True) -> List[Dict[Any, Any]]:
    """preprocessor.py:15-67."""
    eng = _eng.default_engine()
    batch = _eng.pack_documents([segments])
    params = dict(merge_same_speaker=bool(merge_same_speaker), max_segment_duration=max_segment_duration,
                  preserve_timestamps=bool(preserve_timestamps))
    serial = eng.preprocess(batch, want_merged_text=True, **params)
    P = eng.fetch_processed()
    out = ProcessedSegments(_eng.build_processed(batch, P, 0))
    n_kept = len(P["k2r"])
    if merge_same_speaker and n_kept:
        uniq = len({batch.speakers[batch.speaker[int(r)]] for r in P["k2r"]})
        print(f"Preprocessing: Found {uniq} unique speakers in transcript")                    # :130
        print(f"Preprocessing: Combined {n_kept} segments into {len(out)} segments")           # :161
        print(f"Preprocessing: Compression ratio: {len(out) / n_kept:.2f}")                    # :163
    out._trc = {"engine": eng, "serial": serial, "batch": batch, "params": params, "P": P,
                "ids": [id(d) for d in out], "fp": [_fingerprint(d) for d in out]}
    if time_interval_seconds and out:
        return aggregate_by_time_interval(out, time_interval_seconds)
    return out


def clean_text(text: str) -> str:
    """preprocessor.py:69-89 (one segment through the same kernel, on a scratch context: the batch that
    preprocess_transcript left resident on the default engine stays where it is)."""
    eng = _eng.scratch_engine()
    batch = _eng.pack_documents([[{"start": 0, "end": 0, "text": text, "speaker": ""}]])
    eng.preprocess(batch, merge_same_speaker=False, want_merged_text=False)
    P_clean = eng.array("CLEAN", "u1")
    return P_clean.tobytes().decode("utf-8")


def format_timestamp(seconds: float) -> str:
    """preprocessor.py:91-107: one small kernel (trc_format_timestamps, trc_format.cuh); the resident batch is not
    disturbed.  int() of the reference: truncation toward zero, bool is an int, NaN / infinity raise as in Python."""
    return _eng.default_engine().format_timestamps(_eng._finite_f64([seconds]))[0]


def combine_same_speaker_segments(segments: List[Dict[Any, Any]], max_duration: Optional[int] = 120,
                                  preserve_timestamps: bool = True) -> List[Dict[Any, Any]]:
    """preprocessor.py:109-165.  Input: already processed (cleaned) segment dicts; only the merge runs."""
    if not segments:
        return []
    eng = _eng.default_engine()
    pieces = [{"start": s["start"], "end": s["end"], "text": s["text"], "speaker": s["speaker"]} for s in segments]
    batch = _eng.pack_documents([pieces])
    batch.groups = {}                       # already processed: texts verbatim, grouping computed by the merge kernels
    eng.preprocess(batch, merge_same_speaker=True, max_segment_duration=max_duration,
                   preserve_timestamps=bool(preserve_timestamps), want_merged_text=True)
    P = eng.fetch_processed()
    out = _eng.build_processed(batch, P, 0, piece_dicts=segments)
    print(f"Preprocessing: Found {len({s['speaker'] for s in segments})} unique speakers in transcript")   # :130
    print(f"Preprocessing: Combined {len(segments)} segments into {len(out)} segments")                     # :161
    print(f"Preprocessing: Compression ratio: {len(out) / len(segments):.2f}")                              # :163
    return out


def create_combined_segment(segments: List[Dict[Any, Any]], preserve_timestamps: bool = True) -> Dict[Any, Any]:
    """preprocessor.py:167-215: the given segments as ONE combined segment."""
    if not segments:
        return {}
    if len(segments) == 1:
        return segments[0]
    eng = _eng.default_engine()
    sp = segments[0]["speaker"]
    pieces = [{"start": s["start"], "end": s["end"], "text": s["text"], "speaker": sp} for s in segments]
    batch = _eng.pack_documents([pieces])
    batch.groups = {}
    eng.preprocess(batch, merge_same_speaker=True, max_segment_duration=None,
                   preserve_timestamps=bool(preserve_timestamps), want_merged_text=True)
    return _eng.build_processed(batch, eng.fetch_processed(), 0, piece_dicts=segments)[0]


def aggregate_by_time_interval(segments: List[Dict[Any, Any]], interval_seconds: int) -> List[Dict[Any, Any]]:
    """preprocessor.py:217-324: the interval membership tests, the two sorts and the text assembly run in
    trc_aggregate; this wrapper prints the reference's lines and builds the dicts."""
    if not segments:
        return []
    start_time, end_time = segments[0]["start"], segments[-1]["end"]
    num_intervals = math.ceil((end_time - start_time) / interval_seconds)                        # :236
    h = resident_handle(segments)
    if h is not None:
        eng, batch = h["engine"], h["batch"]
        if eng.batch_serial != h["serial"]:
            h["serial"] = eng.preprocess(batch, want_merged_text=True, **h["params"])
        P = h["P"]
    else:
        eng = _eng.default_engine()
        batch = _eng.pack_processed(segments, for_aggregate=True)
        P = None
        for preserve in (True, False):           # how the combined texts were joined (:190-200) does not matter here,
            eng.preprocess(batch, merge_same_speaker=True, max_segment_duration=None,   # but start/end must be the pieces'
                           preserve_timestamps=preserve, want_merged_text=True)
            if int(eng.array("DOC_ERR", "<i4")[0]) != -1:
                P = eng.fetch_processed()
                break
        if P is None:
            raise NotImplementedError("processed segments whose own start/end/text differ from what their "
                                      "segment_timestamps imply are outside the device path")
    gts = P["gts"].tobytes()
    o = (len(segments) - 1) * 48
    print(f"Creating {num_intervals} time intervals of {interval_seconds} seconds each")        # :240
    print(f"Total time range: {gts[1:1 + gts[0]].decode('ascii')} - {gts[o + 25:o + 25 + gts[o + 24]].decode('ascii')} "
          f"({end_time - start_time:.2f}s)")                                                     # :241
    out = []
    if num_intervals > 0:
        eng.aggregate(interval_seconds)
        out = _eng.build_aggregated(batch, P, eng.fetch_aggregated(), segments, interval_seconds, 0)
    print(f"Created {len(out)} time-interval segments (some intervals may be empty)")           # :323
    return out


def extract_speakers(segments: List[Dict[Any, Any]]) -> List[str]:
    """preprocessor.py:326-342."""
    return sorted({s["speaker"] for s in segments if "speaker" in s and s["speaker"]})


def get_transcript_duration(segments: List[Dict[Any, Any]]) -> Tuple[float, str]:
    """preprocessor.py:344-361."""
    if not segments:
        return 0.0, "00:00"
    duration = segments[-1]["end"] - segments[0]["start"]
    return duration, format_timestamp(duration)
