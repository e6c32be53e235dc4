"""Drop-in for the reference's `preprocessor` module (/root/reference/preprocessor.py).

`preprocess_transcript`, `clean_text`, `format_timestamp`, `combine_same_speaker_segments` and
`create_combined_segment` keep the reference's names, arguments, printed lines and dict formats; the
work runs in the CUDA kernels of libtrchunk.so (stage P, include/trchunk.h: trc_preprocess).
`aggregate_by_time_interval` (preprocessor.py:217-324) is outside the hot path (main.py never passes
time_interval_seconds) and raises NotImplementedError rather than running on the CPU.
`extract_speakers` / `get_transcript_duration` (:326-361) are trivial host accessors.
"""
from typing import Any, Dict, List, Optional, Tuple

from . import engine as _eng


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
                "ids": [id(d) for d in out]}
    if time_interval_seconds and out:
        return aggregate_by_time_interval(out, time_interval_seconds)
    return out


def clean_text(text: str) -> str:
    """preprocessor.py:69-89 (one segment through the same kernel)."""
    eng = _eng.default_engine()
    batch = _eng.pack_documents([[{"start": 0, "end": 0, "text": text, "speaker": ""}]])
    eng.preprocess(batch, merge_same_speaker=False, want_merged_text=False)
    P_clean = eng.array("CLEAN", "u1")
    return P_clean.tobytes().decode("utf-8")


def format_timestamp(seconds: float) -> str:
    """preprocessor.py:91-107 (formatted on the device; see trc_format.cuh)."""
    eng = _eng.default_engine()
    batch = _eng.pack_documents([[{"start": seconds, "end": seconds, "text": "x", "speaker": ""}]])
    eng.preprocess(batch, merge_same_speaker=False, want_merged_text=False)
    g = eng.array("GTS", "u1").tobytes()
    return g[1:1 + g[0]].decode("ascii")


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
    """preprocessor.py:217-324 - not on the hot path (SURVEY.md §8f 'next')."""
    raise NotImplementedError("aggregate_by_time_interval is outside the accelerated path (SURVEY.md §8f)")


def extract_speakers(segments: List[Dict[Any, Any]]) -> List[str]:
    """preprocessor.py:326-342."""
    return sorted({s["speaker"] for s in segments if "speaker" in s and s["speaker"]})


def get_transcript_duration(segments: List[Dict[Any, Any]]) -> Tuple[float, str]:
    """preprocessor.py:344-361."""
    if not segments:
        return 0.0, "00:00"
    duration = segments[-1]["end"] - segments[0]["start"]
    return duration, format_timestamp(duration)
