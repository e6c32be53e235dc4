"""Preprocess oracle (TEST INFRASTRUCTURE): restates /root/reference/preprocessor.py:15-215 in the
"local rule" form the CUDA kernels implement, so the tests compare like with like.
Pinned against the real preprocessor.py by tools/make_golden.py and tests/test_oracle_vs_golden.py.
"""
from . import tokenizer as _tk

_PROPS = None


def _props():
    global _PROPS
    if _PROPS is None:
        _PROPS = _tk.load_props()
    return _PROPS


def clean_text(text):
    """preprocessor.py:69-89 as one pass: whitespace collapse (:80), duplicate-word deletion (:84),
    space insertion after [.!?] before an ASCII letter (:87)."""
    P = _props()
    # pass 1: words = maximal runs of non-isspace code points, joined by one U+0020
    s = []
    pending = False
    for ch in text:
        if P[ord(ch)] & 1:
            pending = True
        else:
            if pending and s:
                s.append(" ")
            pending = False
            s.append(ch)
    s = "".join(s)
    # pass 2 (local rule): a \w-run equal to the \w-run that ends exactly one space before it is dropped
    # together with that space.  pass 3 is applied while emitting.
    out = []
    n = len(s)
    i = 0
    prev_word = None          # (start, end) in s of the last \w-run
    while i < n:
        ch = s[i]
        if P[ord(ch)] & 2:
            j = i
            while j < n and P[ord(s[j])] & 2:
                j += 1
            if (prev_word is not None and prev_word[1] == i - 1 and s[i - 1] == " "
                    and s[prev_word[0]:prev_word[1]] == s[i:j]):
                out.pop()                       # the single space
            else:
                _emit(out, s[i:j])
            prev_word = (i, j)
            i = j
        else:
            _emit(out, ch)
            i += 1
    return "".join(out)


def _emit(out, piece):
    # pass 3: [.!?] immediately followed by [A-Za-z] gets a space in between
    if out and out[-1] and out[-1][-1] in ".!?" and piece[0].isascii() and piece[0].isalpha():
        out.append(" ")
    out.append(piece)


def format_ts(x):
    """preprocessor.py:91-107 == big_chunkeroosky.py:234-242."""
    t = int(x)
    h, r = t // 3600, t % 3600
    m, s = r // 60, r % 60
    return "%02d:%02d:%02d" % (h, m, s) if h > 0 else "%02d:%02d" % (m, s)


def group_bounds(segs, max_duration):
    """preprocessor.py:132-158 -> list of (first, last+1) index pairs."""
    out = []
    a = 0
    cur = segs[0]["end"] - segs[0]["start"]
    for i in range(1, len(segs)):
        d = segs[i]["end"] - segs[i]["start"]
        if segs[i]["speaker"] != segs[a]["speaker"] or (max_duration is not None and cur + d > max_duration):
            out.append((a, i))
            a, cur = i, d
        else:
            cur += d
    out.append((a, len(segs)))
    return out


def preprocess(segments, merge_same_speaker=True, time_interval_seconds=None, max_segment_duration=120,
               preserve_timestamps=True):
    P = _props()
    proc = []
    for seg in segments:
        raw = seg.get("text", "")
        if all(P[ord(c)] & 1 for c in raw):
            continue
        proc.append({"start": seg.get("start", 0), "end": seg.get("end", 0),
                     "start_formatted": format_ts(seg.get("start", 0)),
                     "end_formatted": format_ts(seg.get("end", 0)),
                     "speaker": seg.get("speaker", ""), "text": clean_text(raw)})
    if not proc:
        return proc
    out = _merge(proc, max_segment_duration, preserve_timestamps) if merge_same_speaker else proc
    return aggregate(out, time_interval_seconds) if time_interval_seconds else out


def _merge(proc, max_segment_duration, preserve_timestamps):
    out = []
    for a, b in group_bounds(proc, max_segment_duration):
        if b - a == 1:
            out.append(proc[a])
            continue
        grp = proc[a:b]
        if preserve_timestamps:
            text = " ".join("[" + format_ts(g["start"]) + "] " + g["text"] for g in grp)
        else:
            text = " ".join(g["text"] for g in grp)
        out.append({"start": grp[0]["start"], "end": grp[-1]["end"],
                    "start_formatted": format_ts(grp[0]["start"]), "end_formatted": format_ts(grp[-1]["end"]),
                    "speaker": grp[0]["speaker"], "text": text, "is_combined": True,
                    "original_segments": b - a,
                    "segment_timestamps": [{"start": g["start"], "end": g["end"], "text": g["text"]} for g in grp]})
    return out


def _in_interval(lo, hi, s, e):
    """preprocessor.py:253-256 / :270-273: starts inside [lo, hi), or starts at/before lo and ends after lo."""
    return (lo <= s < hi) or (s <= lo and e > lo)


def aggregate(segments, interval_seconds):
    """preprocessor.py:217-324 (aggregate_by_time_interval), prints left out.  The joined speaker string takes
    the iteration order of a Python set of str, like the reference (:296,:311) - identical inside one process."""
    import math
    if not segments:
        return []
    t0, t1 = segments[0]["start"], segments[-1]["end"]
    out = []
    for i in range(math.ceil((t1 - t0) / interval_seconds)):
        lo = t0 + (i * interval_seconds)
        hi = min(lo + interval_seconds, t1)
        members = []
        for seg in segments:
            if not _in_interval(lo, hi, seg["start"], seg["end"]):
                continue
            if "segment_timestamps" not in seg:
                members.append(dict(seg))
                continue
            kept = [ts for ts in seg["segment_timestamps"] if _in_interval(lo, hi, ts["start"], ts["end"])]
            if kept:
                m = dict(seg)
                m["segment_timestamps"] = kept
                m["text"] = " ".join("[" + format_ts(ts["start"]) + "] " + ts["text"]
                                     for ts in sorted(kept, key=lambda ts: ts["start"]))
                members.append(m)
        if not members:
            continue
        who = set(m["speaker"] for m in members)
        ordered = sorted(members, key=lambda m: m["start"])
        out.append({"start": lo, "end": hi, "start_formatted": format_ts(lo), "end_formatted": format_ts(hi),
                    "speaker": ", ".join(who) if len(who) > 1 else next(iter(who)),
                    "text": "\n\n".join("[" + format_ts(m["start"]) + " " + m["speaker"] + "] " + m["text"] for m in ordered),
                    "is_aggregated": True, "interval_index": i, "original_segments": len(members),
                    "segment_timestamps": [{"start": m["start"], "end": m["end"], "speaker": m["speaker"], "text": m["text"]}
                                           for m in ordered]})
    return out
