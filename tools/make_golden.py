#!/usr/bin/env python
"""Generate tests/golden/* by executing the UNMODIFIED reference (/root/reference) in the build
container (nltk shim + repo rank table, oracle/refharness.py).  The GPU box has no /root/reference,
so the vectors are committed.  Re-run: python tools/make_golden.py
"""
import gzip, hashlib, json, os, random, sys

ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from oracle import refharness  # noqa: E402
import fuzzgen  # noqa: E402

G = os.path.join(ROOT, "tests", "golden")


def sha(obj):
    return hashlib.sha256(json.dumps(obj, sort_keys=True).encode()).hexdigest()


def dump_gz(path, obj):
    with gzip.GzipFile(path, "wb", mtime=0) as f:
        f.write(json.dumps(obj, separators=(",", ":")).encode())


def main():
    rp, rc = refharness.load()
    segs = json.load(open("/root/reference/transcript-example.json"))["segments"]
    dump_gz(os.path.join(G, "example_segments.json.gz"), segs)
    ex = {}
    for name, kw in {"default": dict(max_tokens=4000), "t2000": dict(max_tokens=2000),
                     "t400_nomerge": dict(max_tokens=400, merge=False),
                     "t500_cap30": dict(max_tokens=500, max_dur=30),
                     "t300_capNone": dict(max_tokens=300, max_dur=None),
                     "t1000_noctx_plain": dict(max_tokens=1000, preserve=False, add_context=False)}.items():
        proc, chunks = refharness.run(segs, **kw)
        ex[name] = {"kw": kw, "n_proc": len(proc), "n_chunks": len(chunks), "proc_sha256": sha(proc),
                    "chunks_sha256": sha(chunks), "token_counts": [c["token_count"] for c in chunks],
                    "end_times": [c["end_time"] for c in chunks],
                    "pct": [c["position_percentage"] for c in chunks],
                    "first_header": chunks[0]["text_with_context"][:260],
                    "n_sub": sum(1 for c in chunks for s in c["segments"] if s.get("is_sub_chunk"))}
    json.dump(ex, open(os.path.join(G, "example_expected.json"), "w"), indent=1)

    # known-answer vectors, produced by the real functions
    ct = ["  the the  cat\t\nsat ", "The the the end", "is is is is", "a a.b", "go go.Go", "website.com", "e.g.this",
          "3.5 vs 3.x", "ok!!yes", "x\xa0\xa0y z　w", "a\x1cb\x85c", "café café", "你好 你好", "1 1 2 2",
          "_a _a", "no no-no", "ha ha ha. ha", "word word, word", "the​the the", "abc ab ab", "x́ x́",
          ".A", "that that's", "what's s s", "non-stop stop", "don't don't", "ab ab abc", "A.B.C.D", "wait...what",
          "hi?Ünï", "", " ", "a", "a a", "a  a", "-a a", "a a_", "xab ab", "aab ab", "é é é", "a.a a"]
    ft = [0, 59.999, 3599.99, 3600, 360000, 1e6, -1, -61, -0.5, True, 26561.26, 86399.5, 7.0, 12345678901]
    merges = []
    for name, segl, cap in [
        ("7x40", [{"start": 40 * i, "end": 40 * (i + 1), "text": "x%d" % i, "speaker": "A"} for i in range(7)], 120),
        ("3000x0.1", [{"start": 0.1 * i, "end": 0.1 * (i + 1), "text": "w", "speaker": "A"} for i in range(3000)], 120),
        ("500_1_1", [{"start": 0, "end": 500, "text": "a", "speaker": "A"}, {"start": 500, "end": 501, "text": "b", "speaker": "A"},
                     {"start": 501, "end": 502, "text": "c", "speaker": "A"}], 120),
        ("capNone", [{"start": 0, "end": 500, "text": "a", "speaker": "A"}, {"start": 500, "end": 501, "text": "b", "speaker": "A"},
                     {"start": 501, "end": 502, "text": "c", "speaker": "A"}], None),
        ("missing_spk", [{"start": 0, "end": 1, "text": "a", "speaker": "A"}, {"start": 1, "end": 2, "text": "b"},
                         {"start": 2, "end": 3, "text": "  ", "speaker": "A"}, {"start": 3, "end": 4, "text": "c", "speaker": "A"}], 120),
    ]:
        import contextlib, io
        with contextlib.redirect_stdout(io.StringIO()):
            out = rp.preprocess_transcript(segl, True, None, cap)
        merges.append({"name": name, "segments": segl, "cap": cap,
                       "group_sizes": [o.get("original_segments", 1) for o in out], "sha256": sha(out)})
    json.dump({"clean_text": [[s, rp.clean_text(s)] for s in ct],
               "format_timestamp": [[x, rp.format_timestamp(x)] for x in ft],
               "merge": merges}, open(os.path.join(G, "kat.json"), "w"), ensure_ascii=True, indent=0)

    # differential fuzz cases with full reference outputs
    rng = random.Random(20261106)
    cases = []
    for it in range(160):
        s = fuzzgen.rand_transcript(rng, long_p=0.08 if it % 2 else 0.0)
        kw = dict(max_tokens=rng.choice([160, 170, 200, 260, 400, 4000]), max_dur=rng.choice([1, 30, 120, None]),
                  merge=rng.random() < 0.85, preserve=rng.random() < 0.9, add_context=rng.random() < 0.9)
        try:
            proc, chunks = refharness.run(s, **kw)
            err = None
        except ValueError as e:
            rp_, _ = refharness.load()
            import contextlib, io
            with contextlib.redirect_stdout(io.StringIO()):
                proc = rp_.preprocess_transcript(s, kw["merge"], None, kw["max_dur"], kw["preserve"])
            chunks, err = None, "ValueError"
        cases.append({"segments": s, "kw": kw, "proc": proc, "chunks": chunks, "error": err})
    dump_gz(os.path.join(G, "fuzz_cases.json.gz"), cases)

    # aggregate_by_time_interval (preprocessor.py:217-324) through preprocess_transcript(time_interval_seconds=...).
    # The 'speaker' of an interval is a join over a Python set (hash order, differs between processes): the
    # fixture keeps the reference's string and tests compare it as a set of names.
    import contextlib, io
    rng = random.Random(20261107)
    agg = []
    for it in range(90):
        s = fuzzgen.rand_transcript(rng, long_p=0.05)
        kw = dict(interval=rng.choice([1, 5, 10, 30, 60, 300, 7.5, 0.5, 100000, 3]), merge=rng.random() < 0.7,
                  max_dur=rng.choice([1, 30, 120, None]), preserve=rng.random() < 0.8)
        with contextlib.redirect_stdout(io.StringIO()):
            out = rp.preprocess_transcript(s, kw["merge"], kw["interval"], kw["max_dur"], kw["preserve"])
        agg.append({"segments": s, "kw": kw, "agg": out})
    ex_agg = {}
    for iv in (60, 300, 45.5):
        with contextlib.redirect_stdout(io.StringIO()):
            out = rp.preprocess_transcript(segs, True, iv)
        for o in out:
            o["speaker"] = sorted(o["speaker"].split(", "))
        ex_agg[str(iv)] = {"n": len(out), "sha256": sha(out)}
    dump_gz(os.path.join(G, "agg_cases.json.gz"), {"fuzz": agg, "example": ex_agg})
    print("golden written:", os.listdir(G))


if __name__ == "__main__":
    main()
