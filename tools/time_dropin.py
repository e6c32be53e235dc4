#!/usr/bin/env python
"""Wall time of the drop-in Python API (dict in, dict out) on the example transcript, next to the oracle port."""
import contextlib, gzip, io, json, os, sys, time
ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, ROOT)
from llm_map_reduce_summarizer_b200 import preprocessor as PP, big_chunkeroosky as BC
from oracle import preprocess as OP, chunker as OC
segs = json.load(gzip.open(os.path.join(ROOT, "tests", "golden", "example_segments.json.gz")))
def ours():
    with contextlib.redirect_stdout(io.StringIO()):
        p = PP.preprocess_transcript(segs); c = BC.BigChunkeroosky(); return c.postprocess_chunks(c.chunk_transcript(p))
def oracle():
    p = OP.preprocess(segs); return OC.ChunkerOracle().chunk(p)
for name, f in (("drop-in (GPU)", ours), ("oracle port (CPU)", oracle)):
    f(); f()
    ts = []
    for _ in range(10):
        t = time.perf_counter(); r = f(); ts.append(time.perf_counter() - t)
    print("%-20s median %.2f ms  (%d chunks)" % (name, 1e3 * sorted(ts)[5], len(r)))
