#!/usr/bin/env python
"""Determinism soak: the same batch through the device path several times (fresh context each round) must give
byte-identical results - tile scheduling, look-back order, memo races and atomics do not leak into the output."""
import hashlib, os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from llm_map_reduce_summarizer_b200 import engine as E, synth
hs = set()
for cfg, scale in ((3, 0.3), (4, 0.1), (5, 0.2)):
    b, mt, desc = synth.config(cfg, scale)
    seen = set()
    for rnd in range(3):
        e = E.Engine(0)
        for k in range(4):
            e.preprocess(b, want_merged_text=False); e.chunk(mt)
            seen.add(hashlib.sha256(b"".join(e.array(k2, "u1").tobytes() for k2 in ("TWC", "CTOK", "CHUNKS", "SUBS", "COFF"))).hexdigest())
        e.close()
    print(desc, "scale", scale, "- distinct results over 12 runs:", len(seen))
    assert len(seen) == 1
print("soak ok")
