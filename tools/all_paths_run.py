#!/usr/bin/env python
"""Small end-to-end pass over every path of the device code (fused kernel incl. slow lane and giants, sentence path,
aggregation, token service, both other rank tables) - meant for a memory checker where one is available
(compute-sanitizer is closed on the pool this was developed on; the overflow paths are covered by
tests/test_gpu_parity.py::test_dense_tiles and ::test_other_rank_tables against the oracle instead)."""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests"))
import numpy as np
from llm_map_reduce_summarizer_b200 import engine as E, synth

eng = E.Engine(0)
for num, scale in ((3, 0.004), (5, 0.003), (2, 0.02), (4, 0.002)):
    b, mt, desc = synth.config(num, scale)
    eng.preprocess(b)
    eng.chunk(mt)
    c = eng.counts()
    eng.aggregate(45)
    print(desc, {k: c[k] for k in ("n_seg", "n_groups", "n_chunks", "tokens", "n_sent", "n_giants")}, eng.counts()["n_agg"])
cnt, sp = eng.count_tokens(["hello world", "a" * 300, "<|endoftext|> x", "naïve café 你好", ""])
print(cnt.tolist(), sp.tolist())
import parity_common as pc
for name, ranks in pc.small_rank_tables():
    toks = sorted(ranks, key=lambda t: ranks[t])
    e2 = E.Engine(0, rank_table=(name, toks, [ranks[t] for t in toks]))
    b, mt, desc = synth.config(3, 0.003)
    e2.preprocess(b)
    e2.chunk(mt)
    print(name, e2.counts()["tokens"])
    e2.close()
print("sanitize run done")
