#!/usr/bin/env python
"""Device time of trc_aggregate on a configuration: python tools/time_aggregate.py --config 3 --interval 60"""
import argparse, os, sys, time
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import torch
from llm_map_reduce_summarizer_b200 import engine as E, synth
ap = argparse.ArgumentParser()
ap.add_argument("--config", type=int, default=3)
ap.add_argument("--scale", type=float, default=1.0)
ap.add_argument("--interval", type=float, default=60)
a = ap.parse_args()
eng = E.Engine(0)
b, mt, desc = synth.config(a.config, a.scale)
eng.preprocess(b, want_merged_text=False)
for _ in range(2):
    eng.aggregate(a.interval)
eng.set_profiling(True)
torch.cuda.synchronize(); t = time.time()
for _ in range(3):
    eng.aggregate(a.interval)
torch.cuda.synchronize(); wall = (time.time() - t) / 3
c = eng.counts()
print(desc, "interval", a.interval, {k: c[k] for k in ("n_groups", "n_agg", "n_agg_mem", "b_agg")}, "wall ms %.3f" % (wall * 1e3))
for k, v in sorted(eng.profile().items(), key=lambda kv: -kv[1][0])[:10]:
    print("  %-20s %.3f ms (%d launches)" % (k, v[0] / 3, v[1] // 3))
