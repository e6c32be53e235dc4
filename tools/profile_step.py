#!/usr/bin/env python
"""Short driver for ncu: W warm passes + K passes of the hot path on a device-resident batch.
    python tools/profile_step.py [--config 3] [--scale 1.0] [--warm 2] [--steps 1]
"""
import argparse, os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import torch
from llm_map_reduce_summarizer_b200 import engine as E, synth

ap = argparse.ArgumentParser()
ap.add_argument("--config", type=int, default=3)
ap.add_argument("--scale", type=float, default=1.0)
ap.add_argument("--warm", type=int, default=2)
ap.add_argument("--steps", type=int, default=1)
a = ap.parse_args()
eng = E.Engine(0)
b, mt, desc = synth.config(a.config, a.scale)
dev = torch.device("cuda", 0)
dt = {k: torch.from_numpy(getattr(b, k)).to(dev) for k in ("text", "seg_off", "start", "end", "speaker", "doc_off", "spk_bytes", "spk_off")}
ptrs = {k: v.data_ptr() for k, v in dt.items()}
for i in range(a.warm + a.steps):
    eng.preprocess(b, want_merged_text=False, device_ptrs=ptrs)
    eng.chunk(mt)
torch.cuda.synchronize()
print(desc, eng.counts(), "launches", eng.launches())
