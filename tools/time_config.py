#!/usr/bin/env python
"""Device time of one configuration (events around every kernel): python tools/time_config.py --config 4"""
import argparse, os, sys, time
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import torch
from llm_map_reduce_summarizer_b200 import engine as E, synth
ap = argparse.ArgumentParser()
ap.add_argument("--config", type=int, default=4)
ap.add_argument("--scale", type=float, default=1.0)
ap.add_argument("--steps", type=int, default=3)
ap.add_argument("--lexicon", type=int, default=0)
a = ap.parse_args()
eng = E.Engine(0)
t = time.time(); b, mt, desc = synth.config(a.config, a.scale, lexicon=a.lexicon) if a.lexicon else synth.config(a.config, a.scale); print(desc, "generated in %.1fs" % (time.time() - t), b.n_seg, len(b.text))
dev = torch.device("cuda", 0)
dt = {k: torch.from_numpy(getattr(b, k)).to(dev) for k in ("text", "seg_off", "start", "end", "speaker", "doc_off", "spk_bytes", "spk_off")}
ptrs = {k: v.data_ptr() for k, v in dt.items()}
for i in range(2):
    eng.preprocess(b, want_merged_text=False, device_ptrs=ptrs); eng.chunk(mt)
eng.set_profiling(True)
torch.cuda.synchronize(); t = time.time()
for i in range(a.steps):
    eng.preprocess(b, want_merged_text=False, device_ptrs=ptrs); eng.chunk(mt)
torch.cuda.synchronize(); wall = (time.time() - t) / a.steps
prof = eng.profile()
c = eng.counts()
print(c)
print("wall ms/step %.3f  tokens/s %.3e" % (wall * 1e3, c["tokens"] / wall))
for k, v in sorted(prof.items(), key=lambda kv: -kv[1][0])[:int(os.environ.get("TOP", "12"))]:
    print("  %-24s %.3f ms/step (%d launches)" % (k, v[0] / a.steps, v[1] // a.steps))
