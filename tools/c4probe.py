import sys, time, os
sys.path.insert(0, '/root/repo')
import torch
from llm_map_reduce_summarizer_b200 import engine as E, synth
scale = float(sys.argv[1])
eng = E.Engine(0)
t = time.time(); b, mt, desc = synth.config(4, scale); print("gen %.1fs" % (time.time() - t), b.n_seg, len(b.text), flush=True)
eng.set_profiling(True)
t = time.time(); eng.preprocess(b, want_merged_text=False); torch.cuda.synchronize(); print("preprocess %.3fs" % (time.time() - t), flush=True)
print({k: round(v[0], 3) for k, v in sorted(eng.profile().items(), key=lambda kv: -kv[1][0])[:6]}, flush=True)
t = time.time(); eng.chunk(mt); torch.cuda.synchronize(); print("chunk %.3fs" % (time.time() - t), flush=True)
print({k: round(v[0], 3) for k, v in sorted(eng.profile().items(), key=lambda kv: -kv[1][0])[:10]}, flush=True)
print(eng.counts(), flush=True)
