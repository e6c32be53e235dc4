#!/usr/bin/env python
"""Device-resident time of every BASELINE.json configuration (CUDA events around K steps, inputs resident in HBM):
    python tools/bench_configs.py > profiles/rNN_configs.json
config 2 is ONE 10k-segment meeting (a latency case: ~0.4 MB of text); it is also run as a batch of reseeded
instances.  config 5 is run at the size of one GPU's shard of the 10 GB corpus at 8 GPUs (1.2 GB of text)."""
import json, os, sys, time
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import numpy as np
import torch
from llm_map_reduce_summarizer_b200 import engine as E, synth

eng = E.Engine(0)
dev = torch.device("cuda", 0)
out = []


def run(name, b, mt, steps=5):
    dt = {k: torch.from_numpy(getattr(b, k)).to(dev) for k in ("text", "seg_off", "start", "end", "speaker", "doc_off", "spk_bytes", "spk_off")}
    ptrs = {k: v.data_ptr() for k, v in dt.items()}
    for _ in range(3):
        eng.preprocess(b, want_merged_text=False, device_ptrs=ptrs); eng.chunk(mt)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        eng.preprocess(b, want_merged_text=False, device_ptrs=ptrs); eng.chunk(mt)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    c = eng.counts()
    rec = {"config": name, "docs": b.n_doc, "segments": c["n_seg"], "raw_text_mb": round(c["b_raw"] / 1e6, 2), "tokens": c["tokens"],
           "chunks": c["n_chunks"], "sentence_records": c["n_sent"], "ms_per_step": round(ms, 3),
           "tokens_per_s": round(c["tokens"] / (ms / 1e3)), "raw_gb_per_s": round(c["b_raw"] / ms / 1e6, 2)}
    out.append(rec)
    print(json.dumps(rec), file=sys.stderr)
    del dt
    torch.cuda.empty_cache()


b, mt, d = synth.config(2, 1.0); run("2: " + d + " (one document)", b, mt, 20)
bb = synth.make_corpus(512, synth.BASE_SEED + 2, segs_per_doc=10000, words_lambda=3.0, n_speakers=4, turn_mean=8.0, sec_per_word=0.09)
run("2 x 512: 512 reseeded 1-hour meetings in one batch", bb, np.full(512, 4000, dtype=np.int32), 5)
del bb
b, mt, d = synth.config(3, 1.0); run("3: " + d, b, mt, 10)
b, mt, d = synth.config(4, 1.0); run("4: " + d, b, mt, 5)
b, mt, d = synth.config(5, 10.0); run("5: " + d + " (one GPU's 1.2 GB shard of the 10 GB corpus)", b, mt, 3)
print(json.dumps(out, indent=1))
