import sys, torch
sys.path.insert(0, ".")
from llm_map_reduce_summarizer_b200 import engine as E, synth
eng = E.Engine(0)
b, mt, d = synth.config(3, 0.25, lexicon=500000)
dev = torch.device("cuda", 0)
dt = {k: torch.from_numpy(getattr(b, k)).to(dev) for k in ("text", "seg_off", "start", "end", "speaker", "doc_off", "spk_bytes", "spk_off")}
ptrs = {k: v.data_ptr() for k, v in dt.items()}
for i in range(2):
    eng.preprocess(b, want_merged_text=False, device_ptrs=ptrs); eng.chunk(mt)
torch.cuda.synchronize(); print(eng.counts())
