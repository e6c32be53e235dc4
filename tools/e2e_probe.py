#!/usr/bin/env python
"""Where the end-to-end time goes: H2D alone, kernels alone, D2H alone, then N contexts in turn.
python tools/e2e_probe.py [--contexts 3] [--steps 48]"""
import argparse, os, sys, threading, time
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import torch
from llm_map_reduce_summarizer_b200 import engine as E, synth, _native as N
import bench as B
ap = argparse.ArgumentParser()
ap.add_argument("--contexts", type=int, default=3)
ap.add_argument("--steps", type=int, default=48)
a = ap.parse_args()
b, mt, desc = synth.config(3, 1.0)
dev = torch.device("cuda", 0)
keys = ("text", "seg_off", "start", "end", "speaker", "doc_off")
pin = {k: torch.from_numpy(getattr(b, k)).pin_memory() for k in keys}
dd = {k: torch.empty_like(v, device=dev) for k, v in pin.items()}
def ev():
    return torch.cuda.Event(enable_timing=True)
def timeit(fn, n=10):
    fn(); torch.cuda.synchronize()
    e0, e1 = ev(), ev(); e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n
h2d_bytes = sum(v.numel() * v.element_size() for v in pin.values())
t = timeit(lambda: [dd[k].copy_(pin[k], non_blocking=True) for k in keys])
print("H2D %.1f MB: %.3f ms = %.1f GB/s" % (h2d_bytes / 1e6, t, h2d_bytes / t / 1e6))
eng = E.Engine(0)
hb = E.PackedBatch(*[pin[k].numpy() for k in keys], b.speakers, None)
eng.preprocess(hb, want_merged_text=False); eng.chunk(mt)
sizes = {name: eng.array_into(name, None, 0) for name, _ in B.OUT_ARRAYS}
print("D2H arrays MB:", {k: round(v / 1e6, 2) for k, v in sizes.items() if v > 1e5})
po = {k: torch.empty(max(v, 1), dtype=torch.uint8).pin_memory() for k, v in sizes.items()}
def d2h():
    for k, v in sizes.items():
        if v: eng.array_into(k, po[k].data_ptr(), po[k].numel(), N.TRC_MEM_HOST, 0)
t = timeit(d2h)
tot = sum(sizes.values())
print("D2H %.1f MB: %.3f ms = %.1f GB/s" % (tot / 1e6, t, tot / t / 1e6))
big = torch.empty(tot, dtype=torch.uint8, device=dev); pbig = torch.empty(tot, dtype=torch.uint8).pin_memory()
t = timeit(lambda: pbig.copy_(big, non_blocking=True))
print("D2H one copy of the same size: %.3f ms = %.1f GB/s" % (t, tot / t / 1e6))
t0 = time.perf_counter()
for _ in range(10):
    eng.preprocess(hb, want_merged_text=False); eng.chunk(mt)
torch.cuda.synchronize()
print("H2D + kernels through the ABI: %.3f ms/step" % ((time.perf_counter() - t0) * 100))
for nctx in sorted({1, 2, a.contexts, a.contexts + 1}):
    ws = [(eng if i == 0 else E.Engine(0), torch.cuda.Stream(), {k: torch.empty(max(v, 1) + v // 8, dtype=torch.uint8).pin_memory() for k, v in sizes.items()}) for i in range(nctx)]
    def worker(i, n):
        en, s, p = ws[i]
        torch.cuda.set_device(0)
        for _ in range(n):
            en.preprocess(hb, want_merged_text=False, stream=s.cuda_stream); en.chunk(mt, stream=s.cuda_stream)
            for k, _ in B.OUT_ARRAYS:
                nb = en.array_into(k, None, 0)
                if nb: en.array_into(k, p[k].data_ptr(), p[k].numel(), N.TRC_MEM_HOST, s.cuda_stream)
    def run(n):
        ts = [threading.Thread(target=worker, args=(i, (n + nctx - 1 - i) // nctx)) for i in range(nctx)]
        [t.start() for t in ts]; [t.join() for t in ts]; torch.cuda.synchronize()
    run(2 * nctx)
    t0 = time.perf_counter(); run(a.steps); dt = (time.perf_counter() - t0) * 1e3 / a.steps
    print("%d contexts, %d steps: %.3f ms/step" % (nctx, a.steps, dt))
    del ws
