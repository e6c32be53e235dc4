#!/usr/bin/env python
"""Bare-copy probe for the end-to-end leg: pinned cudaMemcpyAsync H2D 174 MB + D2H 187 MB per step and
per GPU, nothing else, on 1 / 2 / 4 / 8 GPUs at once (one process per GPU under torchrun, or --gpus N threads of one
process).  Prints GB/s per GPU and in total: what the host can deliver to this many GPUs concurrently, i.e. the ceiling of
bench.py's e2e number.    torchrun --nproc-per-node N tools/copy_probe.py     or     python tools/copy_probe.py --active 1,2,4,8
With --active the world is N = max and the probe repeats with only the first k ranks copying."""
import argparse, json, os, time
import torch
import torch.distributed as dist

ap = argparse.ArgumentParser()
ap.add_argument("--h2d-mb", type=float, default=174.5)
ap.add_argument("--d2h-mb", type=float, default=187.0)
ap.add_argument("--steps", type=int, default=20)
ap.add_argument("--active", default="")
a = ap.parse_args()
rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)
nh, nd = int(a.h2d_mb * 1e6), int(a.d2h_mb * 1e6)
hin, hout = torch.empty(nh, dtype=torch.uint8).pin_memory(), torch.empty(nd, dtype=torch.uint8).pin_memory()
din, dout = torch.empty(nh, dtype=torch.uint8, device=dev), torch.empty(nd, dtype=torch.uint8, device=dev)
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()


def barrier():
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()


def run(mode, active):
    me = rank < active
    barrier()
    t0 = time.perf_counter()
    if me:
        for _ in range(a.steps):
            if mode in ("h2d", "both"):
                with torch.cuda.stream(s1):
                    din.copy_(hin, non_blocking=True)
            if mode in ("d2h", "both"):
                with torch.cuda.stream(s2):
                    hout.copy_(dout, non_blocking=True)
        torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    t = torch.tensor([dt if me else 0.0], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


actives = [int(x) for x in a.active.split(",")] if a.active else [world]
rows = []
for act in actives:
    if act > world:
        continue
    for mode in ("h2d", "d2h", "both"):
        run(mode, act)
        dt = run(mode, act)
        nbytes = ((nh if mode in ("h2d", "both") else 0) + (nd if mode in ("d2h", "both") else 0)) * a.steps
        rows.append({"gpus_copying": act, "mode": mode, "ms_per_step": 1e3 * dt / a.steps, "gb_s_per_gpu": nbytes / dt / 1e9,
                     "gb_s_total": act * nbytes / dt / 1e9})
if rank == 0:
    print(json.dumps({"world": world, "h2d_mb": a.h2d_mb, "d2h_mb": a.d2h_mb, "rows": rows}))
if world > 1:
    dist.destroy_process_group()
