"""N>1 path on CPU: two gloo ranks shard one corpus by document, run the pipeline on their shards (host
test build of the stages) and the union of their results equals the single-rank result."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from llm_map_reduce_summarizer_b200 import build, engine as E, shard, synth


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _worker(rank, world, port, lib, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    b, mt, _ = synth.config(5, 0.004)
    sb, (a, z) = shard.rank_batch(b, rank, world)
    e = E.Engine(0, lib)
    e.preprocess(sb, want_merged_text=False)
    e.chunk(mt[a:z])
    R = e.fetch_chunks()
    per_doc = np.diff(R["cdoc_off"]).tolist()
    tok = torch.tensor([e.counts()["tokens"]], dtype=torch.int64)
    dist.all_reduce(tok)                                   # the only cross-rank step: summing counters
    gathered = [None] * world
    dist.all_gather_object(gathered, (a, z, per_doc, R["twc"].tobytes()))
    if rank == 0:
        out.put((int(tok.item()), gathered))
    dist.barrier()
    dist.destroy_process_group()


def test_two_ranks_match_single_rank():
    lib = build.build_hostsim()
    b, mt, _ = synth.config(5, 0.004)
    e = E.Engine(0, lib)
    e.preprocess(b, want_merged_text=False)
    e.chunk(mt)
    ref_tokens = e.counts()["tokens"]
    R = e.fetch_chunks()
    ref_per_doc = np.diff(R["cdoc_off"]).tolist()
    ref_twc = R["twc"].tobytes()
    ranges = shard.shard_ranges(b, 2)
    assert ranges[0][0] == 0 and ranges[-1][1] == b.n_doc and ranges[0][1] == ranges[1][0]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, lib, q)) for r in range(2)]
    for p in procs:
        p.start()
    tokens, gathered = q.get(timeout=300)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert tokens == ref_tokens
    gathered.sort()
    assert sum((g[2] for g in gathered), []) == ref_per_doc
    assert b"".join(g[3] for g in gathered) == ref_twc
