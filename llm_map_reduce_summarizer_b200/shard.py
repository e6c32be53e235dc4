"""Document sharding across ranks: contiguous document ranges balanced by raw text
bytes.  Documents are independent, so ranks never exchange data; the only cross-rank step is summing
throughput counters."""
import numpy as np

from . import synth


def shard_ranges(batch, world):
    """[(doc_begin, doc_end)] per rank, contiguous, balanced by bytes of raw text."""
    return shard_ranges_from_offsets(batch.seg_off[batch.doc_off], world)


def shard_ranges_from_offsets(doc_byte_off, world):
    """The same from the byte offset of every document start (n_doc + 1 entries)."""
    seg_bytes = np.asarray(doc_byte_off)
    n_doc = len(seg_bytes) - 1
    total = int(seg_bytes[-1])
    cuts = [0]
    for r in range(1, world):
        target = total * r // world
        cuts.append(int(np.searchsorted(seg_bytes, target, side="left")))
    cuts.append(n_doc)
    cuts = [min(max(c, cuts[i - 1] if i else 0), n_doc) for i, c in enumerate(cuts)]
    return [(cuts[r], cuts[r + 1]) for r in range(world)]


def rank_batch(batch, rank, world):
    a, b = shard_ranges(batch, world)[rank]
    return synth.sub_batch(batch, a, b), (a, b)


class MultiGpu:
    """One process, several GPUs: the documents of a batch are cut into contiguous ranges balanced by
    bytes, and every GPU runs the identical single-GPU pipeline on its range from its own host thread, context and
    stream.  No collective, no peer access; results stay per shard (documents keep their order: shard r holds
    documents [ranges[r][0], ranges[r][1]))."""

    def __init__(self, devices=None, lib_path=None, tokenizer_name="cl100k_base", rank_table=None):
        from . import engine as E
        if devices is None:
            import torch
            devices = list(range(torch.cuda.device_count()))
        if not devices:
            raise E.TrcError("no CUDA device: this library has no CPU path")
        self.devices = list(devices)
        self.engines = [E.Engine(d, lib_path, tokenizer_name, rank_table) for d in self.devices]
        self.ranges = None

    def run(self, batch, doc_max_tokens, fetch=False, **preprocess_kw):
        """preprocess + chunk of every shard, concurrently.  Returns one dict per shard: its document range, counters
        and, with fetch=True, the stage arrays (Engine.fetch_processed / fetch_chunks)."""
        import threading
        world = len(self.engines)
        self.ranges = shard_ranges(batch, world)
        mt = np.asarray(doc_max_tokens)
        out, errs = [None] * world, []

        def work(r):
            try:
                x0, x1 = self.ranges[r]
                eng = self.engines[r]
                res = {"docs": (x0, x1), "device": self.devices[r]}
                if x1 > x0:
                    sb = synth.sub_batch(batch, x0, x1)
                    eng.preprocess(sb, **preprocess_kw)
                    eng.chunk(np.ascontiguousarray(mt[x0:x1]))
                    res["counts"] = eng.counts()
                    if fetch:
                        res["batch"], res["processed"], res["chunks"] = sb, eng.fetch_processed(), eng.fetch_chunks()
                out[r] = res
            except Exception as e:      # re-raised in the caller's thread
                errs.append(e)

        ts = [threading.Thread(target=work, args=(r,)) for r in range(world)]
        for t in ts:
            t.start()
        for t in ts:
            t.join()
        if errs:
            raise errs[0]
        return out

    def close(self):
        for e in self.engines:
            e.close()
