"""Document sharding across ranks: contiguous document ranges balanced by raw text
bytes.  Documents are independent, so ranks never exchange data; the only cross-rank step is summing
throughput counters."""
import numpy as np

from . import synth


def shard_ranges(batch, world):
    """[(doc_begin, doc_end)] per rank, contiguous, balanced by bytes of raw text."""
    seg_bytes = batch.seg_off[batch.doc_off]                # byte offset at every document start
    total = int(seg_bytes[-1])
    cuts = [0]
    for r in range(1, world):
        target = total * r // world
        cuts.append(int(np.searchsorted(seg_bytes, target, side="left")))
    cuts.append(batch.n_doc)
    cuts = [min(max(c, cuts[i - 1] if i else 0), batch.n_doc) for i, c in enumerate(cuts)]
    return [(cuts[r], cuts[r + 1]) for r in range(world)]


def rank_batch(batch, rank, world):
    a, b = shard_ranges(batch, world)[rank]
    return synth.sub_batch(batch, a, b), (a, b)
