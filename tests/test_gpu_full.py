"""Exhaustive parity at the benchmarked sizes (last paragraph): EVERY document of config 3 (10 000
podcasts, the bench workload), of config 4 (the 1 M-segment transcript with its 1 000 monologues) and of the
config-5 base (the same corpus under mixed 2k/8k/32k limits) is rebuilt from the CUDA result arrays in the
reference's dict format and compared, as canonical JSON, with the oracle run on the same document.

The CUDA pass runs once in this process; the result arrays are then shared (fork) with one worker per host core,
and each worker materialises its documents from the arrays and runs the oracle on them, so the whole comparison
takes about a minute on the GPU box.  Workers never touch CUDA."""
import hashlib, json, multiprocessing as mp, os

import numpy as np
import pytest

from llm_map_reduce_summarizer_b200 import engine as E, synth
from oracle import chunker as OC, preprocess as OP

pytestmark = pytest.mark.gpu
_CTX = {}


class _Specials:                      # build_chunks only needs .specials, and only to raise
    specials = ["<|endoftext|>", "<|fim_prefix|>", "<|fim_middle|>", "<|fim_suffix|>", "<|endofprompt|>"]


def _compare(doc_range):
    b, mt, P, R = _CTX["b"], _CTX["mt"], _CTX["P"], _CTX["R"]
    bad, h = [], hashlib.sha256()
    for x in doc_range:
        segs = synth.doc_to_segments(b, x)
        op = OP.preprocess(segs)
        oc = OC.ChunkerOracle(int(mt[x])).chunk(op)
        pp = E.build_processed(b, P, x)
        got = json.dumps(pp) + json.dumps(E.build_chunks(_Specials, b, P, R, pp, x))
        want = json.dumps(op) + json.dumps(oc)
        if got != want:
            bad.append(x)
        h.update(hashlib.sha256(got.encode()).digest())
    return bad, h.hexdigest(), len(doc_range)


def _every_document(e, b, mt):
    e.preprocess(b)
    e.chunk(mt)
    P, R = e.fetch_processed(), e.fetch_chunks()
    E._text_of(P, "clean"); E._text_of(P, "merged"); E._text_of(R, "twc")      # decoded once, shared by the workers
    b.originals = (b.start.tolist(), b.end.tolist())
    _CTX.update(b=b, mt=mt, P=P, R=R)
    procs = max(1, min(os.cpu_count() or 1, b.n_doc))
    # contiguous ranges balanced by bytes
    cum = b.seg_off[b.doc_off].astype(np.float64)
    cuts = np.searchsorted(cum, np.linspace(0, cum[-1], 4 * procs + 1)[1:-1]).tolist()
    edges = sorted(set([0, b.n_doc] + [int(c) for c in cuts]))
    ranges = [range(a, z) for a, z in zip(edges[:-1], edges[1:]) if z > a]
    if procs == 1:
        results = [_compare(r) for r in ranges]
    else:
        with mp.get_context("fork").Pool(procs) as pool:
            results = pool.map(_compare, ranges)
    _CTX.clear()
    bad = [x for r in results for x in r[0]]
    assert sum(r[2] for r in results) == b.n_doc
    assert not bad, "documents that differ from the oracle: %s" % bad[:20]
    return e.counts()


def test_config3_every_document(gpu_engine):
    b, mt, _ = synth.config(3, 1.0)
    c = _every_document(gpu_engine, b, mt)
    assert c["n_doc"] == 10000 and c["tokens"] > 4e7


def test_config4_every_document(gpu_engine):
    """ONE document: 1 M segments, 1 000 monologues through the sentence / clause / 20-word paths."""
    b, mt, _ = synth.config(4, 1.0)
    c = _every_document(gpu_engine, b, mt)
    assert c["n_seg"] == 1000000 and c["n_sent"] > 100000 and c["n_subs"] > 1000


def test_config5_base_every_document(gpu_engine):
    """The 100 MB base of the 10 GB corpus (replicas of it differ only by seed), limits 2k / 8k / 32k by doc_id % 3."""
    b, mt, _ = synth.config(5, 1.0)
    c = _every_document(gpu_engine, b, mt)
    assert c["n_doc"] == 10000 and sorted(set(mt.tolist())) == [2000, 8000, 32000]
