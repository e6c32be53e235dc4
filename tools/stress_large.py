#!/usr/bin/env python
"""One large shard (config 3 x SCALE, default 10 = 1.2 GB raw text, 21.7 M segments) through the device
path with size-independent invariants and a per-document comparison of a sample against the oracle.
    python tools/stress_large.py [--scale 10]"""
import argparse, json, os, sys, time
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import numpy as np
import torch
from llm_map_reduce_summarizer_b200 import engine as E, synth
from oracle import chunker as OC, preprocess as OP

ap = argparse.ArgumentParser(); ap.add_argument("--scale", type=float, default=10.0); a = ap.parse_args()
t = time.time(); b, mt, desc = synth.config(3, a.scale); print("generated %.1fs" % (time.time() - t), b.n_doc, b.n_seg, len(b.text), flush=True)
e = E.Engine(0)
for it in range(2):
    t = time.time(); e.preprocess(b, want_merged_text=False); e.chunk(mt); torch.cuda.synchronize(); print("step %.3fs" % (time.time() - t), flush=True)
c = e.counts(); print(c, flush=True)
ch = e.array("CHUNKS", [("item0", "<i4"), ("item1", "<i4"), ("doc", "<i4"), ("chunk_index", "<i4"), ("token_count", "<i8"), ("start_time", "<f8"), ("end_time", "<f8"), ("position_percentage", "<f8"), ("pct_is_int_zero", "<i4"), ("start_kind", "<i4"), ("start_ref", "<i4"), ("n_speakers", "<i4"), ("hdr_len", "<i4"), ("err", "<i4"), ("text_len", "<i8"), ("twc_off", "<i8")])
assert (ch["item0"][1:] == ch["item1"][:-1]).all() and ch["item0"][0] == 0 and ch["item1"][-1] == c["n_items"]
assert int(ch["token_count"].sum()) == c["tokens"]
hdr = np.where(ch["hdr_len"] > 0, ch["hdr_len"] + 2, 0)
assert (ch["twc_off"][1:] == (ch["twc_off"] + hdr + ch["text_len"])[:-1]).all()
assert int((ch["twc_off"] + hdr + ch["text_len"])[-1]) == c["b_twc"]
over = ch["token_count"] > (mt - 150)[ch["doc"]]
assert ((ch["item1"] - ch["item0"])[over] == 1).all()
print("invariants ok", flush=True)
# a sample of documents against the oracle (re-run on the sample only; whole-batch arrays are large)
for x in (0, b.n_doc // 2, b.n_doc - 1):
    sb = synth.sub_batch(b, x, x + 1)
    e.preprocess(sb); e.chunk(mt[x:x + 1])
    P, R = e.fetch_processed(), e.fetch_chunks()
    segs = synth.doc_to_segments(sb, 0)
    op = OP.preprocess(segs); oc = OC.ChunkerOracle(int(mt[x])).chunk(op)
    pp = E.build_processed(sb, P, 0)
    assert json.dumps(pp) == json.dumps(op) and json.dumps(E.build_chunks(e, sb, P, R, pp, 0)) == json.dumps(oc), x
print("sample documents match the oracle", flush=True)
# the same documents taken out of the WHOLE batch's result arrays (offsets beyond 2^31 for scale >= 20)
t = time.time(); e.preprocess(b, want_merged_text=True); e.chunk(mt)
P, R = e.fetch_processed(), e.fetch_chunks()
print("whole batch fetched in %.1fs (TWC %.2f GB)" % (time.time() - t, len(R["twc"]) / 1e9), flush=True)
for x in (0, b.n_doc // 3, b.n_doc // 2, b.n_doc - 2, b.n_doc - 1):
    segs = synth.doc_to_segments(b, x)
    op = OP.preprocess(segs); oc = OC.ChunkerOracle(int(mt[x])).chunk(op)
    pp = E.build_processed(b, P, x)
    assert json.dumps(pp) == json.dumps(op), x
    assert json.dumps(E.build_chunks(e, b, P, R, pp, x)) == json.dumps(oc), x
print("documents inside the whole batch match the oracle (last chunk at byte %d of TWC)" % int(R["chunks"]["twc_off"][-1]), flush=True)
