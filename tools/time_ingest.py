#!/usr/bin/env python
"""Ingest: trc_ingest_json against json.loads + engine.pack_documents on a synthetic corpus serialised as JSON
(host-only; uses the CUDA library when present, else the host test build).
    python tools/time_ingest.py [--scale 0.1] [--threads N]"""
import argparse, json, os, sys, time
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from llm_map_reduce_summarizer_b200 import _native as N, build, engine as E, synth
ap = argparse.ArgumentParser()
ap.add_argument("--scale", type=float, default=0.1)
ap.add_argument("--threads", type=int, default=os.cpu_count())
a = ap.parse_args()
lib = N.load_library(N.DEFAULT_LIB if os.path.exists(N.DEFAULT_LIB) else build.build_hostsim())
b, mt, desc = synth.config(3, a.scale)
docs = [json.dumps({"segments": synth.doc_to_segments(b, x)}).encode() for x in range(b.n_doc)]
nbytes = sum(map(len, docs))
print(desc, "scale", a.scale, "-", b.n_doc, "documents,", b.n_seg, "segments, %.1f MB of JSON" % (nbytes / 1e6))
t = time.perf_counter(); ref = E.pack_documents([json.loads(d)["segments"] for d in docs]); t_ref = time.perf_counter() - t
for nt in (1, a.threads):
    t = time.perf_counter(); got = E.ingest_json(docs, n_threads=nt, lib=lib); dt = time.perf_counter() - t
    assert got.text.tobytes() == ref.text.tobytes() and (got.start == ref.start).all() and (got.speaker == ref.speaker).all()
    print("trc_ingest_json, %2d threads: %.3f s = %.0f MB/s   (json.loads + pack_documents: %.3f s = %.0f MB/s, x%.1f)"
          % (nt, dt, nbytes / dt / 1e6, t_ref, nbytes / t_ref / 1e6, t_ref / dt))
