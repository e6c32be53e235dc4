#!/usr/bin/env python
"""bench.py - transcript tokens/s of preprocess+chunk (BASELINE.json metric) on N B200s.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--config 3] [--scale 1.0]
  python bench.py --impl reference ...     # the reference-equivalent CPU path (oracle port) on host cores

One step = one pass of the whole hot path (trc_preprocess + trc_chunk) over one batch of synthetic
transcripts.  `value` is measured with the inputs already resident in HBM; `e2e` is the same metric
through the C ABI with pinned HOST buffers (H2D of the inputs and D2H of every result array inside the
timed region).  One process per GPU; documents shard across ranks with no collective,
so torch.distributed is only used for the barrier and the max-over-ranks of the device times.
"""
import argparse, json, os, subprocess, sys, threading, time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "transcript tokens/sec preprocess+chunk"
UNIT = "tokens/s"
# every array a caller needs to rebuild the processed segments and the chunks: the cleaned piece texts are
# slices of TWC (PDST + COFF lengths) and so are the merged texts (ITEM_DST; a piece's offset inside its merged
# text is PDST minus that), so neither the cleaned-text buffer nor the merged-text offsets are copied back
# CLEN32 carries the piece lengths (COFF's differences); K2R is the identity when no segment was blank, which the
# counters of the call say, so it is only downloaded when segments were dropped
OUT_ARRAYS = [("K2R", "<i4"), ("CLEN32", "<i4"), ("PDST", "<i8"), ("GSTART", "<i4"), ("GTS", "u1"),
              ("GTOK", "<i8"), ("ITEM_GROUP", "<i4"), ("ITEM_SUB", "<i4"), ("ITEM_TOK", "<i4"), ("ITEM_LEN", "<i8"),
              ("ITEM_DST", "<i8"), ("SUBS", None), ("CHUNKS", None), ("SPK_LIST", "<i4"), ("TWC", "u1"), ("DOC_ERR", "<i4"),
              ("GDOC_OFF", "<i8"), ("CDOC_OFF", "<i8")]


def workload(args, rank):
    from llm_map_reduce_summarizer_b200 import synth
    lex = getattr(args, "lexicon", 0) or None
    b, mt, desc = synth.config(args.config, args.scale, lexicon=lex)
    if rank:
        # weak scaling: every rank gets its own shard of the same shape (different seed)
        saved = synth.BASE_SEED
        synth.BASE_SEED = saved + 1000 * rank
        try:
            b, mt, desc = synth.config(args.config, args.scale, lexicon=lex)
        finally:
            synth.BASE_SEED = saved
    return b, mt, desc


class _DevBatch:
    """What Engine.preprocess needs besides the device pointers."""

    def __init__(self, n_seg, n_doc, speakers):
        self.n_seg, self.n_doc, self.speakers = int(n_seg), int(n_doc), speakers


def resident_ms(torch, eng, hb, mt, dev, steps, warm=3, tensors=None):
    """ms per step of preprocess + chunk with the batch resident on the device (CUDA events on the current stream)."""
    if tensors is None:
        tensors = {k: torch.from_numpy(getattr(hb, k)).to(dev) for k in ("text", "seg_off", "start", "end", "speaker", "doc_off", "spk_bytes", "spk_off")}
    ptrs = {k: v.data_ptr() for k, v in tensors.items()}
    st = torch.cuda.current_stream()

    def step():
        eng.preprocess(hb, want_merged_text=False, stream=st.cuda_stream, device_ptrs=ptrs)
        eng.chunk(mt, stream=st.cuda_stream)

    for _ in range(warm):
        step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(st)
    for _ in range(steps):
        step()
    e1.record(st)
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / steps, eng.counts()


def config5_shard(torch, dev, rank, world, corpus_bytes=10e9):
    """This rank's shard of the 10 GB corpus of config 5 (the 100 MB base tiled to 10 GB at distinct
    addresses): the virtual corpus is the base repeated, shard.shard_ranges cuts it by bytes, and the rank's document
    range is assembled on the device from contiguous slices of the base.  Limits 2k / 8k / 32k by (corpus doc id) % 3."""
    from llm_map_reduce_summarizer_b200 import shard as SH, synth
    base, _, _ = synth.config(5, 1.0)
    nb = base.n_doc
    reps = max(1, int(round(corpus_bytes / base.text.nbytes)))
    doc_bytes = np.diff(base.seg_off[base.doc_off])
    voff = np.concatenate([[0], np.cumsum(np.tile(doc_bytes, reps))])
    x0, x1 = SH.shard_ranges_from_offsets(voff, world)[rank]
    T = {k: torch.from_numpy(getattr(base, k)).to(dev) for k in ("text", "seg_off", "start", "end", "speaker", "doc_off")}
    seg_len = T["seg_off"][1:] - T["seg_off"][:-1]
    doc_segs = T["doc_off"][1:] - T["doc_off"][:-1]
    parts = {k: [] for k in ("text", "seg_len", "start", "end", "speaker", "doc_segs")}
    x = x0
    while x < x1:
        d0 = x % nb
        d1 = min(nb, d0 + (x1 - x))
        s0, s1 = int(base.doc_off[d0]), int(base.doc_off[d1])
        t0, t1 = int(base.seg_off[s0]), int(base.seg_off[s1])
        parts["text"].append(T["text"][t0:t1]); parts["seg_len"].append(seg_len[s0:s1]); parts["start"].append(T["start"][s0:s1])
        parts["end"].append(T["end"][s0:s1]); parts["speaker"].append(T["speaker"][s0:s1]); parts["doc_segs"].append(doc_segs[d0:d1])
        x += d1 - d0
    z = torch.zeros(1, dtype=torch.int64, device=dev)
    out = {"text": torch.cat(parts["text"]), "start": torch.cat(parts["start"]), "end": torch.cat(parts["end"]),
           "speaker": torch.cat(parts["speaker"]),
           "seg_off": torch.cat([z, torch.cumsum(torch.cat(parts["seg_len"]), 0)]),
           "doc_off": torch.cat([z, torch.cumsum(torch.cat(parts["doc_segs"]), 0)]),
           "spk_bytes": torch.from_numpy(base.spk_bytes).to(dev), "spk_off": torch.from_numpy(base.spk_off).to(dev)}
    del T, parts
    mt = np.array([2000, 8000, 32000], dtype=np.int32)[np.arange(x0, x1) % 3]
    hb = _DevBatch(out["start"].numel(), x1 - x0, base.speakers)
    return hb, mt, out, {"corpus_docs": int(nb * reps), "corpus_bytes": int(voff[-1]), "docs": [int(x0), int(x1)]}


def extra_configs(args, torch, dist, eng, dev, rank, world):
    """Driver-visible numbers for the other BASELINE.json configs, measured in the same run (device-resident, CUDA
    events): N = 1: config 2 (one meeting: a latency case), config 4 (one 1 M-segment transcript: does not shard) and an
    open-vocabulary variant of config 3; N > 1: config 5, the 10 GB corpus cut by shard.shard_ranges, max over ranks."""
    from llm_map_reduce_summarizer_b200 import synth
    rec = {}
    if world == 1:
        for num, key in ((2, "config2_one_meeting"), (4, "config4_one_long_transcript")):
            b, mt, desc = synth.config(num, 1.0)
            ms, c = resident_ms(torch, eng, b, mt, dev, steps=5)
            rec[key] = {"workload": desc, "ms_per_step": ms, "tokens_per_s": c["tokens"] / (ms / 1e3), "tokens": c["tokens"],
                        "segments": c["n_seg"], "raw_text_bytes": c["b_raw"], "chunks": c["n_chunks"], "sentences": c["n_sent"],
                        "sub_chunks": c["n_subs"]}
        b, mt, desc = synth.config(3, 0.25, lexicon=500000)
        ms, c = resident_ms(torch, eng, b, mt, dev, steps=5)
        rec["open_vocabulary"] = {"workload": desc + " (scale 0.25)", "ms_per_step": ms, "tokens_per_s": c["tokens"] / (ms / 1e3),
                                  "tokens": c["tokens"], "raw_text_bytes": c["b_raw"], "words": c["n_words"],
                                  "words_first_sighting": c["n_word_first"], "words_generic_matcher": c["n_word_slow"],
                                  "long_spans": c["n_long_spans"]}
        b, mt, desc = synth.config(3, 0.25)
        ms, c = resident_ms(torch, eng, b, mt, dev, steps=5)
        rec["open_vocabulary"]["closed_vocabulary_same_size"] = {"ms_per_step": ms, "tokens_per_s": c["tokens"] / (ms / 1e3), "words": c["n_words"],
                                                                 "words_first_sighting": c["n_word_first"], "words_generic_matcher": c["n_word_slow"]}
    else:
        hb, mt, tensors, info = config5_shard(torch, dev, rank, world)
        ms, c = resident_ms(torch, eng, hb, mt, dev, steps=3, warm=2, tensors=tensors)
        t = torch.tensor([ms, float(c["tokens"]), float(c["b_raw"])], device=dev, dtype=torch.float64)
        tmax = t.clone()
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        rec["config5_corpus"] = dict(info, workload="10 GB corpus sharded by document over %d GPUs, limits 2k/8k/32k by doc id %% 3" % world,
                                     ms_per_step=float(tmax[0].item()), tokens=float(t[1].item()), raw_text_bytes=float(t[2].item()),
                                     tokens_per_s=float(t[1].item()) / (float(tmax[0].item()) / 1e3), scaling="strong (the corpus is fixed)")
        del tensors
        torch.cuda.empty_cache()
    return rec


def dropin_config1(eng):
    """config 1 end to end through the drop-in modules - the example transcript as Python dicts in,
    the reference's dict formats out, wall clock, median of 10 (packing, H2D, kernels, D2H, dict building)."""
    import contextlib, gzip, io
    from llm_map_reduce_summarizer_b200 import big_chunkeroosky as BC, engine as E, preprocessor as PP
    p = os.path.join(ROOT, "tests", "golden", "example_segments.json.gz")
    if not os.path.exists(p):
        return None
    segs = json.load(gzip.open(p))
    E.set_default_engine(eng)

    def once():
        with contextlib.redirect_stdout(io.StringIO()):
            proc = PP.preprocess_transcript(segs)
            c = BC.BigChunkeroosky()
            return c.postprocess_chunks(c.chunk_transcript(proc))

    once(); once()
    ts = []
    for _ in range(10):
        t0 = time.perf_counter()
        r = once()
        ts.append(time.perf_counter() - t0)
    ms = 1e3 * sorted(ts)[5]
    tok = sum(c["token_count"] for c in r)
    return {"workload": "config 1: transcript-example.json (4778 segments) through preprocess_transcript + chunk_transcript + "
                        "postprocess_chunks of the drop-in modules, dicts in and out", "ms": ms, "chunks": len(r), "tokens": tok,
            "tokens_per_s": tok / (ms / 1e3)}


def algorithmic_bytes(c):
    """compulsory traffic of the whole path, each byte once."""
    return c["b_raw"] + 28 * c["n_seg"] + c["b_clean"] + c["b_twc"] + 40 * c["n_groups"] + 64 * c["n_chunks"]


KERNEL_BYTES = {  # algorithmic bytes of one launch of the named kernel
    "k_count_units": lambda c: c["b_clean"] + 16 * c["n_kept"] + 8 * c["n_kept"],
    "k_clean_count": lambda c: c["b_raw"] + 8 * c["n_seg"] + 16 * c["n_seg"],
    "k_clean_write": lambda c: c["b_raw"] + c["b_clean"] + 12 * c["n_kept"],
    "k_emit_chunk_pieces": lambda c: c["b_clean"] + c["b_twc"] + 24 * c["n_kept"],
    "k_emit_merged": lambda c: c["b_clean"] + c["b_merged"] + 24 * c["n_kept"],
    "k_clean_tok_fused": lambda c: c["b_raw"] + c["b_clean"] + 28 * c["n_seg"],
    "k_clean_tok_fused3": lambda c: c["b_raw"] + c["b_clean"] + 28 * c["n_seg"],
    "k_clean_tok_w32": lambda c: c["b_raw"] + c["b_clean"] + 28 * c["n_seg"],
    "k_clean_tok_c32": lambda c: c["b_raw"] + c["b_clean"] + 28 * c["n_seg"],
}


class ClockSampler:
    Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.rows, self.p = [], None
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                                       "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.p = None

    def _read(self):
        for line in self.p.stdout:
            self.rows.append(line.strip().split(", "))

    def stop(self):
        if not self.p:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.p.terminate()
        sm, mx, reasons = [], [], set()
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx.append(float(r[1]))
                for name, v in zip(["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"], r[2:6]):
                    if v.strip() == "Active":
                        reasons.add(name)
            except Exception:
                pass
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def oracle_docs(args_tuple):
    """Worker: the oracle PORT of the path (oracle/preprocess.py + oracle/chunker.py) on a list of documents."""
    docs, mts = args_tuple
    from oracle import chunker as OC, preprocess as OP
    tok = 0
    for segs, mt in zip(docs, mts):
        proc = OP.preprocess(segs)
        ch = OC.ChunkerOracle(int(mt)).chunk(proc)
        tok += sum(c["token_count"] for c in ch)
    return tok


def reference_docs(args_tuple):
    """Worker: the UNMODIFIED reference (oracle/_ref copies of preprocessor.py + big_chunkeroosky.py,
    main.py:139-149's three calls) on a list of documents."""
    docs, mts = args_tuple
    from oracle import refimpl
    return refimpl.run_docs(docs, mts)[0]


def cpu_sample(b, mt, max_docs, max_bytes):
    from llm_map_reduce_summarizer_b200 import synth
    docs, mts, nbytes = [], [], 0
    for x in range(min(b.n_doc, max_docs)):
        docs.append(synth.doc_to_segments(b, x))
        mts.append(int(mt[x]))
        nbytes += int(b.seg_off[b.doc_off[x + 1]] - b.seg_off[b.doc_off[x]])
        if nbytes >= max_bytes:
            break
    return docs, mts, nbytes


def workload_config(args, b, mt, desc, table):
    """The `config` object of the JSON line.  Both arms print exactly this (it only depends on the synthetic
    workload and the rank table, not on what ran), so the driver can tell that they measured the same thing."""
    in_bytes = int(b.text.nbytes + 4 * b.n_seg + 8 * b.n_seg + 8 * b.n_seg + 4 * b.n_seg)
    return {"workload": "config %d: %s" % (args.config, desc), "scale": args.scale,
            "docs_per_gpu": int(b.n_doc), "segments_per_gpu": int(b.n_seg), "raw_text_bytes_per_gpu": int(b.text.nbytes),
            "max_tokens_per_chunk": [int(x) for x in np.unique(np.asarray(mt))],
            "rank_table": table,
            "l2": "inputs larger than L2 (text + segment arrays > 126 MB)" if in_bytes > 126e6
                  else "inputs smaller than L2: numbers are L2-warm",
            "merged_text": "not materialised in the timed step (want_merged_text=0): merged texts are slices of the "
                           "text_with_context stream, which is written in full",
            "parallelism": "documents sharded by rank, no collective"}


def pin_near_gpu(index):
    """Keep this process (and the threads and pinned buffers it creates from here on) on the host cores next to its GPU:
    with one rank per GPU on a two-socket box, a rank on the far socket pays the socket link for every pinned copy and
    every mapped counter read.  A box with one NUMA node lists all cores and nothing changes.  Returns a note for the line."""
    try:
        import pynvml
        pynvml.nvmlInit()
        bus = pynvml.nvmlDeviceGetPciInfo(pynvml.nvmlDeviceGetHandleByIndex(index)).busId
        bus = bus.decode() if isinstance(bus, bytes) else bus
        dom, rest = bus.lower().split(":", 1)
        path = "/sys/bus/pci/devices/%s:%s/local_cpulist" % (dom[-4:], rest)
        cpus = set()
        for part in open(path).read().strip().split(","):
            if part:
                a, _, b = part.partition("-")
                cpus.update(range(int(a), int(b or a) + 1))
        have = os.sched_getaffinity(0)
        want = cpus & have
        if want and want != have:
            os.sched_setaffinity(0, want)
            return "%d cores next to the GPU (of %d)" % (len(want), len(have))
        return "unchanged (%d cores, all local)" % len(have)
    except Exception as e:       # no NVML, no sysfs entry: run where the launcher put us
        return "unchanged (%s)" % type(e).__name__


def table_name():
    from llm_map_reduce_summarizer_b200 import vocab
    return vocab.resolve_table_name()


def pool_rate(fn, docs, mts, procs, warmup, steps):
    """tokens/s of `fn` over the sample with `procs` worker processes (documents dealt round robin)."""
    import multiprocessing as mp
    parts = [(docs[i::procs], mts[i::procs]) for i in range(procs)] if len(docs) >= procs else [(docs, mts)]
    # modules and the rank table are loaded once, before the fork (the GPU arm does not time its vocabulary load either)
    from oracle import tokenizer as TK
    TK.encoding()
    if fn is reference_docs:
        from oracle import refimpl
        refimpl.load()
    else:
        from oracle import chunker, preprocess  # noqa: F401
    ctx = mp.get_context("fork")
    with ctx.Pool(len(parts)) as pool:
        for _ in range(warmup):
            pool.map(fn, parts)
        t0 = time.perf_counter()
        tok = 0
        for _ in range(steps):
            tok = sum(pool.map(fn, parts))
        dt = time.perf_counter() - t0
    return tok * steps / dt, dt / steps, len(parts), tok


def hostsim_rate(b, mt, max_bytes):
    """this repo's own stage code compiled by g++ (tests/hostsim, one host thread) on the sample.
    Test build only - reported as a native-CPU comparator, never loaded by the product."""
    try:
        from llm_map_reduce_summarizer_b200 import build, engine as E, synth, shard as SH
        lib = build.build_hostsim()
        x1, nbytes = 0, 0
        while x1 < b.n_doc and nbytes < max_bytes:
            nbytes += int(b.seg_off[b.doc_off[x1 + 1]] - b.seg_off[b.doc_off[x1]])
            x1 += 1
        sb = synth.sub_batch(b, 0, x1)
        eng = E.Engine(0, lib)
        t0 = time.perf_counter()
        eng.preprocess(sb, want_merged_text=False)
        eng.chunk(np.ascontiguousarray(mt[:x1]))
        dt = time.perf_counter() - t0
        tok = eng.counts()["tokens"]
        eng.close()
        return {"value": tok / dt, "unit": UNIT, "cores": 1, "sample": "%d documents, %.1f MB" % (x1, nbytes / 1e6)}
    except Exception as e:      # the comparator is optional
        return {"unavailable": str(e)[:200]}


def run_reference(args):
    """--impl reference: the reference's own CPU implementation of the path on the box's host cores.
    oracle/_ref holds byte-identical copies of /root/reference/preprocessor.py and big_chunkeroosky.py (staged by
    __graft_entry__.build() in the build container, shipped like the built .so); they run unmodified, one process
    per host core over a stated sample of the SAME synthetic workload, with tiktoken's Rust CoreBPE on the same rank
    table as the GPU arm and the nltk shim for Punkt.  Falls back to the oracle port (kind "port") only when the
    copies are missing."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import refimpl
    b, mt, desc = workload(args, 0)
    cores = os.cpu_count() or 1
    docs, mts, nbytes = cpu_sample(b, mt, 100000, args.cpu_sample_mb * 1e6)
    is_ref = refimpl.available()
    fn = reference_docs if is_ref else oracle_docs
    val, sec, used, tok = pool_rate(fn, docs, mts, cores, args.warmup, args.steps)
    extra = {}
    if not args.no_cpu_baseline:
        # the other CPU legs, each on the same sample, one pass
        other = oracle_docs if is_ref else None
        if other:
            v2, _, u2, _ = pool_rate(other, docs, mts, cores, 0, 1)
            extra["oracle_port"] = {"value": v2, "unit": UNIT, "cores": u2}
        v1, _, _, _ = pool_rate(fn, docs[:max(1, len(docs) // max(cores, 1))], mts[:max(1, len(docs) // max(cores, 1))], 1, 0, 1)
        extra["single_process"] = {"value": v1, "unit": UNIT, "cores": 1}
        extra["own_cpp_build"] = hostsim_rate(b, mt, args.cpu_sample_mb * 1e6 / 4)
    sample = ("%d of %d documents (%.1f MB of %.1f MB raw text, %d tokens) per step, %s, one process per core"
              % (len(docs), b.n_doc, nbytes / 1e6, b.text.nbytes / 1e6, tok,
                 "unmodified reference preprocessor.py + big_chunkeroosky.py (tiktoken CoreBPE, nltk shim)" if is_ref
                 else "oracle port (Python + tiktoken CoreBPE)"))
    line = {"impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * sec, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": workload_config(args, b, mt, desc, table_name()),
            "cpu_baseline": dict({"value": val, "unit": UNIT, "cores": used, "kind": "reference" if is_ref else "port",
                                  "sample": sample}, **extra),
            "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours")
    ap.add_argument("--config", type=int, default=3)
    ap.add_argument("--scale", type=float, default=1.0)
    ap.add_argument("--cpu-sample-mb", type=float, default=12.0)
    ap.add_argument("--lexicon", type=int, default=0, help="draw words from an open vocabulary of this many forms (sensitivity run)")
    ap.add_argument("--no-extra-configs", action="store_true", help="skip the sub-records of configs 2, 4, 5 and the open-vocabulary run")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--e2e-offsets64", action="store_true", help="upload the int64 segment offsets instead of int32 lengths")
    ap.add_argument("--e2e-repeats", type=int, default=3, help="repetitions of the K-step end-to-end leg (median reported)")
    ap.add_argument("--e2e-parts", type=int, default=1, help="sub-batches a step's documents are cut into in the end-to-end leg")
    ap.add_argument("--e2e-contexts", type=int, default=6, help="contexts (host threads) that take batches in turn in the end-to-end leg")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl != "reference" else args.warmup
    if args.impl == "reference":
        return run_reference(args)

    import torch
    import torch.distributed as dist
    from llm_map_reduce_summarizer_b200 import build, engine as E, _native as N
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a B200: the path has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    affinity = pin_near_gpu(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    build.build()
    eng = E.Engine(local)
    b, mt, desc = workload(args, rank)
    dev = torch.device("cuda", local)
    st = torch.cuda.current_stream()
    sp = st.cuda_stream

    def dev_t(a):
        return torch.from_numpy(a).to(dev)

    dt = {k: dev_t(getattr(b, k)) for k in ("text", "seg_off", "start", "end", "speaker", "doc_off", "spk_bytes", "spk_off")}
    ptrs = {k: v.data_ptr() for k, v in dt.items()}

    def step_resident():
        eng.preprocess(b, want_merged_text=False, stream=sp, device_ptrs=ptrs)
        eng.chunk(mt, stream=sp)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        e0.record(st)
        for _ in range(steps):
            fn()
        e1.record(st)
        barrier()
        ms = e0.elapsed_time(e1)
        if world > 1:
            t = torch.tensor([ms], device=dev, dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms

    for _ in range(args.warmup):
        step_resident()
    counts = eng.counts()
    clocks = ClockSampler(local)
    l0 = eng.launches()
    ms = timed(step_resident, args.steps)
    launches = eng.launches() - l0
    clk = clocks.stop()
    tokens_local = counts["tokens"]
    if world > 1:
        t = torch.tensor([tokens_local], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        tokens_all = float(t.item())
    else:
        tokens_all = float(tokens_local)
    value = tokens_all * args.steps / (ms / 1e3)

    # per-kernel device times (events around each launch) -> roofline of the dominant kernel
    eng.set_profiling(True)
    for _ in range(args.steps):
        step_resident()
    prof = eng.profile()
    eng.set_profiling(False)
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    peak_src = "measured (MEASURED_PEAKS.json hbm_gbs)" if "hbm_gbs" in peaks else "fallback 6650 GB/s (B200_PROFILING.md)"
    tot_ms = sum(v[0] for v in prof.values()) or 1.0
    top = max(prof.items(), key=lambda kv: kv[1][0])
    tname, (tms, tn) = top
    kb = KERNEL_BYTES.get(tname, lambda c: 0)(counts)
    per_launch_ms = tms / max(tn, 1)
    ach = kb / (per_launch_ms / 1e3) / 1e9 if per_launch_ms > 0 else 0.0
    A = algorithmic_bytes(counts)
    pipe_gbs = A * args.steps / (ms / 1e3) / 1e9 if world == 1 else None
    traffic = None
    try:        # DRAM bytes of that kernel from the committed ncu --set full capture (same workload only)
        tr = json.load(open(os.path.join(ROOT, "profiles", "traffic.json"))).get(tname)
        if tr and tr.get("config") == args.config and tr.get("scale") == args.scale:
            traffic = tr["dram_bytes_per_launch"]
    except Exception:
        pass
    roofline = {"bound": "hbm", "kernel": tname, "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
                "traffic": traffic, "peak_source": peak_src, "algorithmic_bytes_per_launch": kb,
                "kernel_ms_per_launch": per_launch_ms, "kernel_share_of_step": tms / tot_ms,
                "pipeline": {"algorithmic_bytes_per_step": A, "achieved_gbs": pipe_gbs,
                             "frac": (pipe_gbs / peak) if pipe_gbs else None},
                "kernels_ms_per_step": {k: v[0] / args.steps for k, v in sorted(prof.items(), key=lambda kv: -kv[1][0])}}

    # end to end through the C ABI with host buffers.  The step's documents are independent, so the host side works
    # like a serving loop: the batch is cut into a few sub-batches (by documents, balanced by bytes) and a few
    # contexts, each on its own host thread and stream, take sub-batches in turn - one sub-batch's result download
    # overlaps the next one's upload and kernels.  Every step still uploads all of its inputs from pinned host
    # memory and downloads all of its result arrays.
    from llm_map_reduce_summarizer_b200 import shard as SH, synth
    nctx, nparts = max(1, args.e2e_contexts), max(1, args.e2e_parts)
    bounds = SH.shard_ranges(b, nparts)
    parts = []
    for (x0, x1) in bounds:
        sb = synth.sub_batch(b, x0, x1)
        pin = {k: torch.from_numpy(np.ascontiguousarray(getattr(sb, k))).pin_memory()
               for k in ("text", "seg_off", "start", "end", "speaker", "doc_off")}
        # the segment lengths go up as int32 (trc_batch.seg_len32) instead of the int64 offsets
        pin["seg_len32"] = torch.from_numpy(np.diff(sb.seg_off).astype(np.int32)).pin_memory()
        parts.append((E.PackedBatch(pin["text"].numpy(), pin["seg_off"].numpy(), pin["start"].numpy(), pin["end"].numpy(),
                                    pin["speaker"].numpy(), pin["doc_off"].numpy(), b.speakers, None),
                      np.ascontiguousarray(mt[x0:x1]), pin))
    workers = [(eng if wi == 0 else E.Engine(local), torch.cuda.Stream()) for wi in range(nctx)]
    pin_outs = [dict() for _ in range(nctx)]
    qlock = threading.Lock()

    def e2e_worker(wi, state, res, own):
        en, stream = workers[wi]
        torch.cuda.set_device(local)
        po = pin_outs[wi]
        while True:
            with qlock:
                if own is not None:              # warm-up: a fixed number of batches for THIS context
                    if own[0] <= 0:
                        return
                    own[0] -= 1
                    item = own[0]
                else:
                    item = state[0]
                    if item >= state[1]:
                        return
                    state[0] = item + 1
            hb_, mt_, pin_ = parts[item % nparts]
            en.preprocess(hb_, want_merged_text=False, stream=stream.cuda_stream, seg_len32=None if args.e2e_offsets64 else pin_["seg_len32"].numpy())
            en.chunk(mt_, stream=stream.cuda_stream)
            cn = en.counts()
            names = [name for name, _ in OUT_ARRAYS if not (name == "K2R" and cn["n_kept"] == cn["n_seg"])]
            _, nb, slab = en.arrays_into(names, None, 0)            # layout only
            if "slab" not in po or po["slab"].numel() < slab:
                po["slab"] = torch.empty(slab + slab // 8 + 4096, dtype=torch.uint8).pin_memory()
            # every result array of the step with ONE call: a copy per array into the pinned slab, one synchronisation
            en.arrays_into(names, po["slab"].data_ptr(), po["slab"].numel(), N.TRC_MEM_HOST, stream.cuda_stream)
            total = int(nb.sum())
            with qlock:
                res[item % nparts] = total

    def step_e2e_pair(nsteps, per_context=None):
        res, state = [0] * nparts, [0, nsteps * nparts]
        ts = [threading.Thread(target=e2e_worker, args=(wi, state, res, None if per_context is None else [per_context]))
              for wi in range(nctx)]
        for t in ts:
            t.start()
        for t in ts:
            t.join()
        return sum(res)

    # warm-up: every context needs three batches of its own before its arena is one block (grow, fuse, steady);
    # cudaMalloc / cudaFree inside the timed region would stall all contexts
    d2h = step_e2e_pair(0, per_context=3 * nparts)
    # K steps, timed with the host clock (the work runs on the workers' own streams), max over ranks; repeated a few
    # times because the PCIe leg shares the host with other tenants: the median run is reported, all runs are listed
    runs = []
    for _ in range(max(1, args.e2e_repeats)):
        barrier()
        t0 = time.perf_counter()
        step_e2e_pair(args.steps)
        torch.cuda.synchronize()
        barrier()
        ms_run = (time.perf_counter() - t0) * 1e3
        if world > 1:
            t = torch.tensor([ms_run], device=dev, dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms_run = float(t.item())
        runs.append(ms_run)
    ms_e2e = sorted(runs)[len(runs) // 2]
    h2d = int(sum(v.numel() * v.element_size() for _, _, pin in parts for k, v in pin.items() if k != "seg_off")
              + nparts * (b.spk_bytes.nbytes + b.spk_off.nbytes))
    e2e = {"value": tokens_all * args.steps / (ms_e2e / 1e3), "unit": UNIT, "h2d_bytes_per_step": h2d,
           "d2h_bytes_per_step": int(d2h), "ms_per_step": ms_e2e / args.steps,
           "ms_per_step_runs": [r / args.steps for r in runs],
           "api": "trc_preprocess + trc_chunk + trc_get_arrays (C ABI, pinned host buffers, one download call per step); every step's documents go "
                  "through as %d sub-batches taken in turn by %d contexts on %d host threads; timed with the host "
                  "clock around K steps" % (nparts, nctx, nctx)}

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        # the reference's own code (oracle/_ref, unmodified) on ONE host core over a bounded sample of this workload
        from oracle import refimpl
        docs, mts, nbytes = cpu_sample(b, mt, 100000, args.cpu_sample_mb * 1e6)
        is_ref = refimpl.available()
        if is_ref:
            refimpl.load()
            refimpl.run_docs(docs[:2], mts[:2])
        t0 = time.perf_counter()
        tok = reference_docs((docs, mts)) if is_ref else oracle_docs((docs, mts))
        dtc = time.perf_counter() - t0
        cpu = {"value": tok / dtc, "unit": UNIT, "cores": 1, "kind": "reference" if is_ref else "port",
               "sample": "first %d documents (%.1f MB raw text) of the workload, %.1f s, %s"
                         % (len(docs), nbytes / 1e6, dtc, "unmodified reference preprocessor.py + big_chunkeroosky.py "
                            "(tiktoken CoreBPE, nltk shim)" if is_ref else "oracle port (Python + tiktoken CoreBPE)")}

    extra = None
    if not args.no_extra_configs and args.config == 3 and not args.lexicon:
        for w_eng, _ in workers[1:]:
            w_eng.close()
        torch.cuda.empty_cache()
        extra = extra_configs(args, torch, dist if world > 1 else None, eng, dev, rank, world)
        if rank == 0 and world == 1:
            extra["config1_dropin"] = dropin_config1(eng)

    if rank == 0:
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "u8", "data": "synthetic",
                "config": workload_config(args, b, mt, desc, eng.vocab_name),
                "counts": {"tokens_per_gpu_per_step": int(tokens_local), "chunks_per_gpu": counts["n_chunks"],
                           "merged_segments_per_gpu": counts["n_groups"], "algorithmic_bytes_per_step": int(A)},
                "clocks": clk, "e2e": e2e, "gpu_launches": int(launches), "roofline": roofline}
        if cpu:
            line["cpu_baseline"] = cpu
        if extra:
            line["configs"] = extra
        line["cpu_affinity"] = affinity
        line["counts"].update({"words": counts["n_words"], "words_first_sighting": counts["n_word_first"],
                               "words_generic_matcher": counts["n_word_slow"]})
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
