"""Shared comparison helpers: run an Engine (CUDA or hostsim) through the drop-in modules and compare
with golden vectors / the oracle."""
import json

from conftest import quiet
from llm_map_reduce_summarizer_b200 import engine as E, preprocessor as PP, big_chunkeroosky as BC


def run_dropin(eng, segments, max_tokens=4000, merge=True, max_dur=120, preserve=True, add_context=True):
    E.set_default_engine(eng)
    with quiet():
        proc = PP.preprocess_transcript(segments, merge, None, max_dur, preserve)
        c = BC.BigChunkeroosky(max_tokens_per_chunk=max_tokens)
        chunks = c.postprocess_chunks(c.chunk_transcript(proc, add_context))
    return proc, chunks


def same(a, b):
    return json.dumps(a) == json.dumps(b)


def first_diff(a, b):
    if a is None or b is None or len(a) != len(b):
        return "len %s vs %s" % (None if a is None else len(a), None if b is None else len(b))
    for i, (x, y) in enumerate(zip(a, b)):
        if json.dumps(x) != json.dumps(y):
            for k in x:
                if json.dumps(x[k]) != json.dumps(y.get(k)):
                    return "item %d key %s: %r != %r" % (i, k, repr(x[k])[:200], repr(y.get(k))[:200])
            return "item %d differs" % i
    return ""


def check_fuzz_cases(eng, cases, require_all=True, foreign=False):
    """foreign=True: chunk_transcript gets a plain deep copy of the REFERENCE's processed list (a list
    this package did not produce), exercising the already-processed entry of the C ABI."""
    import copy
    n_ok = 0
    for i, cs in enumerate(cases):
        kw = cs["kw"]
        E.set_default_engine(eng)
        if foreign:
            proc = copy.deepcopy(cs["proc"])
        else:
            with quiet():
                proc = PP.preprocess_transcript(cs["segments"], kw["merge"], None, kw["max_dur"], kw["preserve"])
            assert same(list(proc), cs["proc"]), "case %d preprocess: %s" % (i, first_diff(list(proc), cs["proc"]))
        err, chunks = None, None
        try:
            with quiet():
                c = BC.BigChunkeroosky(max_tokens_per_chunk=kw["max_tokens"])
                chunks = c.postprocess_chunks(c.chunk_transcript(proc, kw["add_context"]))
        except ValueError:
            err = "ValueError"
        except NotImplementedError:
            if require_all:
                raise
            continue
        assert err == cs["error"], "case %d error %r != %r" % (i, err, cs["error"])
        assert same(chunks, cs["chunks"]), "case %d chunks: %s" % (i, first_diff(chunks, cs["chunks"]))
        n_ok += 1
    return n_ok


def sentence_fuzz(eng, n_cases, seed, giant=False):
    """Over-budget un-merged segments: Punkt sentences, clause splits, 20-word splits, nested greedy.
    giant=True: segments of 5-10 KB, i.e. larger than a fused tile's share (pre-tokenised word by word in stage P;
    sentence and clause token counts are then differences of the per-word running sum, partial edge words apart)."""
    import random
    import fuzzgen
    from oracle import chunker as OC, preprocess as OP
    rng = random.Random(seed)
    extra = ["!!!", "?!", ".)", '."', ".'", "..", "...", "--", "Mr.", "a.", "b.c.", "3.", "-3.5.", "1,2.", "x.y", "(z)", "\"q\"",
             "end.", "End.", "É.", "é", "no;", "yes:", "so,", "A", "the", "of", "and"]
    bad = 0
    for it in range(n_cases):
        nseg = rng.randint(1, 4)
        segs = []
        t = 0.0
        for k in range(nseg):
            words = [rng.choice(fuzzgen.WORDS + extra) for _ in range(rng.randint(900, 1800) if giant else rng.randint(30, 260))]
            if rng.random() < 0.15:
                words = [w.strip(",.;:?!") or "w" for w in words]          # forces the 20-word path
            elif rng.random() < 0.2:
                words = [w.replace(".", "").replace("?", "").replace("!", "") or "w" for w in words]   # clause path
            d = rng.choice([10.0, 61, 333.33, 0.5])
            segs.append({"start": t if rng.random() < 0.5 else int(t), "end": t + d, "text": " ".join(words),
                         "speaker": rng.choice(["A", "B", ""])})
            t += d
        mt = rng.choice([158, 165, 180, 220, 300] + ([600, 1500] if giant else []))
        op = OP.preprocess(segs, rng.random() < 0.5)
        merge = "is_combined" in json.dumps(op)
        proc, chunks = run_dropin(eng, segs, mt, merge=merge)
        op = OP.preprocess(segs, merge)
        oc = OC.ChunkerOracle(mt).chunk(op)
        assert same(list(proc), op), "case %d preprocess" % it
        assert same(chunks, oc), "case %d: %s" % (it, first_diff(chunks, oc))
    return n_cases


def agg_norm(agg):
    """Time-interval segments with the 'speaker' join (a Python set's iteration order, preprocessor.py:296,311:
    it differs from process to process) reduced to the set it was joined from."""
    out = []
    for a in agg:
        names = set(t["speaker"] for t in a["segment_timestamps"])
        b = dict(a)
        if len(names) > 1:
            assert len(a["speaker"]) == sum(map(len, names)) + 2 * (len(names) - 1) and all(n in a["speaker"] for n in names)
            b["speaker"] = sorted(names)
        else:
            assert a["speaker"] == next(iter(names))
        out.append(b)
    return out


def check_agg_cases(eng, cases, foreign=False):
    """Drop-in preprocess_transcript(time_interval_seconds=...) against the reference's outputs; foreign=True
    feeds aggregate_by_time_interval a plain deep copy of the oracle's processed list instead."""
    import copy
    from oracle import preprocess as OP
    for i, cs in enumerate(cases):
        kw = cs["kw"]
        E.set_default_engine(eng)
        with quiet():
            if foreign:
                proc = OP.preprocess(cs["segments"], kw["merge"], None, kw["max_dur"], kw["preserve"])
                got = PP.aggregate_by_time_interval(copy.deepcopy(proc), kw["interval"]) if proc else []
            else:
                got = PP.preprocess_transcript(cs["segments"], kw["merge"], kw["interval"], kw["max_dur"], kw["preserve"])
        assert same(agg_norm(got), agg_norm(cs["agg"])), "case %d: %s" % (i, first_diff(agg_norm(got), agg_norm(cs["agg"])))
    return len(cases)


def check_agg_batch(eng, batch, interval, docs):
    """Engine.aggregate on a multi-document batch against the oracle, document by document (same process, so
    the speaker joins agree too)."""
    from llm_map_reduce_summarizer_b200 import synth
    from oracle import preprocess as OP
    eng.preprocess(batch)
    P = eng.fetch_processed()
    eng.aggregate(interval)
    AG = eng.fetch_aggregated()
    c = eng.counts()
    assert c["n_agg"] == len(AG["agg"]) and c["n_agg_mem"] == len(AG["mem"]) and c["b_agg"] == len(AG["text"])
    for x in docs:
        segs = synth.doc_to_segments(batch, x)
        want = OP.preprocess(segs, True, interval)
        proc = E.build_processed(batch, P, x)
        got = E.build_aggregated(batch, P, AG, proc, interval, x) if proc else []
        assert same(got, want), "doc %d: %s" % (x, first_diff(got, want))


def small_rank_tables():
    """Other byte-pair tables for the table-driven kernels: the 256 single bytes (every token is one byte, SURVEY.md
    Appendix C.2) and the first 1 200 ranks of the repo table (a prefix of a BPE table is a BPE table; nearly every
    word then needs byte-pair merging, long spans included)."""
    from oracle import tokenizer as TK
    full = TK.load_ranks()
    by_rank = sorted(full.items(), key=lambda kv: kv[1])
    out = []
    for name, n in (("bytes256", 256), ("prefix1200", 1200)):
        sub = dict(by_rank[:n])
        assert all(len(t) == 1 for t, r in by_rank[:256])
        out.append((name, sub))
    return out


def check_other_vocab(make_engine, name, ranks, n_cases=25, seed=3):
    """Drop-in vs oracle under the rank table `ranks` (dict bytes -> rank)."""
    import random
    import fuzzgen
    from oracle import chunker as OC, preprocess as OP, tokenizer as TK
    toks = sorted(ranks, key=lambda t: ranks[t])
    eng = make_engine((name, toks, [ranks[t] for t in toks]))
    enc = TK.encoding(ranks, name)
    count = lambda s_: len(enc.encode(s_))
    strs = ["", " ", "hello world", "a" * 120, "ab" * 70 + " x", "!" * 90, "don't stop-me now", "naïve café 你好", "1234567 89",
            "The quick brown fox jumps over the lazy dog.", "  lead", "trail  ", "x\ny\r\n z", "supercalifragilisticexpialidocious"]
    cnt, _ = eng.count_tokens(strs)
    for s_, c in zip(strs, cnt):
        assert c == len(enc.encode_ordinary(s_)), (name, s_)
    rng = random.Random(seed)
    done = 0
    for it in range(n_cases):
        segs = fuzzgen.rand_transcript(rng, long_p=0.1)
        mt = rng.choice([170, 260, 400, 4000])
        merge = rng.random() < 0.8
        want_p = OP.preprocess(segs, merge)
        try:
            want_c = OC.ChunkerOracle(mt, count=count).chunk(want_p)
        except ValueError:
            want_c = "ValueError"
        try:
            got_p, got_c = run_dropin(eng, segs, mt, merge)
        except ValueError:
            got_p, got_c = want_p, "ValueError"
        assert same(list(got_p), want_p), (name, it)
        assert same(got_c, want_c), (name, it, first_diff(got_c, want_c) if isinstance(want_c, list) and isinstance(got_c, list) else (got_c, want_c))
        done += 1
    return done


def dense_tile_docs():
    """Transcripts that overrun the fused kernel's per-tile capacities (trc_fused3.cuh: word list, duplicate-candidate
    list, output buffer, input buffer) so that every overflow path runs: one-letter words, repeated words, a missing
    space after every second byte, and segments sized around the input capacity of a 256-segment tile."""
    import random
    rng = random.Random(99)
    docs = []
    def doc(texts, spk=2):
        return [{"start": 1.5 * i, "end": 1.5 * i + 1.5, "speaker": "S%d" % (i // 7 % spk), "text": t} for i, t in enumerate(texts)]
    docs.append(doc([" ".join(rng.choice("abcdefgh") for _ in range(35)) for _ in range(600)]))            # ~9k words per tile
    docs.append(doc([" ".join(["go"] * rng.randint(20, 34)) + " x" for _ in range(600)]))                   # every word a duplicate
    docs.append(doc([".".join(rng.choice("abcXYZ") for _ in range(36)) for _ in range(600)]))               # output 1.5x the input
    docs.append(doc(["w" * 71 for _ in range(300)] + ["v" * 72 for _ in range(300)] + ["u" * 73 for _ in range(300)]))   # at / over the input capacity
    docs.append(doc([("é " * 30).strip() for _ in range(300)] + ["你好 你好 " * 9 for _ in range(300)]))     # multi-byte everywhere
    docs.append(doc([("ab " * 20) + "\t\n " * 5 + ("cd  " * 5) for _ in range(520)]))                      # whitespace runs
    return docs


def check_dense_tiles(eng):
    from llm_map_reduce_summarizer_b200 import engine as E2
    from oracle import chunker as OC, preprocess as OP
    docs = dense_tile_docs()
    b = E2.pack_documents(docs)
    import numpy as np
    mt = np.array([4000, 700, 4000, 2000, 4000, 1000], dtype=np.int32)
    eng.preprocess(b)
    eng.chunk(mt)
    P, R = eng.fetch_processed(), eng.fetch_chunks()
    for x, segs in enumerate(docs):
        op = OP.preprocess(segs)
        oc = OC.ChunkerOracle(int(mt[x])).chunk(op)
        pp = E2.build_processed(b, P, x)
        assert same(pp, op), "doc %d processed: %s" % (x, first_diff(pp, op))
        got = E2.build_chunks(eng, b, P, R, pp, x)
        assert same(got, oc), "doc %d chunks: %s" % (x, first_diff(got, oc))
    return len(docs)


def check_prompts(eng):
    from llm_map_reduce_summarizer_b200 import synth
    b, mt, _ = synth.config(5, 0.004)
    eng.preprocess(b)
    eng.chunk(mt)
    R = eng.fetch_chunks()
    twc = R["twc"].tobytes()
    ch = R["chunks"]
    texts = []
    for c in ch:
        n = int(c["text_len"]) + (int(c["hdr_len"]) + 2 if c["hdr_len"] > 0 else 0)
        texts.append(twc[int(c["twc_off"]):int(c["twc_off"]) + n].decode("utf-8"))
    for tpl, kw in (("Summarise ({summary_type}):\n\n{transcript}\n\nEnd {{braces}} é.", {"summary_type": "brief"}),
                    ("{transcript}", {}), ("no field at all", {}), ("{transcript}|{transcript}|{x}", {"x": 3}), ("", {})):
        buf, off = eng.assemble_prompts(tpl, **kw)
        raw = buf.tobytes()
        assert len(off) == len(ch) + 1
        for k, t in enumerate(texts):
            assert raw[int(off[k]):int(off[k + 1])].decode("utf-8") == tpl.format(transcript=t, **kw), (tpl, k)
    import pytest
    with pytest.raises(NotImplementedError):
        eng.assemble_prompts("{transcript!r}")
    with pytest.raises(KeyError):
        eng.assemble_prompts("{transcript} {missing}")


def check_len32_input(eng):
    """trc_batch.seg_len32: int32 lengths uploaded instead of the int64 offsets give the same batch."""
    import numpy as np
    from llm_map_reduce_summarizer_b200 import synth
    b, mt, _ = synth.config(5, 0.004)
    eng.preprocess(b)
    eng.chunk(mt)
    want = {k: eng.array(k, "u1").tobytes() for k in ("CLEAN", "COFF", "K2R", "CTOK", "GSTART", "TWC", "CHUNKS")}
    eng.preprocess(b, seg_len32=np.diff(b.seg_off).astype(np.int32))
    eng.chunk(mt)
    for k, v in want.items():
        assert eng.array(k, "u1").tobytes() == v, k


def check_edge_batches(eng):
    """Ragged and extreme batches in ONE call: empty documents, all-blank documents, one 300 KB segment, thousands of
    speakers, a 1 KB speaker name, NUL bytes, only multi-byte whitespace - every document against the oracle."""
    import random
    import numpy as np
    from llm_map_reduce_summarizer_b200 import engine as E2
    from oracle import chunker as OC, preprocess as OP
    rng = random.Random(17)
    words = ["alpha", "beta,", "gamma.", "delta", "it's", "3.5", "go", "go", "é", "naïve", "ok!", "so;", "and"]
    def seg(t, text, spk):
        return {"start": t, "end": t + 2.5, "text": text, "speaker": spk}
    docs = [
        [],
        [seg(0, "", "A"), seg(1, " \t\n", "B"), seg(2, "\u00a0\u2003", "A")],
        [seg(0, " ".join(rng.choice(words) for _ in range(60000)), "MONO")],
        [seg(i * 2.5, "word%d same same" % i, "speaker_%05d" % (i * 7919 % 3001)) for i in range(3001)],
        [seg(0, "hello there", "N" * 1000), seg(3, "again again", "N" * 1000), seg(6, "x", "")],
        [seg(0, "nul\x00inside \x00 \x00", "A"), seg(1, "\x00", "A")],
        [seg(0, "\u3000\u2028lead", "A"), seg(1, "trail\u2029\u3000", "A"), seg(2, "a\u00a0\u00a0b", "B")],
        [],
        [seg(5, "last one.", "Z")],
    ]
    b = E2.pack_documents(docs)
    mt = np.array([4000, 4000, 2000, 300, 4000, 4000, 4000, 4000, 200], dtype=np.int32)
    eng.preprocess(b)
    eng.chunk(mt)
    P, R = eng.fetch_processed(), eng.fetch_chunks()
    for x, segs in enumerate(docs):
        op = OP.preprocess(segs)
        oc = OC.ChunkerOracle(int(mt[x])).chunk(op)
        pp = E2.build_processed(b, P, x)
        assert same(pp, op), "doc %d processed: %s" % (x, first_diff(pp, op))
        got = E2.build_chunks(eng, b, P, R, pp, x)
        assert same(got, oc), "doc %d chunks: %s" % (x, first_diff(got, oc))
    # an entirely empty batch
    e0 = E2.pack_documents([[], []])
    eng.preprocess(e0)
    eng.chunk(np.array([4000, 4000], dtype=np.int32))
    c = eng.counts()
    assert c["n_seg"] == 0 and c["n_chunks"] == 0 and c["tokens"] == 0
    return len(docs)
