"""Parity tests proper: the CUDA kernels (through the C ABI) against the golden vectors of the real
reference and against the oracle.  Bit-exact: chunk boundaries, token counts, timestamps, headers,
float64 percentages."""
import json
import random

import numpy as np
import pytest

import parity_common as pc
from conftest import load_golden
from llm_map_reduce_summarizer_b200 import engine as E, preprocessor as PP, synth
from oracle import chunker as OC, preprocess as OP, tokenizer as TK
from test_oracle_golden import sha

pytestmark = pytest.mark.gpu


def test_example_all_configs(gpu_engine):
    segs = load_golden("example_segments.json.gz")
    for name, e in load_golden("example_expected.json").items():
        kw = e["kw"]
        proc, chunks = pc.run_dropin(gpu_engine, segs, kw["max_tokens"], kw.get("merge", True), kw.get("max_dur", 120),
                                     kw.get("preserve", True), kw.get("add_context", True))
        assert sha(list(proc)) == e["proc_sha256"], name
        assert [c["token_count"] for c in chunks] == e["token_counts"], name
        assert sha(chunks) == e["chunks_sha256"], name


def test_kats(gpu_engine):
    E.set_default_engine(gpu_engine)
    kat = load_golden("kat.json")
    for s, want in kat["clean_text"]:
        assert PP.clean_text(s) == want, repr(s)
    for x, want in kat["format_timestamp"]:
        assert PP.format_timestamp(x) == want, x
    for m in kat["merge"]:
        proc, _ = pc.run_dropin(gpu_engine, m["segments"], 100000, True, m["cap"])
        assert sha(list(proc)) == m["sha256"], m["name"]


def test_fuzz_cases(gpu_engine):
    n = pc.check_fuzz_cases(gpu_engine, load_golden("fuzz_cases.json.gz"), require_all=True)
    assert n == 160


def test_token_counts_vs_corebpe(gpu_engine):
    enc = TK.encoding()
    rng = random.Random(11)
    alpha = ["a", "b", "Z", "é", "1", "2", "٣", " ", "  ", "\n", "\r", "\t", " ", "'", "'s", "'LL", "'ſ", ".", ",",
             "!", "?", "-", "(", "你", "_", "x0", " ", "​", "[00:05]", " don't", "hello", " world", "ing", "tion"]
    strs = ["".join(rng.choice(alpha) for _ in range(rng.randint(0, 40))) for _ in range(20000)]
    # long spans: block-wide merge kernel with global scratch (k_bpe_long)
    strs += ["a" * 200, "ab" * 300, "!" * 100, " " * 70, "x" * 49, "y" * 48, "0123456789" * 9, "né" * 60,
             "the" * 700, "z" * 5000, "".join(rng.choice("abcdefghij") for _ in range(3000))]
    for lead in (False, True):
        cnt, _ = gpu_engine.count_tokens(strs, leading_space=lead)
        for s, c in zip(strs, cnt):
            assert c == len(enc.encode_ordinary((" " if lead else "") + s)), repr(s[:60])


def test_special_tokens_raise_value_error(gpu_engine):
    segs = [{"start": 0, "end": 1, "text": "hello <|endoftext|> there", "speaker": "A"}]
    with pytest.raises(ValueError):
        pc.run_dropin(gpu_engine, segs)


def _check_docs(e, b, mt, docs):
    e.preprocess(b)
    e.chunk(mt)
    P, R = e.fetch_processed(), e.fetch_chunks()
    for x in docs:
        segs = synth.doc_to_segments(b, x)
        op = OP.preprocess(segs)
        oc = OC.ChunkerOracle(int(mt[x])).chunk(op)
        pp = E.build_processed(b, P, x)
        assert json.dumps(pp) == json.dumps(op), x
        assert json.dumps(E.build_chunks(e, b, P, R, pp, x)) == json.dumps(oc), x


def test_config2_meeting(gpu_engine):
    b, mt, _ = synth.config(2, 1.0)
    _check_docs(gpu_engine, b, mt, [0])


def test_config3_podcasts_sample(gpu_engine):
    b, mt, _ = synth.config(3, 0.05)
    _check_docs(gpu_engine, b, mt, range(0, b.n_doc, 5))


def test_config5_mixed_limits_sample(gpu_engine):
    b, mt, _ = synth.config(5, 0.03)
    _check_docs(gpu_engine, b, mt, range(0, b.n_doc, 3))


def test_full_size_properties(gpu_engine):
    """Config 3 at full size (10k documents, ~128 MB): size-independent invariants + oracle on a sample."""
    e = gpu_engine
    b, mt, _ = synth.config(3, 1.0)
    e.preprocess(b)
    e.chunk(mt)
    c = e.counts()
    R = e.fetch_chunks()
    ch = R["chunks"]
    assert c["n_doc"] == b.n_doc and c["n_kept"] <= c["n_seg"]
    assert (e.array("CLEN32", "<i4") == np.diff(e.array("COFF", "<i8"))).all()
    # every item lands in exactly one chunk, chunks tile the item list, token sums agree
    assert (ch["item0"][1:] == ch["item1"][:-1]).all() and ch["item0"][0] == 0 and ch["item1"][-1] == c["n_items"]
    assert int(ch["token_count"].sum()) == int(R["item_tok"].sum()) == c["tokens"]
    # no chunk exceeds its budget unless it is a single over-budget item
    E_doc = (mt - 150)[ch["doc"]]
    over = ch["token_count"] > E_doc
    assert ((ch["item1"] - ch["item0"])[over] == 1).all()
    # the output stream is tiled by the chunks
    assert int(ch["twc_off"][0]) == 0
    hdr = np.where(ch["hdr_len"] > 0, ch["hdr_len"] + 2, 0)
    assert (ch["twc_off"][1:] == (ch["twc_off"] + hdr + ch["text_len"])[:-1]).all()
    assert int((ch["twc_off"] + hdr + ch["text_len"])[-1]) == c["b_twc"]
    P = e.fetch_processed()
    for x in range(0, b.n_doc, 997):
        segs = synth.doc_to_segments(b, x)
        op = OP.preprocess(segs)
        oc = OC.ChunkerOracle(int(mt[x])).chunk(op)
        pp = E.build_processed(b, P, x)
        assert json.dumps(pp) == json.dumps(op), x
        assert json.dumps(E.build_chunks(e, b, P, R, pp, x)) == json.dumps(oc), x


def test_idempotent_rerun(gpu_engine):
    e = gpu_engine
    b, mt, _ = synth.config(3, 0.02)
    e.preprocess(b); e.chunk(mt)
    a = e.array("TWC", "u1").tobytes()
    e.preprocess(b); e.chunk(mt)
    assert a == e.array("TWC", "u1").tobytes()


def test_sentence_path_fuzz(gpu_engine):
    pc.sentence_fuzz(gpu_engine, 300, 1234)
    pc.sentence_fuzz(gpu_engine, 40, 77, giant=True)


def test_config4_monologues(gpu_engine):
    """One long transcript: huge same-speaker groups (combined-piece path) + 20-80 KB monologues
    (sentence, clause and 20-word paths); the slow lane of the fused kernel handles the giant segments."""
    b, mt, _ = synth.config(4, 0.03)
    _check_docs(gpu_engine, b, mt, [0])
    c = gpu_engine.counts()
    assert c["n_sent"] > 1000 and c["n_subs"] > 100


def test_foreign_processed_lists(gpu_engine):
    """chunk_transcript on lists this package did not produce (here: the reference's own processed lists)."""
    n = pc.check_fuzz_cases(gpu_engine, load_golden("fuzz_cases.json.gz"), require_all=True, foreign=True)
    assert n == 160
    segs = load_golden("example_segments.json.gz")
    from oracle import preprocess as OP2, chunker as OC2
    proc = OP2.preprocess(segs)
    want = OC2.ChunkerOracle(2000).chunk(proc)
    E.set_default_engine(gpu_engine)
    from conftest import quiet
    from llm_map_reduce_summarizer_b200 import big_chunkeroosky as BC2
    with quiet():
        c = BC2.BigChunkeroosky(max_tokens_per_chunk=2000)
        got = c.postprocess_chunks(c.chunk_transcript(proc))
    assert pc.same(got, want)
    assert all(a is b for ch in got for a, b in zip(ch["segments"], ch["segments"])) and got[0]["segments"][0] is proc[0]
    # an inconsistent combined segment is refused, not approximated
    import copy, pytest
    bad = copy.deepcopy(proc)
    bad[0]["text"] = bad[0]["text"] + " edited"
    with quiet(), pytest.raises(NotImplementedError):
        c.chunk_transcript(bad)


def test_aggregate_by_time_interval(gpu_engine):
    """preprocessor.py:217-324 on the device, against the reference's outputs and the oracle."""
    g = load_golden("agg_cases.json.gz")["fuzz"]
    assert pc.check_agg_cases(gpu_engine, g) == 90
    assert pc.check_agg_cases(gpu_engine, g[::3], foreign=True) == 30
    b, _, _ = synth.config(3, 0.02)
    for iv in (60, 300, 7.5):
        pc.check_agg_batch(gpu_engine, b, iv, range(0, b.n_doc, 2))
    b, _, _ = synth.config(2, 0.2)
    pc.check_agg_batch(gpu_engine, b, 30, range(b.n_doc))


@pytest.mark.gpu
@pytest.mark.parametrize("table", pc.small_rank_tables(), ids=lambda t: t[0])
def test_other_rank_tables(table):
    """The kernels are table driven: the same parity under two other byte-pair tables (nearly every
    word needs byte-pair merging under the small one: slow-word kernel, long-span kernel, memo of failures)."""
    name, ranks = table
    assert pc.check_other_vocab(lambda rt: E.Engine(0, rank_table=rt), name, ranks, n_cases=40) == 40
    # and a synthetic batch through the fused kernel, token totals against the oracle tokenizer
    from oracle import tokenizer as TK
    toks = sorted(ranks, key=lambda t: ranks[t])
    eng = E.Engine(0, rank_table=(name, toks, [ranks[t] for t in toks]))
    b, mt, _ = synth.config(3, 0.01)
    eng.preprocess(b)
    P = eng.fetch_processed()
    enc = TK.encoding(ranks, name)
    clean = P["clean"].tobytes()
    coff = P["coff"]
    for j in range(0, len(P["k2r"]), 7):
        t = clean[int(coff[j]):int(coff[j + 1])].decode("utf-8")
        assert int(P["ctok"][j]) == len(enc.encode_ordinary(" " + t)), (name, j)


def test_dense_tiles(gpu_engine):
    """Tiles that overrun the fused kernel's word / candidate / output / input capacities (overflow paths)."""
    assert pc.check_dense_tiles(gpu_engine) == 6


@pytest.mark.parametrize("env", ["TRCHUNK_FUSED=2", "TRCHUNK_FUSED=3", "TRCHUNK_FUSED=4", "TRCHUNK_NO_FUSED=1", "TRCHUNK_NO_MEMO=1"])
def test_alternate_kernel_paths(monkeypatch, env):
    """The older kernel generations kept for A/B runs and the fused kernel without its word memo:
    the switches are read when a context is created."""
    k, v = env.split("=")
    monkeypatch.setenv(k, v)
    eng = E.Engine(0)
    try:
        b, mt, _ = synth.config(5, 0.01)
        _check_docs(eng, b, mt, range(0, b.n_doc, 2))
        assert pc.check_dense_tiles(eng) == 6
        pc.sentence_fuzz(eng, 30, 5)
    finally:
        eng.close()


def test_prompt_assembly(gpu_engine):
    pc.check_prompts(gpu_engine)


def test_len32_input(gpu_engine):
    pc.check_len32_input(gpu_engine)


def test_edge_batches(gpu_engine):
    assert pc.check_edge_batches(gpu_engine) == 9
