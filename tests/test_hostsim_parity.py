"""Stage logic + host orchestration (product sources compiled for the host, tests/hostsim) against the
golden vectors and the oracle.  CPU only; the GPU parity tests proper are tests/test_gpu_parity.py."""
import json

import numpy as np
import pytest

import parity_common as pc
from conftest import load_golden
from llm_map_reduce_summarizer_b200 import engine as E, preprocessor as PP, synth
from oracle import chunker as OC, preprocess as OP, tokenizer as TK
from test_oracle_golden import sha


def test_example_all_configs(hostsim_engine):
    segs = load_golden("example_segments.json.gz")
    for name, e in load_golden("example_expected.json").items():
        kw = e["kw"]
        proc, chunks = pc.run_dropin(hostsim_engine, segs, kw["max_tokens"], kw.get("merge", True), kw.get("max_dur", 120),
                                     kw.get("preserve", True), kw.get("add_context", True))
        assert sha(list(proc)) == e["proc_sha256"], name
        assert sha(chunks) == e["chunks_sha256"], name


def test_kats(hostsim_engine):
    E.set_default_engine(hostsim_engine)
    kat = load_golden("kat.json")
    for s, want in kat["clean_text"]:
        assert PP.clean_text(s) == want, repr(s)
    for x, want in kat["format_timestamp"]:
        assert PP.format_timestamp(x) == want, x
    for m in kat["merge"]:
        proc, _ = pc.run_dropin(hostsim_engine, m["segments"], 100000, True, m["cap"])
        assert sha(list(proc)) == m["sha256"], m["name"]


def test_fuzz_cases(hostsim_engine):
    n = pc.check_fuzz_cases(hostsim_engine, load_golden("fuzz_cases.json.gz"), require_all=True)
    assert n == 160


@pytest.mark.parametrize("tile", ["1", "3", "0"])
def test_fuzz_cases_tile_edges(hostsim_engine, monkeypatch, tile):
    """The cleaner's tile edges: tiles of 1 and 3 segments put every segment boundary on a tile boundary, with
    the neighbouring tiles' bytes visible past the span, as in the fused kernel; 0 = one tile per batch."""
    monkeypatch.setenv("TRCHUNK_HOSTSIM_TILE", tile)
    cases = load_golden("fuzz_cases.json.gz")
    assert pc.check_fuzz_cases(hostsim_engine, cases[::2], require_all=True) == 80
    b, mt, _ = synth.config(5, 0.004)
    hostsim_engine.preprocess(b)
    P = hostsim_engine.fetch_processed()
    for x in range(b.n_doc):
        assert json.dumps(E.build_processed(b, P, x)) == json.dumps(OP.preprocess(synth.doc_to_segments(b, x))), x


@pytest.mark.parametrize("tile", ["1", "3", "32", "0"])
def test_chunk32_cleaner_tile_edges(hostsim_engine, monkeypatch, tile):
    """The 32-byte-chunk formulation of clean_text that k_clean_tok_w32 runs (trc_w32.cuh), tiled as in the kernel:
    KATs, the reference's fuzz cases, synthetic documents and the dense tiles, with tile edges on every segment
    boundary (1), every third (3), the kernel's own 32, and one tile per batch (0)."""
    monkeypatch.setenv("TRCHUNK_HOSTSIM_CLEAN", "w32")
    monkeypatch.setenv("TRCHUNK_HOSTSIM_TILE", tile)
    E.set_default_engine(hostsim_engine)
    for s, want in load_golden("kat.json")["clean_text"]:
        assert PP.clean_text(s) == want, repr(s)
    cases = load_golden("fuzz_cases.json.gz")
    assert pc.check_fuzz_cases(hostsim_engine, cases[1::2], require_all=True) == 80
    b, mt, _ = synth.config(5, 0.004)
    hostsim_engine.preprocess(b)
    P = hostsim_engine.fetch_processed()
    for x in range(b.n_doc):
        assert json.dumps(E.build_processed(b, P, x)) == json.dumps(OP.preprocess(synth.doc_to_segments(b, x))), x
    assert pc.check_dense_tiles(hostsim_engine) == 6


def test_token_counts_vs_corebpe(hostsim_engine):
    import random
    enc = TK.encoding()
    rng = random.Random(11)
    alpha = ["a", "b", "Z", "é", "1", "2", "٣", " ", "  ", "\n", "\r", "\t", " ", "'", "'s", "'LL", "'ſ", ".", ",",
             "!", "?", "-", "(", "你", "_", "x0", " ", "​", "[00:05]", " don't", "hello", " world", "ing", "tion"]
    strs = ["".join(rng.choice(alpha) for _ in range(rng.randint(0, 40))) for _ in range(4000)]
    strs += ["a" * 200, "ab" * 300, "!" * 100, " " * 70, "x" * 49, "y" * 48, "0123456789" * 9, "né" * 60]
    for lead in (False, True):
        cnt, _ = hostsim_engine.count_tokens(strs, leading_space=lead)
        for s, c in zip(strs, cnt):
            assert c == len(enc.encode_ordinary((" " if lead else "") + s)), repr(s)


def test_synthetic_docs_vs_oracle(hostsim_engine):
    e = hostsim_engine
    for num, scale in ((3, 0.004), (5, 0.004), (2, 0.05)):
        b, mt, _ = synth.config(num, scale)
        e.preprocess(b)
        e.chunk(mt)
        P, R = e.fetch_processed(), e.fetch_chunks()
        assert (e.array("CLEN32", "<i4") == np.diff(P["coff"])).all()
        for x in range(b.n_doc):
            segs = synth.doc_to_segments(b, x)
            op = OP.preprocess(segs)
            oc = OC.ChunkerOracle(int(mt[x])).chunk(op)
            pp = E.build_processed(b, P, x)
            assert json.dumps(pp) == json.dumps(op), (num, x)
            assert json.dumps(E.build_chunks(e, b, P, R, pp, x)) == json.dumps(oc), (num, x)


def test_empty_inputs(hostsim_engine):
    proc, chunks = pc.run_dropin(hostsim_engine, [])
    assert list(proc) == [] and chunks == []
    proc, chunks = pc.run_dropin(hostsim_engine, [{"start": 0, "end": 1, "text": "  \t", "speaker": "A"}])
    assert list(proc) == [] and chunks == []


def test_sentence_path_fuzz(hostsim_engine):
    pc.sentence_fuzz(hostsim_engine, 300, 99)
    pc.sentence_fuzz(hostsim_engine, 10, 77, giant=True)


def test_foreign_processed_lists(hostsim_engine):
    """chunk_transcript on lists this package did not produce (here: the reference's own processed lists)."""
    n = pc.check_fuzz_cases(hostsim_engine, load_golden("fuzz_cases.json.gz"), require_all=True, foreign=True)
    assert n == 160
    segs = load_golden("example_segments.json.gz")
    from oracle import preprocess as OP2, chunker as OC2
    proc = OP2.preprocess(segs)
    want = OC2.ChunkerOracle(2000).chunk(proc)
    E.set_default_engine(hostsim_engine)
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


def test_combine_helpers(hostsim_engine):
    """combine_same_speaker_segments / create_combined_segment on already cleaned dicts
    (preprocessor.py:109-215): clean-then-combine must equal preprocess_transcript."""
    import random
    import fuzzgen
    from conftest import quiet
    E.set_default_engine(hostsim_engine)
    rng = random.Random(77)
    for _ in range(40):
        segs = fuzzgen.rand_transcript(rng)
        cap = rng.choice([1, 30, 120, None])
        pres = rng.random() < 0.8
        want = OP.preprocess(segs, True, None, cap, pres)
        unmerged = OP.preprocess(segs, False)
        with quiet():
            got = PP.combine_same_speaker_segments(unmerged, cap, pres)
        assert json.dumps(got) == json.dumps(want)
        assert all(g is u for g in got if "is_combined" not in g for u in unmerged if u is g) 
        if len(unmerged) >= 2:
            with quiet():
                one = PP.create_combined_segment(unmerged[:4], pres)
            assert one["original_segments"] == min(4, len(unmerged)) and one["is_combined"] is True
            assert one["text"].count(unmerged[0]["text"]) >= 1


def test_token_counter_service(hostsim_engine):
    import pytest
    from llm_map_reduce_summarizer_b200.token_counter import TokenCounter
    E.set_default_engine(hostsim_engine)
    enc = TK.encoding()
    tc = TokenCounter()
    texts = ["A short summary.", "", "Bullet points:\n- one\n- two\n\nDone.", "naïve café 你好 123456", "x" * 300]
    assert tc.count_batch(texts) == [len(enc.encode(t)) for t in texts]
    assert tc.total(texts) == sum(len(enc.encode(t)) for t in texts)
    with pytest.raises(ValueError):
        tc.count("bad <|endoftext|> text")


def test_aggregate_by_time_interval(hostsim_engine):
    g = load_golden("agg_cases.json.gz")["fuzz"]
    assert pc.check_agg_cases(hostsim_engine, g) == 90
    assert pc.check_agg_cases(hostsim_engine, g[::3], foreign=True) == 30
    b, _, _ = synth.config(3, 0.002)
    for iv in (60, 300, 7.5, 100000):
        pc.check_agg_batch(hostsim_engine, b, iv, range(b.n_doc))
    # error behaviour of the reference's own expression (:236)
    E.set_default_engine(hostsim_engine)
    with pytest.raises(ZeroDivisionError):
        PP.aggregate_by_time_interval([{"start": 0, "end": 5, "speaker": "A", "text": "x"}], 0)
    with pc.quiet():
        assert PP.aggregate_by_time_interval([{"start": 0, "end": 5, "speaker": "A", "text": "x"}], -3) == []
        assert PP.aggregate_by_time_interval([], 10) == []
        assert PP.preprocess_transcript([{"start": 0, "end": 5, "speaker": "A", "text": " "}], True, 10) == []


@pytest.mark.parametrize("table", pc.small_rank_tables(), ids=lambda t: t[0])
def test_other_rank_tables(table):
    """The kernels are table driven: the same parity under two other byte-pair tables."""
    import conftest
    name, ranks = table
    assert pc.check_other_vocab(lambda rt: E.Engine(0, lib_path=conftest.hostsim_library(), rank_table=rt), name, ranks) == 25


def test_dense_tiles(hostsim_engine):
    """Tiles that overrun the fused kernel's word / candidate / output / input capacities (overflow paths)."""
    assert pc.check_dense_tiles(hostsim_engine) == 6


def test_prompt_assembly(hostsim_engine):
    """llm_executor.py:187-190: prompt_template.format(transcript=chunk['text_with_context'], summary_type=...)."""
    pc.check_prompts(hostsim_engine)


def test_len32_input(hostsim_engine):
    pc.check_len32_input(hostsim_engine)


def test_edge_batches(hostsim_engine):
    assert pc.check_edge_batches(hostsim_engine) == 9
