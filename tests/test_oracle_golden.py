"""The oracle against the golden vectors produced by the UNMODIFIED reference (tools/make_golden.py)."""
import hashlib, json, random

from conftest import load_golden
from oracle import chunker as OC, preprocess as OP, tokenizer as TK, punkt


def sha(o):
    return hashlib.sha256(json.dumps(o, sort_keys=True).encode()).hexdigest()


def test_example_all_configs():
    segs = load_golden("example_segments.json.gz")
    exp = load_golden("example_expected.json")
    for name, e in exp.items():
        kw = e["kw"]
        proc = OP.preprocess(segs, kw.get("merge", True), None, kw.get("max_dur", 120), kw.get("preserve", True))
        chunks = OC.ChunkerOracle(kw["max_tokens"]).chunk(proc, kw.get("add_context", True))
        assert len(proc) == e["n_proc"] and sha(proc) == e["proc_sha256"], name
        assert [c["token_count"] for c in chunks] == e["token_counts"], name
        assert sha(chunks) == e["chunks_sha256"], name


def test_survey_kat_sha():
    # sha256 of the 171 merged segments of the example
    assert load_golden("example_expected.json")["default"]["proc_sha256"] == \
        "2acb391256fefb6dc5e59876fd6d767a88ece5a3b4b38fe58de39445a7fb508e"


def test_kats():
    kat = load_golden("kat.json")
    for s, want in kat["clean_text"]:
        assert OP.clean_text(s) == want, repr(s)
    for x, want in kat["format_timestamp"]:
        assert OP.format_ts(x) == want, x
    for m in kat["merge"]:
        out = OP.preprocess(m["segments"], True, None, m["cap"])
        assert [o.get("original_segments", 1) for o in out] == m["group_sizes"], m["name"]
        assert sha(out) == m["sha256"], m["name"]


def test_fuzz_cases():
    for i, cs in enumerate(load_golden("fuzz_cases.json.gz")):
        kw = cs["kw"]
        proc = OP.preprocess(cs["segments"], kw["merge"], None, kw["max_dur"], kw["preserve"])
        assert json.dumps(proc) == json.dumps(cs["proc"]), i
        try:
            chunks, err = OC.ChunkerOracle(kw["max_tokens"]).chunk(proc, kw["add_context"]), None
        except ValueError:
            chunks, err = None, "ValueError"
        assert err == cs["error"], i
        assert json.dumps(chunks) == json.dumps(cs["chunks"]), i


def test_aggregate_cases():
    """aggregate_by_time_interval (preprocessor.py:217-324) against the reference's outputs."""
    import parity_common as pc
    g = load_golden("agg_cases.json.gz")
    nonempty = multi = 0
    for i, cs in enumerate(g["fuzz"]):
        kw = cs["kw"]
        got = OP.preprocess(cs["segments"], kw["merge"], kw["interval"], kw["max_dur"], kw["preserve"])
        assert pc.same(pc.agg_norm(got), pc.agg_norm(cs["agg"])), i
        nonempty += bool(got)
        multi += sum(1 for a in got if a["original_segments"] > 1)
    assert nonempty > 60 and multi > 200
    segs = load_golden("example_segments.json.gz")
    for iv, e in g["example"].items():
        out = OP.preprocess(segs, True, float(iv) if "." in iv else int(iv))
        for o in out:
            o["speaker"] = sorted(o["speaker"].split(", "))
        assert len(out) == e["n"] and sha(out) == e["sha256"], iv


def test_python_bpe_restatement_matches_corebpe():
    ranks = TK.load_ranks()
    enc = TK.encoding()
    rng = random.Random(5)
    alpha = ["a", "b", "Z", "é", "1", "2", "٣", " ", "  ", "\n", "\r", "\t", " ", "'", "'s", "'LL", "'ſ", ".", ",",
             "!", "?", "-", "(", "你", "_", "x0", " ", "​", "[00:05]", " don't", "hello", " world"]
    for _ in range(3000):
        s = "".join(rng.choice(alpha) for _ in range(rng.randint(0, 14)))
        assert TK.count_python(ranks, s) == len(enc.encode_ordinary(s)), repr(s)


def test_punkt_documented_examples():
    p = punkt.PunktSentenceTokenizer()
    # docstring example of nltk 3.8.1 PunktSentenceTokenizer._match_potential_end_contexts
    assert p.tokenize("Very bad acting!!! I promise.") == ["Very bad acting!!!", "I promise."]
    assert p.tokenize("J. Bach wrote it. He left.") == ["J. Bach wrote it.", "He left."]
    assert p.tokenize('He said "No." Then left') == ['He said "No."', "Then left"]
    assert p.tokenize("Ok... fine. 3.5 is. 3. then x") == ["Ok... fine.", "3.5 is.", "3. then x"]
