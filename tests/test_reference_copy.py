"""oracle/_ref: the timed CPU arm must be the reference itself, not a restatement."""
import json, os, sys

import pytest

ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, ROOT)
from oracle import refimpl  # noqa: E402


@pytest.mark.skipif(not os.path.isdir(refimpl.REF_SRC), reason="/root/reference only exists in the build container")
def test_staged_copies_are_the_reference():
    assert refimpl.stage()
    for f in refimpl.FILES:
        assert open(os.path.join(refimpl.REF_SRC, f), "rb").read() == open(os.path.join(refimpl.REF_DIR, f), "rb").read(), f


@pytest.mark.skipif(not refimpl.available(), reason="oracle/_ref not staged")
def test_reference_run_equals_oracle_port():
    """The staged reference and the oracle port give the same chunks on synthetic documents (same rank table)."""
    from llm_map_reduce_summarizer_b200 import synth
    from oracle import chunker as OC, preprocess as OP
    rp, rc = refimpl.load()
    b, mt, _ = synth.config(5, 0.002)
    import contextlib, io
    tok = 0
    for x in range(b.n_doc):
        segs = synth.doc_to_segments(b, x)
        with contextlib.redirect_stdout(io.StringIO()):
            proc = rp.preprocess_transcript(segs)
            ch = rc.BigChunkeroosky(max_tokens_per_chunk=int(mt[x]))
            chunks = ch.postprocess_chunks(ch.chunk_transcript(proc))
        op = OP.preprocess(segs)
        oc = OC.ChunkerOracle(int(mt[x])).chunk(op)
        assert json.dumps(proc) == json.dumps(op) and json.dumps(chunks) == json.dumps(oc), x
        tok += sum(c["token_count"] for c in chunks)
    docs = [synth.doc_to_segments(b, x) for x in range(b.n_doc)]
    assert refimpl.run_docs(docs, [int(v) for v in mt])[0] == tok


def test_unknown_encoding_names_raise():
    from llm_map_reduce_summarizer_b200 import vocab
    for name in ("gpt2", "p50k_base", "o200k_base", "nope"):
        with pytest.raises(ValueError):
            vocab.load_rank_table(name)
    assert vocab.resolve_table_name("cl100k_base") in ("cl100k_base", "trbpe_v1")
