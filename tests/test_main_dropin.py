"""Config 1: transcript-example.json through the reference's own main.py (mock LLM executor) with this
package's modules aliased in, as INTEGRATION.md describes.  Needs /root/reference (build container only)."""
import json, os, subprocess, sys

import pytest

HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.mark.skipif(not os.path.isdir("/root/reference"), reason="the reference tree only exists in the build container")
def test_main_py_runs_with_dropin_modules(tmp_path):
    outs = {}
    for mode in ("reference", "dropin"):
        d = tmp_path / mode
        d.mkdir()
        r = subprocess.run([sys.executable, os.path.join(HERE, "run_reference_main.py"), mode, str(d)],
                           capture_output=True, text=True, timeout=600)
        assert r.returncode == 0, r.stderr[-2000:]
        outs[mode] = json.load(open(d / "chunks.json"))["chunks"]
    a, b = outs["reference"], outs["dropin"]
    assert len(a) == len(b) > 10
    assert [(c["chunk_index"], c["start_time"], c["end_time"]) for c in a] == \
           [(c["chunk_index"], c["start_time"], c["end_time"]) for c in b]


@pytest.mark.skipif(not os.path.isdir("/root/reference"), reason="the reference tree only exists in the build container")
def test_save_chunks_writer_reproduces_the_reference_file(tmp_path):
    """egress.save_chunk_records (main.py:178-201) on this package's chunk dicts, with the map step's two fields taken
    from the reference run, writes the reference's --save-chunks file byte for byte."""
    import gzip
    sys.path.insert(0, os.path.join(HERE, ".."))
    from conftest import quiet
    from llm_map_reduce_summarizer_b200 import build, engine as E, preprocessor as PP, big_chunkeroosky as BC, egress
    d = tmp_path / "ref"
    d.mkdir()
    r = subprocess.run([sys.executable, os.path.join(HERE, "run_reference_main.py"), "reference", str(d)], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    want_text = open(d / "chunks.json", encoding="utf-8").read()
    want = json.loads(want_text)
    E.set_default_engine(E.Engine(0, build.build_hostsim()))
    segs = json.load(gzip.open(os.path.join(HERE, "golden", "example_segments.json.gz")))
    with quiet():
        c = BC.BigChunkeroosky(max_tokens_per_chunk=2000)
        chunks = c.postprocess_chunks(c.chunk_transcript(PP.preprocess_transcript(segs)))
    assert len(chunks) == len(want["chunks"])
    for ch, w in zip(chunks, want["chunks"]):
        ch["summary"], ch["tokens_used"] = w["summary"], w["tokens_used"]
    out = tmp_path / "ours.json"
    egress.save_chunk_records(chunks, str(out), timestamp=want["timestamp"])
    assert open(out, encoding="utf-8").read() == want_text
