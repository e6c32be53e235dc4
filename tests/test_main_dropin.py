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
