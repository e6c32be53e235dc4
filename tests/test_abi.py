"""The C-ABI library loads and exports every symbol include/trchunk.h declares (no compute calls)."""
import ctypes, os, re

from llm_map_reduce_summarizer_b200 import _native, build

ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")


def declared():
    src = open(os.path.join(ROOT, "include", "trchunk.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(trc_[a-z_]+)\s*\(", src)))


def test_header_matches_binding_list():
    assert declared() == sorted(_native.EXPORTS)


def test_cuda_library_builds_and_exports():
    lib = ctypes.CDLL(build.build())
    for name in declared():
        assert hasattr(lib, name), name


def test_no_device_fails_loudly():
    import pytest, torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from llm_map_reduce_summarizer_b200 import engine
    with pytest.raises(engine.TrcError):
        engine.Engine(0)


def test_struct_layouts_match_the_header(tmp_path):
    """sizeof of every struct of include/trchunk.h (compiled by gcc as C) against the ctypes / numpy mirrors."""
    import subprocess
    import numpy as np
    src = tmp_path / "sz.c"
    src.write_text('#include <stdio.h>\n#include "trchunk.h"\nint main(void) { printf("%zu %zu %zu %zu %zu %zu\\n", sizeof(trc_batch), '
                   'sizeof(trc_sub), sizeof(trc_chunk_rec), sizeof(trc_sent), sizeof(trc_agg), sizeof(trc_agg_mem)); return 0; }\n')
    exe = tmp_path / "sz"
    subprocess.run(["gcc", "-std=c99", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)], check=True)
    got = [int(x) for x in subprocess.run([str(exe)], check=True, capture_output=True, text=True).stdout.split()]
    want = [ctypes.sizeof(_native.Batch), np.dtype(_native.SUB_DTYPE).itemsize, np.dtype(_native.CHUNK_DTYPE).itemsize,
            np.dtype(_native.SENT_DTYPE).itemsize, np.dtype(_native.AGG_DTYPE).itemsize, np.dtype(_native.AGG_MEM_DTYPE).itemsize]
    assert got == want
