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
