import contextlib, gzip, io, json, os, sys

import pytest

ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")


def load_golden(name):
    p = os.path.join(GOLDEN, name)
    if name.endswith(".gz"):
        return json.load(gzip.open(p))
    return json.load(open(p))


@contextlib.contextmanager
def quiet():
    with contextlib.redirect_stdout(io.StringIO()):
        yield


def hostsim_library():
    from llm_map_reduce_summarizer_b200 import build
    return build.build_hostsim()


@pytest.fixture(scope="session")
def hostsim_engine():
    """TEST INFRASTRUCTURE: the product sources compiled by g++ (tests/hostsim), so the CPU suite can
    exercise the stage logic + host orchestration without a GPU.  Never used by the package itself."""
    from llm_map_reduce_summarizer_b200 import build, engine
    lib = build.build_hostsim()
    e = engine.Engine(0, lib)
    yield e
    e.close()


@pytest.fixture(scope="session")
def gpu_engine():
    import torch
    assert torch.cuda.is_available(), "-m gpu tests need a CUDA device"
    from llm_map_reduce_summarizer_b200 import build, engine
    build.build()
    e = engine.Engine(0)
    yield e
    e.close()
