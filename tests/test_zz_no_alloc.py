"""The allocation-free span counter (trc_bpe_span_count_static) against the regular entry point.

Collected last on purpose: if this extra entry point misbehaves, a `-x` run has already exercised every
other test.  Who needs it: embedders that may not allocate GPU memory while they run (their own pool owns
the device, or a cudaMalloc stall is unacceptable); the regular path may grow the context arena.
The comparison is exact (integer token counts), the same bar as every other parity test here.
"""
import random

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _letters(rng, n):
    # one run of ASCII letters == exactly one pre-tokeniser span (alt 2, \p{L}+), long enough not to be a token
    return "".join(rng.choice("abcdefghijklmnopqrstuvwxyz") for _ in range(n))


def test_static_span_counter_matches_regular_entry_point(gpu_engine):
    import torch
    eng = gpu_engine
    lib = eng.lib
    rng = random.Random(4242)
    spans = [_letters(rng, 3000), _letters(rng, 50000), "the" * 700, _letters(rng, 49), "hello"]
    want, _ = eng.count_tokens(spans, leading_space=False)          # existing entry point (arena scratch)
    dev = torch.device("cuda", eng.device)
    out = torch.zeros(1, dtype=torch.int32, device=dev)
    stream = torch.cuda.current_stream().cuda_stream
    # two rounds back to back over the same shared work area: stale contents must not leak between calls
    for rnd in range(2):
        for s, w in zip(spans, want):
            buf = torch.from_numpy(np.frombuffer(s.encode(), dtype=np.uint8).copy()).to(dev)
            rc = lib.trc_bpe_span_count_static(eng.h, buf.data_ptr(), buf.numel(), out.data_ptr(), stream)
            assert rc == 0, lib.trc_last_error(eng.h)
            torch.cuda.synchronize()
            assert int(out.item()) == int(w), (rnd, len(s), int(out.item()), int(w))


def test_static_span_counter_rejects_oversize(gpu_engine):
    import torch
    eng = gpu_engine
    dev = torch.device("cuda", eng.device)
    buf = torch.zeros(65537, dtype=torch.uint8, device=dev)
    out = torch.zeros(1, dtype=torch.int32, device=dev)
    rc = eng.lib.trc_bpe_span_count_static(eng.h, buf.data_ptr(), 65537, out.data_ptr(), None)
    assert rc == -2
    assert b"65536" in eng.lib.trc_last_error(eng.h)
    assert eng.lib.trc_bpe_span_count_static(eng.h, buf.data_ptr(), 0, out.data_ptr(), None) == -2
