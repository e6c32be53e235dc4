"""The entry points that the CPU suite only exercises through the g++ test build (tests/hostsim), run here through
libtrchunk.so on the GPU: the same test bodies, the CUDA engine instead of the host build.  The host build does
not contain the fused clean+tokenise kernel, so without these the CUDA product path of these entries is untested."""
import random

import numpy as np
import pytest

from llm_map_reduce_summarizer_b200 import engine as E
from oracle import tokenizer as TK
from test_hostsim_parity import test_combine_helpers as _combine_helpers, test_token_counter_service as _token_counter
from test_ingest import test_ingested_batch_through_the_path as _ingested_batch

pytestmark = pytest.mark.gpu


def test_combine_helpers_cuda(gpu_engine):
    """combine_same_speaker_segments / create_combined_segment (preprocessor.py:109-215)."""
    _combine_helpers(gpu_engine)


def test_token_counter_service_cuda(gpu_engine):
    """TokenCounter (result_aggregator.py:357-392's counting) through trc_count_tokens."""
    _token_counter(gpu_engine)


def test_ingested_batch_through_the_path_cuda(gpu_engine):
    """trc_ingest_json -> trc_preprocess -> trc_chunk on the GPU, against the oracle on json.loads' segments."""
    _ingested_batch(gpu_engine)


def test_long_span_worklist_bound(gpu_engine):
    """Every string is 48 letters that no vocabulary entry covers; with the virtual leading space each becomes a
    49-byte span, i.e. one long span PER string - more than total_bytes/49 of them.  The work list is sized from the
    logical bytes (stored bytes + one per unit), so none is dropped."""
    enc = TK.encoding()
    rng = random.Random(3)
    for n in (50, 64, 257):
        strs = ["".join(rng.choice("qxzjvkw") for _ in range(48)) for _ in range(n)]
        for lead in (True, False):
            cnt, _ = gpu_engine.count_tokens(strs, leading_space=lead)
            assert [int(c) for c in cnt] == [len(enc.encode_ordinary((" " if lead else "") + s)) for s in strs]
    # the same through stage P (segments of one 48-letter word each: virtual leading space on every kept segment)
    segs = [{"start": float(i), "end": float(i + 1), "text": "".join(rng.choice("qxzjvkw") for _ in range(48)), "speaker": "A"}
            for i in range(300)]
    b = E.pack_documents([segs])
    gpu_engine.preprocess(b, merge_same_speaker=False)
    P = gpu_engine.fetch_processed()
    clean, coff = P["clean"].tobytes(), P["coff"]
    want = [len(enc.encode_ordinary(" " + clean[int(coff[j]):int(coff[j + 1])].decode())) for j in range(len(segs))]
    assert P["ctok"].tolist() == want
    assert gpu_engine.counts()["n_long_spans"] >= 300


def test_single_speaker_uncapped_is_linear(gpu_engine):
    """max_segment_duration=None on one 200k-segment single-speaker document: group starts come from the speaker
    changes alone (no forward scan per segment), so this finishes in milliseconds, not minutes."""
    import time
    n = 200000
    text = np.frombuffer(b"ok " * n, dtype=np.uint8).copy()
    seg_off = np.arange(0, 3 * n + 1, 3, dtype=np.int64)
    start = np.arange(n, dtype=np.float64)
    b = E.PackedBatch(text, seg_off, start, start + 1.0, np.ones(n, dtype=np.int32), np.array([0, n], dtype=np.int64), ["", "A"], None)
    gpu_engine.preprocess(b, max_segment_duration=None, want_merged_text=False)      # warm-up (arena growth)
    t0 = time.perf_counter()
    gpu_engine.preprocess(b, max_segment_duration=None, want_merged_text=False)
    c = gpu_engine.counts()
    dt = time.perf_counter() - t0
    assert c["n_groups"] == 1 and c["n_kept"] == n
    assert dt < 0.5, dt
