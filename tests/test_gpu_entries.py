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


def test_save_chunks_records_and_timestamps_cuda(gpu_engine, tmp_path):
    """--save-chunks records (main.py:178-201) from the CUDA engine's chunks equal those from the oracle's chunks;
    format_timestamp / _format_time run as one small kernel and leave the resident batch resident."""
    import json
    from conftest import load_golden, quiet
    from llm_map_reduce_summarizer_b200 import big_chunkeroosky as BC, egress, preprocessor as PP
    from oracle import chunker as OC, preprocess as OP
    E.set_default_engine(gpu_engine)
    segs = load_golden("example_segments.json.gz")
    with quiet():
        proc = PP.preprocess_transcript(segs)
        serial = gpu_engine.batch_serial
        c = BC.BigChunkeroosky(max_tokens_per_chunk=2000)
        assert [c._format_time(x) for x in (0, 59.999, 3600, -1, 360000)] == ["00:00", "00:59", "01:00:00", "59:59", "100:00:00"]
        assert PP.clean_text("  so  so it   is.Then ") == "so it is. Then"
        assert gpu_engine.batch_serial == serial                      # the helpers did not evict the batch
        chunks = c.postprocess_chunks(c.chunk_transcript(proc))
    want = OC.ChunkerOracle(2000).chunk(OP.preprocess(segs))
    a, b = tmp_path / "a.json", tmp_path / "b.json"
    egress.save_chunk_records(chunks, str(a), timestamp="t")
    egress.save_chunk_records(want, str(b), timestamp="t")
    assert open(a).read() == open(b).read()
    assert len(json.load(open(a))["chunks"]) == len(want) > 10
    assert gpu_engine.format_timestamps([1e6, 26561.26, True]) == ["277:46:40", "07:22:41", "00:01"]


def test_multi_gpu_single_process(gpu_engine):
    """shard.MultiGpu: one process, every visible GPU, documents cut by bytes; the shards' results are the single-GPU
    results of the same documents."""
    import torch
    from llm_map_reduce_summarizer_b200 import shard as SH, synth
    b, mt, _ = synth.config(5, 0.01)
    ndev = torch.cuda.device_count()
    mg = SH.MultiGpu(devices=list(range(ndev)) if ndev > 1 else [0, 0])      # two contexts on one GPU when there is only one
    try:
        res = mg.run(b, mt, fetch=True)
        assert [r["docs"] for r in res] == SH.shard_ranges(b, len(res)) and res[0]["docs"][0] == 0 and res[-1]["docs"][1] == b.n_doc
        gpu_engine.preprocess(b)
        gpu_engine.chunk(mt)
        whole = gpu_engine.fetch_chunks()["chunks"]
        got = np.concatenate([r["chunks"]["chunks"]["token_count"] for r in res if "chunks" in r])
        assert (got == whole["token_count"]).all()
        assert sum(r["counts"]["tokens"] for r in res if "counts" in r) == gpu_engine.counts()["tokens"]
    finally:
        mg.close()


def test_published_punkt_examples_through_the_sentence_kernels(gpu_engine):
    """The nltk examples of tests/test_oracle_golden.py as over-budget segments: the CUDA sentence path
    (k_punkt_breaks and the sentence greedy) must cut them where the restated Punkt does."""
    import parity_common as pc
    from oracle import chunker as OC, preprocess as OP
    texts = ['(How does it deal with this parenthesis?)  "It should be part of the\nprevious sentence." "(And the same with this one.)" '
             '(\'And this one!\')\n"(\'(And (this)) \'?)" [(and this. )]',
             "this is a test. . new sentence.", "This. . . That", "This..... That", "This... That", "This.. . That", "This. .. That",
             "This. ,. That", "This!!! That", "This! That", "1. This is R .\n2. This is A .\n3. That's all",
             "1. This is R .\t2. This is A .\t3. That's all", "Hello.\tThere", "Very bad acting!!! I promise.",
             "Punkt knows that the periods in Mr. Smith and Johann S. Bach\ndo not mark sentence boundaries.  And sometimes sentences\n"
             "can start with non-capitalized words.  i is a good variable\nname."]
    for t in texts:
        segs = [{"start": 0.0, "end": 60.0, "text": t, "speaker": "A"}]
        for mx in (153, 156, 170):
            proc, chunks = pc.run_dropin(gpu_engine, segs, mx)
            want = OC.ChunkerOracle(mx).chunk(OP.preprocess(segs))
            assert pc.same(chunks, want), (t, mx)


def test_long_spans_batched_merge(gpu_engine):
    """Spans far beyond the thread-local limit (k_bpe_long): one round per distinct rank.  Letters without a separator
    (one pre-tokeniser span), a base64-like blob, runs of one letter (chains of overlapping equal pairs), against the real
    CoreBPE; a 64 KB span of random letters takes ~30 ms through trc_count_tokens (22 ms of it the single thread that walks the
    span for the pre-tokeniser, 8 ms the merges), where one merge per block-wide reduction took 183 ms."""
    import time
    enc = TK.encoding()
    rng = random.Random(9)
    b64 = "ABCDEFGHIJKLMNOPQRSTUVWXYZabcdefghijklmnopqrstuvwxyz"
    strs = ["a" * 4097, "ab" * 3000, "aaab" * 1500, "".join(rng.choice("etaoinshr") for _ in range(20000)),
            "".join(rng.choice(b64) for _ in range(65536)), "the" * 9000, "z" * 65000]
    cnt, _ = gpu_engine.count_tokens(strs, leading_space=False)
    want = [len(enc.encode_ordinary(s)) for s in strs]
    assert [int(c) for c in cnt] == want
    one = ["".join(rng.choice(b64) for _ in range(65536))]
    gpu_engine.count_tokens(one)
    t0 = time.perf_counter()
    c1, _ = gpu_engine.count_tokens(one)
    dt = time.perf_counter() - t0
    assert int(c1[0]) == len(enc.encode_ordinary(one[0]))
    assert dt < 0.1, dt


def test_sentence_counts_over_a_cut_segment(gpu_engine):
    """An over-budget monologue of 30 KB goes through the sentence stage with token counts taken from the per-word token
    bytes of the cut-mode fused kernel; a 300-letter span (its count is finished later by the long-span kernel, so its byte
    says "not known") and sentences of several KB (counted directly) sit in the middle.  Chunks against the oracle."""
    import parity_common as pc
    from oracle import chunker as OC, preprocess as OP
    rng = random.Random(12)
    words = ["alpha", "beta", "Gamma", "delta", "x", "the", "Mr.", "3.5", "café", "don't", "end"]
    sents = []
    for i in range(900):
        n = rng.randint(3, 14) if i % 97 else 900                       # every 97th sentence is several KB long
        sents.append(" ".join(rng.choice(words) for _ in range(n)).capitalize() + rng.choice([".", "!", "?"]))
    sents.insert(400, "Before " + "qzxv" * 75 + " after.")
    text = " ".join(sents)
    assert len(text.encode()) > 20000
    segs = [{"start": 0.0, "end": 30.0, "text": "Intro words here.", "speaker": "B"},
            {"start": 30.0, "end": 3000.0, "text": text, "speaker": "A"},
            {"start": 3000.0, "end": 3010.0, "text": "Closing words.", "speaker": "B"}]
    for mx in (400, 1000):
        proc, chunks = pc.run_dropin(gpu_engine, segs, mx)
        want = OC.ChunkerOracle(mx).chunk(OP.preprocess(segs))
        assert pc.same(chunks, want), mx
    assert gpu_engine.counts()["n_giants"] == 1 and gpu_engine.counts()["n_sent"] > 800


def test_long_span_in_a_speaker_name(gpu_engine):
    """The long-span check of the speaker strings rides on the fused kernel's pre-pass round trip: a 70-letter speaker name
    (one pre-tokeniser span beyond the thread-local limit) must still be counted by the long-span kernel."""
    import parity_common as pc
    from oracle import chunker as OC, preprocess as OP
    name = "qxzjvkw" * 10
    segs = [{"start": float(i), "end": float(i + 1), "text": "some words to count here %d" % i, "speaker": name if i % 2 else "B"}
            for i in range(40)]
    proc, chunks = pc.run_dropin(gpu_engine, segs, 400)
    want = OC.ChunkerOracle(400).chunk(OP.preprocess(segs))
    assert pc.same(chunks, want)
    assert gpu_engine.counts()["n_long_spans"] >= 1
