"""Random segments through stage P on the GPU: cleaned text and token counts per segment against the oracle.
Shapes that steer the fused kernel through its paths: odd whitespace (tabs, NBSP, U+2003, U+3000, runs of them),
multi-byte words, duplicate chains, `.X` insertions, special-token fragments, empty and whitespace-only segments, and
segment lengths from 0 to several KB (the tile size follows the batch's mean segment length; segments beyond a tile take
the slow lane, beyond 4 KB the pre-pass for oversized segments)."""
import random

import numpy as np
import pytest

from llm_map_reduce_summarizer_b200 import engine as E
from oracle import preprocess as OP, tokenizer as TK

pytestmark = pytest.mark.gpu
ALPHA = ["a", "b", "the", "the", "is", "Z", "é", "1", " ", " ", " ", "  ", "\t", " ", " ", "　", ".", ",", "!", "?", "-",
         "你好", "_", "x0", "​", "word", "end.Next", "café", " \t ", "́", "<|", "'s"]


def _run(eng, segs):
    enc = TK.encoding()
    b = E.pack_documents([segs])
    eng.preprocess(b, merge_same_speaker=False, want_merged_text=False)
    P = eng.fetch_processed()
    clean, coff, k2r, ctok = P["clean"].tobytes(), P["coff"], P["k2r"], P["ctok"]
    want = [(i, OP.clean_text(s["text"])) for i, s in enumerate(segs)]
    want = [(i, t) for i, t in want if t]
    assert [i for i, _ in want] == k2r.tolist()
    for j, (i, t) in enumerate(want):
        got = clean[int(coff[j]):int(coff[j + 1])]
        assert got == t.encode(), (i, segs[i]["text"][:80])
        assert int(ctok[j]) == len(enc.encode_ordinary(" " + t)), (i, t[:80])


@pytest.mark.parametrize("seed", [1, 2, 3])
def test_random_segments(gpu_engine, seed):
    rng = random.Random(seed)
    for trial in range(12):
        nseg = rng.choice([1, 3, 40, 200, 700])
        segs = []
        for i in range(nseg):
            L = rng.choice([0, 1, 5, 30, 60, 200, 900]) if rng.random() < 0.3 else rng.randint(0, 80)
            segs.append({"start": float(i), "end": float(i + 1), "text": "".join(rng.choice(ALPHA) for _ in range(L)), "speaker": "A"})
        _run(gpu_engine, segs)


@pytest.mark.parametrize("mean_len", [400, 1500, 6000])
def test_long_segments(gpu_engine, mean_len):
    """Batches whose segments are hundreds to thousands of bytes: smaller tiles, the slow lane, oversized segments."""
    rng = random.Random(mean_len)
    segs = []
    for i in range(60):
        L = max(1, int(rng.gauss(mean_len, mean_len / 3)))
        segs.append({"start": float(i), "end": float(i + 1), "text": "".join(rng.choice(ALPHA) for _ in range(L // 3)), "speaker": "A"})
    segs.insert(7, {"start": 0.0, "end": 1.0, "text": " \t ", "speaker": "A"})
    segs.insert(9, {"start": 0.0, "end": 1.0, "text": "", "speaker": "A"})
    _run(gpu_engine, segs)


def _tail_batch(seed, n, giant_every=0):
    """Mostly short segments with a heavy tail up to ~4 KB (no fixed tile size fits everywhere), optionally an oversized
    segment every `giant_every` segments."""
    rng = random.Random(seed)
    segs = []
    for i in range(n):
        u = rng.random()
        L = rng.randint(0, 25) if u < 0.9 else (rng.randint(200, 1300) if u < 0.995 else rng.randint(1300, 1360))
        if giant_every and i % giant_every == giant_every // 2:
            L = rng.randint(1500, 9000)
        segs.append({"start": float(i), "end": float(i + 1), "text": "".join(rng.choice(ALPHA) for _ in range(L)), "speaker": "A"})
    return segs


@pytest.mark.parametrize("giant_every", [0, 97])
def test_tiles_of_varying_size(gpu_engine, giant_every):
    """Heavy-tailed lengths, and oversized segments between short ones: the pre-pass finds no fixed tile size that fits
    everywhere (or sees an oversized segment) and the fused kernel runs on tiles cut by byte windows (k_tile_flags)."""
    _run(gpu_engine, _tail_batch(11 + giant_every, 6000, giant_every))


@pytest.mark.parametrize("tsegs", ["128", "40", "3"])
def test_forced_tile_sizes_take_the_slow_lane(tsegs, monkeypatch):
    """TRCHUNK_TSEGS pins the tile size: with 128 segments per tile most tiles of this batch exceed the tile buffer and go
    through the slow lane (thread-per-segment routines), which the automatic choice avoids."""
    monkeypatch.setenv("TRCHUNK_TSEGS", tsegs)
    eng = E.Engine(0)
    try:
        _run(eng, _tail_batch(5, 3000, 0))
        _run(eng, _tail_batch(6, 1500, 200))
    finally:
        eng.close()


def test_oversized_segments_cut_and_uncut(monkeypatch):
    """Oversized segments reach the fused kernel as parts cut in front of whitespace runs (default), or pre-cleaned by the
    byte-parallel pre-pass (TRCHUNK_GIANT_CUT=0, and any batch with a segment that cannot be cut: here 9 KB of one repeated
    word, and 6 KB without ASCII whitespace)."""
    rng = random.Random(77)
    base = _tail_batch(21, 1200, 150)
    words = ["alpha", "beta", "beta", "Gamma.", "x", "the", "the", "end.Next", "café", "你好", "3.5", "don't"]
    for k in range(6):
        base.insert(100 + 150 * k, {"start": 0.0, "end": 1.0, "speaker": "A",
                                    "text": " ".join(rng.choice(words) for _ in range(rng.randint(1500, 9000)))})
    uncuttable = [dict(base[0], text="same " * 1800), dict(base[0], text="你好　" * 900 + "tail")]
    for env, extra in (("1", []), ("0", []), ("1", uncuttable)):
        monkeypatch.setenv("TRCHUNK_GIANT_CUT", env)
        eng = E.Engine(0)
        try:
            _run(eng, base[:400] + extra + base[400:])
            assert eng.counts()["n_giants"] >= 6
        finally:
            eng.close()


def test_special_token_inside_cut_segment(monkeypatch):
    """A special-token literal inside an oversized segment: the parts of a cut segment report their first hit with its
    position, the record keeps the earliest (tiktoken raises for the first literal in the text).  Two documents with the
    two literals in opposite order, cut and uncut: DOC_ERR = 1 + index of the literal that comes first."""
    from llm_map_reduce_summarizer_b200 import vocab
    filler = " ".join(["word", "other", "thing."] * 700)                     # ~13 KB: several parts
    a, b = vocab.SPECIALS[0], vocab.SPECIALS[1]
    docs = [[{"start": 0.0, "end": 9.0, "speaker": "A", "text": filler + " " + x + " " + filler + " " + y + " " + filler}]
            for x, y in ((a, b), (b, a))]
    docs.append([{"start": 0.0, "end": 9.0, "speaker": "A", "text": filler}])
    for env in ("1", "0"):
        monkeypatch.setenv("TRCHUNK_GIANT_CUT", env)
        eng = E.Engine(0)
        try:
            eng.preprocess(E.pack_documents(docs), want_merged_text=False)
            assert eng.counts()["n_giants"] == 3
            assert eng.array("DOC_ERR", "<i4").tolist() == [1, 2, 0], env
        finally:
            eng.close()


def test_cut_segment_through_the_word_cap_fallback(gpu_engine):
    """Oversized segments of one-letter words: a tile holds more words than the fused kernel's word list, so it counts per
    segment (continuing parts without the virtual space, added to the record of the part that started the segment), and
    a 200-letter span inside such a part reaches the long-span kernel by its absolute position."""
    rng = random.Random(5)
    one = " ".join(rng.choice("abcdefgh") for _ in range(4000))
    segs = [{"start": 0.0, "end": 1.0, "speaker": "A", "text": "short one"},
            {"start": 1.0, "end": 2.0, "speaker": "A", "text": one + " " + "qzx" * 70 + " " + one},
            {"start": 2.0, "end": 3.0, "speaker": "B", "text": "x " * 3000 + "end"},
            {"start": 3.0, "end": 4.0, "speaker": "B", "text": "tail segment"}]
    _run(gpu_engine, segs)
    assert gpu_engine.counts()["n_giants"] == 2 and gpu_engine.counts()["n_long_spans"] >= 1
