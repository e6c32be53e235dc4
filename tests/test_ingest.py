"""trc_ingest_json against json.loads + engine.pack_documents, i.e. against what main.py:337-339
followed by the drop-in's own packing produces.  Host code: runs on the CPU test build."""
import json
import random

import numpy as np
import pytest

import conftest
import fuzzgen
from llm_map_reduce_summarizer_b200 import _native as N, engine as E


@pytest.fixture(scope="module")
def lib():
    return N.load_library(conftest.hostsim_library())


def same_batch(a, b):
    assert a.text.tobytes() == b.text.tobytes()
    for k in ("seg_off", "start", "end", "speaker", "doc_off"):
        x, y = getattr(a, k), getattr(b, k)
        assert x.shape == y.shape and (x == y).all(), k
    assert a.speakers == b.speakers
    sa, ea = a.originals
    sb, eb = b.originals
    assert [(type(v), v) for v in sa] == [(type(v), v) for v in sb]
    assert [(type(v), v) for v in ea] == [(type(v), v) for v in eb]


def variants(rng, segs):
    """The same transcript serialised in different legal ways."""
    doc = {"segments": segs}
    if rng.random() < 0.5:
        doc = {"language": "en", "meta": {"a": [1, 2.5e-3, {"b": None}], "segments": "not this one"}, "segments": segs,
               "word_segments": [{"word": "x", "start": 1, "end": 2, "score": 0.5}], "ok": True, "n": -0.0}
    kw = dict(ensure_ascii=rng.random() < 0.5)
    if rng.random() < 0.3:
        kw["indent"] = rng.choice([1, 2, "\t"])
    elif rng.random() < 0.5:
        kw["separators"] = (",", ":")
    return json.dumps(doc, **kw)


def test_fuzz_against_json_loads(lib):
    rng = random.Random(321)
    docs, raws = [], []
    for it in range(120):
        segs = fuzzgen.rand_transcript(rng, long_p=0.05)
        for s in segs:
            r = rng.random()
            if r < 0.10:
                s["start"] = int(s["start"])
            elif r < 0.13:
                s["end"] = True
            elif r < 0.16:
                del s["end"]
            elif r < 0.19 and "speaker" in s:
                s["speaker"] = rng.choice(["Zoë", "说话人", "a\"b\\c", "\U0001F600 host", "tab\tname"])
            if rng.random() < 0.05:
                s["text"] = s["text"] + rng.choice(["\U0001F600", " ", "\x7f", "\\u0041", "\"q\"", "/", "\b\f\n\r\t", "é你"])
            if rng.random() < 0.05:
                s["words"] = [{"word": "w", "start": 0.1, "end": 0.2, "probability": 1e-7}, [], {}, "s", None]
            if rng.random() < 0.03:
                del s["text"]
        raw = variants(rng, segs)
        raws.append(raw)
        docs.append(json.loads(raw)["segments"])
    got = E.ingest_json(raws, n_threads=3, lib=lib)
    same_batch(got, E.pack_documents(docs))
    # single documents, bytes input, UTF-8 signature
    for raw, segs in list(zip(raws, docs))[:20]:
        same_batch(E.ingest_json([b"\xef\xbb\xbf" + raw.encode("utf-8")], lib=lib), E.pack_documents([segs]))


def test_hand_written_documents(lib):
    cases = [
        '{}', '{"segments": []}', '{"segments": {}}', '{"segments": { }, "x": 1}', ' \n{"a": 1, "segments" : [ ] }\t', '{"segments": [{}]}',
        '{"segments": [{"text": "a", "text": "b", "start": 1, "start": 2.5}]}',           # repeated keys: the last one wins
        '{"segments": [{"text": "x"}], "segments": [{"text": "y", "speaker": "S"}]}',
        '{"segments": [{"start": 1e3, "end": 1E+2, "text": "\\ud83d\\ude00 \\u00e9 \\/"}]}',
        '{"segments": [{"start": -0, "end": -0.0, "text": "t"}, {"start": 9007199254740992, "end": -9007199254740992}]}',
        '{"segments": [{"start": 0.1, "end": 25.520000000000003, "text": "t"}, {"start": 1.7976931348623157e308, "end": 5e-324}]}',
        '{"segments": [{"start": false, "end": true, "text": "b"}]}',
    ]
    for raw in cases:
        same_batch(E.ingest_json([raw], lib=lib), E.pack_documents([json.loads(raw).get("segments", [])]))


def test_rejected_documents(lib):
    """Everything here fails in the reference too: in json.load, or at the first use of the field."""
    bad = [
        '', '[]', '{"segments": {"a": 1}}', '{"segments": null}', '{"segments": [1]}', '{"segments": [[]]}', '{"segments": [{"text": null}]}',
        '{"segments": [{"text": 5}]}', '{"segments": [{"speaker": 5}]}', '{"segments": [{"start": null}]}', '{"segments": [{"start": "x"}]}',
        '{"segments": [{"start": NaN}]}', '{"segments": [{"end": Infinity}]}', '{"segments": [{"end": -Infinity}]}',
        '{"segments": [{"start": 1e999}]}',
        '{"segments": [{"text": "a\\ud800b"}]}', '{"segments": [{"text": "a\nb"}]}', '{"segments": [{"text": "a\\xb"}]}',
        '{"segments": [{"text": "a"},]}', '{"segments": [{"text": "a",}]}', '{"segments": [{"text": "a"}]} x', '{"segments": [{"start": 01}]}',
        '{"segments": [{"start": 1.}]}', '{"segments": [{"start": .5}]}', '{"segments": [{"start": 1e}]}', "{'segments': []}",
        '{"segments": [{"text": "a"}', '{"a": tru}', '{"segments": [{"text": "\\u12"}]}',
    ]
    for raw in bad:
        with pytest.raises(ValueError):
            E.ingest_json([raw], lib=lib)
        with pytest.raises(Exception):                       # the reference's ingest + first use
            segs = json.loads(raw).get("segments", [])
            b = E.pack_documents([segs])
            for v in b.originals[0] + b.originals[1]:
                int(v)
    for raw in ('{"segments": [{"start": 9007199254740993}]}', '{"segments": [{"end": 123456789012345678901234}]}',
                '{"segments": [{"start": "1"}]}'):
        with pytest.raises(ValueError):                      # fine for the reference, not for a float64 layout
            E.ingest_json([raw], lib=lib)
    for rawb in (b'{"segments": [{"text": "\xff"}]}', b'{"segments": [{"text": "\xc0\xaf"}]}', b'{"segments": [{"text": "\xed\xa0\x80"}]}'):
        with pytest.raises(ValueError):
            E.ingest_json([rawb], lib=lib)
        with pytest.raises(UnicodeDecodeError):               # main.py:337 opens the file with encoding='utf-8'
            json.loads(rawb.decode("utf-8"))


def test_ingested_batch_through_the_path(hostsim_engine):
    """JSON text -> ingest -> stage P -> stage C -> the reference's dicts, against the oracle on json.loads' segments
    (int / bool times keep their Python types in the output)."""
    from oracle import chunker as OC, preprocess as OP
    import parity_common as pc
    rng = random.Random(5)
    raws = []
    for it in range(12):
        segs = fuzzgen.rand_transcript(rng, long_p=0.05)
        for s in segs:
            if rng.random() < 0.2:
                s["start"] = int(s["start"])
        raws.append(json.dumps({"segments": segs, "language": "en"}, ensure_ascii=it % 2 == 0))
    raws.append(json.dumps({"segments": conftest.load_golden("example_segments.json.gz")}))
    b = E.ingest_json(raws, n_threads=2, lib=hostsim_engine.lib)
    mt = np.full(b.n_doc, 400, dtype=np.int32)
    mt[-1] = 4000
    hostsim_engine.preprocess(b)
    hostsim_engine.chunk(mt)
    P, R = hostsim_engine.fetch_processed(), hostsim_engine.fetch_chunks()
    for x, raw in enumerate(raws):
        op = OP.preprocess(json.loads(raw)["segments"])
        pp = E.build_processed(b, P, x)
        assert pc.same(pp, op), x
        try:
            oc = OC.ChunkerOracle(int(mt[x])).chunk(op)
        except ValueError:
            continue
        assert pc.same(E.build_chunks(hostsim_engine, b, P, R, pp, x), oc), x


def test_mutated_documents_are_never_misread(lib):
    """Soundness under random damage: whenever the ingest ACCEPTS a mutated document, json.loads accepts it too and
    the packed batch is identical (it may refuse more than Python does, never read something else)."""
    rng = random.Random(2024)
    base = []
    for it in range(25):
        segs = fuzzgen.rand_transcript(rng, nseg=rng.randint(1, 6))
        for s in segs:
            if rng.random() < 0.3:
                s["start"] = int(s["start"])
            if rng.random() < 0.2:
                s["extra"] = {"k": [1, "two", None, True, 3.5e-2]}
        base.append(variants(rng, segs))
    alphabet = list('{}[]",:\\ntrufalse0123456789.-+eE \n') + ["\\u00e9", "\\ud83d", "\\ude00", "é", "null", "true", '"text"', '"start"', '"segments"']
    accepted = 0
    for trial in range(4000):
        raw = rng.choice(base)
        for _ in range(rng.randint(1, 3)):
            k = rng.randrange(len(raw) + 1)
            r = rng.random()
            if r < 0.35:
                raw = raw[:k] + raw[k + rng.randint(1, 4):]
            elif r < 0.7:
                raw = raw[:k] + rng.choice(alphabet) + raw[k:]
            else:
                raw = raw[:k] + rng.choice(alphabet) + raw[k + 1:]
        try:
            got = E.ingest_json([raw], lib=lib)
        except ValueError:
            continue
        accepted += 1
        segs = json.loads(raw).get("segments", [])          # must not raise
        same_batch(got, E.pack_documents([list(segs)]))
    assert accepted > 300
