"""Synthetic transcripts of the shapes BASELINE.json names, generated directly in the
packed layout of the C ABI (numpy PCG64; seed = 20261106 + config number).

Words follow the example transcript's unigram distribution (data/lexicon.json, derived from
transcript-example.json by tools/train_vocab.py); segments end with '.' 52 % / '?' 7 % / ',' 6 %;
artefacts: duplicated word 2 %, missing space after '.' 1 %, odd whitespace 5 %, non-ASCII word 1 %.
Times: durations rounded to 0.01 s, start_i = round(cumsum, 2), end_i = start_{i+1}.
"""
import json
import os

import numpy as np

from .engine import PackedBatch

_DATA = os.path.join(os.path.dirname(os.path.abspath(__file__)), "data")
_NONASCII = ["café", "naïve", "über", "señor", "你好", "世界", "déjà", "jalapeño"]
_SEPS = [" ", "  ", " ", "\t", " \t "]
_ENDS = ["", ".", "?", ","]
BASE_SEED = 20261106


class _Atoms:
    def __init__(self, words=None, p=None):
        if words is None:
            lex = json.load(open(os.path.join(_DATA, "lexicon.json")))
            words = lex["words"]
            p = np.array(lex["counts"], dtype=np.float64)
        self.words = words
        self.p = p / p.sum()
        atoms = [w.encode() for w in self.words]
        self.na0 = len(atoms)
        atoms += [w.encode() for w in _NONASCII]
        self.sep0 = len(atoms)
        atoms += [s.encode() for s in _SEPS]
        self.empty = len(atoms)
        atoms.append(b"")
        self.end0 = len(atoms)
        atoms += [s.encode() for s in _ENDS]
        self.dot_nospace = len(atoms)
        atoms.append(b".")
        self.len = np.array([len(a) for a in atoms], dtype=np.int64)
        self.off = np.concatenate([[0], np.cumsum(self.len)])
        self.bytes = np.frombuffer(b"".join(atoms), dtype=np.uint8)


_ATOMS = None


def _atoms():
    global _ATOMS
    if _ATOMS is None:
        _ATOMS = _Atoms()
    return _ATOMS


_OPEN = {}


def open_lexicon(n_forms, seed=BASE_SEED):
    """An OPEN vocabulary of n_forms distinct word forms for the sensitivity bench (bench.py --lexicon N): the example's
    5 390 words plus derived forms - capitalised / upper-cased, inflected, compounded, with digits (years, amounts,
    identifiers), and 5 % non-ASCII (accents, CJK) - under a Zipf law, so that most running words are still common ones
    but the tail never ends: many words miss the word cache and a good share needs byte-pair merging."""
    key = (n_forms, seed)
    if key in _OPEN:
        return _OPEN[key]
    base = _atoms().words
    rng = np.random.Generator(np.random.PCG64(seed + 7))
    forms, seen = list(base), set(base)
    suffixes = ["s", "ed", "ing", "ly", "er", "ness", "able", "ism", "ize", "tion"]
    accents = {"a": "á", "e": "é", "i": "ï", "o": "ö", "u": "ü", "n": "ñ", "c": "ç"}
    cjk = "的一是不了人我在有他这为之大来以个中上们到说国和地也子时道出而要于就下得可你年生自会那后能对着事其里所去行过家十用发天如然作方成者多日都三小军二无同么经法当起与好看学进种将还分此心前面又定见只主没公从"
    while len(forms) < n_forms:
        k = int(rng.integers(0, 8))
        w = base[int(rng.integers(0, len(base)))].strip(",.;:?!'\"") or "uh"
        if k == 0:
            f = w.capitalize()
        elif k == 1:
            f = w.upper()
        elif k == 2:
            f = w + suffixes[int(rng.integers(0, len(suffixes)))]
        elif k == 3:
            f = w + base[int(rng.integers(0, len(base)))].strip(",.;:?!'\"")
        elif k == 4:
            f = str(int(rng.integers(0, 10 ** int(rng.integers(1, 8)))))
        elif k == 5:
            f = w + str(int(rng.integers(0, 1000)))
        elif k == 6:
            f = "".join(accents.get(ch, ch) if rng.random() < 0.5 else ch for ch in w)
            if f == w:
                f = w + "é"
        else:
            f = "".join(cjk[int(rng.integers(0, len(cjk)))] for _ in range(int(rng.integers(1, 4))))
        if rng.random() < 0.15:
            f += [",", ".", "?", "!", ";", ":"][int(rng.integers(0, 6))]
        if f not in seen:
            seen.add(f); forms.append(f)
    p = 1.0 / (np.arange(len(forms), dtype=np.float64) + 10.0)
    order = np.concatenate([np.arange(len(base)), len(base) + rng.permutation(len(forms) - len(base))])
    forms = [forms[i] for i in order]
    _OPEN[key] = _Atoms(forms, p)
    return _OPEN[key]


def _gather(atom_ids, A):
    lens = A.len[atom_ids]
    dst = np.concatenate([[0], np.cumsum(lens)])
    total = int(dst[-1])
    src = np.repeat(A.off[atom_ids] - dst[:-1], lens) + np.arange(total, dtype=np.int64)
    return A.bytes[src], dst


def make_corpus(n_docs, seed, segs_per_doc=None, segs_mu=180.0, segs_sigma=0.6, words_lambda=9.8, n_speakers=2,
                turn_mean=8.0, sec_per_word=0.36, artefacts=True, atoms=None):
    """n_docs transcripts -> PackedBatch (originals=None: times are the float64 arrays themselves)."""
    A = atoms if atoms is not None else _atoms()
    rng = np.random.Generator(np.random.PCG64(seed))
    if segs_per_doc is None:
        nseg = np.maximum(1, np.round(rng.lognormal(np.log(segs_mu), segs_sigma, n_docs))).astype(np.int64)
    else:
        nseg = np.full(n_docs, segs_per_doc, dtype=np.int64)
    doc_off = np.concatenate([[0], np.cumsum(nseg)])
    n = int(doc_off[-1])
    nw = 1 + rng.poisson(words_lambda, n).astype(np.int64)
    woff = np.concatenate([[0], np.cumsum(nw)])
    W = int(woff[-1])
    wid = rng.choice(len(A.words), size=W, p=A.p).astype(np.int64)
    first = np.zeros(W, dtype=bool)
    first[woff[:-1]] = True
    last = np.zeros(W, dtype=bool)
    last[woff[1:] - 1] = True
    sep = np.full(W, A.sep0, dtype=np.int64)
    suffix = np.full(W, A.empty, dtype=np.int64)
    if artefacts:
        u = rng.random(W)
        dup = (u < 0.02) & ~first
        wid[dup] = wid[np.nonzero(dup)[0] - 1]
        na = (u >= 0.02) & (u < 0.03)
        wid[na] = A.na0 + rng.integers(0, len(_NONASCII), int(na.sum()))
        odd = (u >= 0.03) & (u < 0.08)
        sep[odd] = A.sep0 + rng.integers(1, len(_SEPS), int(odd.sum()))
        glue = (u >= 0.08) & (u < 0.09) & ~first       # "end.Next": previous word gets '.', no separator
        sep[glue] = A.dot_nospace
    sep[first] = A.empty
    e = rng.random(n)
    endk = np.where(e < 0.52, 1, np.where(e < 0.59, 2, np.where(e < 0.65, 3, 0)))
    suffix[woff[1:] - 1] = A.end0 + endk
    atoms = np.empty(3 * W, dtype=np.int64)
    atoms[0::3] = sep
    atoms[1::3] = wid
    atoms[2::3] = suffix
    text, aoff = _gather(atoms, A)
    seg_off = aoff[3 * woff]
    # times
    dur = np.round(nw * sec_per_word * rng.uniform(0.6, 1.4, n), 2)
    dur = np.maximum(dur, 0.01)
    start = np.empty(n, dtype=np.float64)
    end = np.empty(n, dtype=np.float64)
    cs = np.cumsum(dur)
    base = np.repeat(np.concatenate([[0.0], cs])[doc_off[:-1]], nseg)
    start[:] = np.round(np.concatenate([[0.0], cs[:-1]]) - base, 2)
    end[:-1] = start[1:]
    lastseg = doc_off[1:] - 1
    end[lastseg] = np.round(start[lastseg] + dur[lastseg], 2)
    # speakers: turns of geometric length
    change = rng.random(n) < (1.0 / turn_mean)
    change[doc_off[:-1]] = True
    turn = np.cumsum(change) - 1
    tspk = rng.integers(0, n_speakers, int(turn[-1]) + 1)
    names = [""] + ["SPEAKER_%02d" % k for k in range(n_speakers)]
    spk = (1 + tspk[turn]).astype(np.int32)
    return PackedBatch(text, seg_off, start, end, spk, doc_off, names, None)


def add_monologues(batch, seed, n_mono, min_bytes=20000, max_bytes=80000, clause_frac=0.10, nopunct_frac=0.02):
    """Replace n_mono evenly spaced segments of a single-document batch by 20-80 KB monologues
    (config 4): 10 % contain a >= 20 KB stretch without .?! (forces the clause path), 2 % have no
    ,.;:?! at all (forces the 20-word path).  Durations 300-1200 s (times re-accumulated)."""
    A = _atoms()
    rng = np.random.Generator(np.random.PCG64(seed))
    n = batch.n_seg
    pos = np.linspace(0, n - 1, n_mono + 2)[1:-1].astype(np.int64)
    texts = []
    for k in range(n_mono):
        target = int(rng.integers(min_bytes, max_bytes))
        u = rng.random()
        words = rng.choice(len(A.words), size=target // 4, p=A.p)
        parts = []
        size = 0
        sent_len = 0
        mode = 2 if u < nopunct_frac else (1 if u < nopunct_frac + clause_frac else 0)
        for i, w in enumerate(words):
            s = A.words[int(w)]
            if mode == 2:
                s = s.strip(",.;:?!") or "uh"
            elif mode == 1 and size < 24000:
                s = s.strip(".?!") or "uh"
                sent_len += 1
                if sent_len >= 12 and rng.random() < 0.2:
                    s += ","
                    sent_len = 0
            else:
                sent_len += 1
                if sent_len >= 8 and rng.random() < 0.15:
                    s = s.rstrip(",.;:?!") + ("." if rng.random() < 0.85 else "?")
                    sent_len = 0
            parts.append(s)
            size += len(s) + 1
            if size >= target:
                break
        texts.append(" ".join(parts).encode())
    # splice
    seg_off = batch.seg_off
    pieces, new_off, cur, prev = [], [0], 0, 0
    old = batch.text
    posset = {int(p): k for k, p in enumerate(pos)}
    lens = np.diff(seg_off)
    newlens = lens.copy()
    for p, k in posset.items():
        newlens[p] = len(texts[k])
    new_seg_off = np.concatenate([[0], np.cumsum(newlens)])
    out = np.empty(int(new_seg_off[-1]), dtype=np.uint8)
    order = sorted(posset)
    a = 0
    for p in order + [n]:
        if p > a:   # copy the untouched run [a, p)
            out[new_seg_off[a]:new_seg_off[p]] = old[seg_off[a]:seg_off[p]]
        if p < n:
            out[new_seg_off[p]:new_seg_off[p + 1]] = np.frombuffer(texts[posset[p]], dtype=np.uint8)
        a = p + 1
    dur = batch.end - batch.start
    dur[pos] = np.round(rng.uniform(300, 1200, n_mono), 2)
    cs = np.cumsum(dur)
    start = np.round(np.concatenate([[0.0], cs[:-1]]), 2)
    end = np.concatenate([start[1:], [np.round(start[-1] + dur[-1], 2)]])
    return PackedBatch(out, new_seg_off, start, end, batch.speaker, batch.doc_off, batch.speakers, None)


def config(num, scale=1.0, lexicon=None):
    """The BASELINE.json configs 2-5 (config 1 is the example transcript fixture).
    Returns (PackedBatch, per-document max_tokens array, description).  lexicon=N: config 3's shape with words drawn
    from an open vocabulary of N forms (open_lexicon) instead of the example's 5 390."""
    seed = BASE_SEED + num
    if lexicon:
        b = make_corpus(int(10000 * scale), seed, n_speakers=2, atoms=open_lexicon(int(lexicon)))
        return b, np.full(b.n_doc, 8000, dtype=np.int32), "10k podcast transcripts, open vocabulary of %d word forms, 8k-token chunks" % int(lexicon)
    if num == 2:
        b = make_corpus(1, seed, segs_per_doc=int(10000 * scale), words_lambda=3.0, n_speakers=4, turn_mean=8.0, sec_per_word=0.09)
        return b, np.full(1, 4000, dtype=np.int32), "1-hour 4-speaker meeting, 10k segments, 4k-token chunks"
    if num == 3:
        b = make_corpus(int(10000 * scale), seed, n_speakers=2)
        return b, np.full(b.n_doc, 8000, dtype=np.int32), "10k podcast transcripts (~100 MB), 8k-token chunks"
    if num == 4:
        n = int(1000000 * scale)
        b = make_corpus(1, seed, segs_per_doc=n, words_lambda=5.0, n_speakers=4, turn_mean=50000.0 if scale >= 0.5 else 5000.0,
                        sec_per_word=0.006)
        b = add_monologues(b, seed + 100, max(1, int(1000 * scale)))
        return b, np.full(1, 4000, dtype=np.int32), "10-hour transcript, 1M segments, 1k monologues of 20-80 KB"
    if num == 5:
        b = make_corpus(int(10000 * scale), seed, n_speakers=2)
        mt = np.array([2000, 8000, 32000], dtype=np.int32)[np.arange(b.n_doc) % 3]
        return b, mt, "corpus shard, mixed 2k/8k/32k chunk limits"
    raise ValueError(num)


def doc_to_segments(batch, x):
    """Document x as the list of dicts the reference's main.py would load."""
    a, b = int(batch.doc_off[x]), int(batch.doc_off[x + 1])
    tb = batch.text.tobytes()
    out = []
    for i in range(a, b):
        out.append({"start": float(batch.start[i]), "end": float(batch.end[i]),
                    "text": tb[int(batch.seg_off[i]):int(batch.seg_off[i + 1])].decode("utf-8"),
                    "speaker": batch.speakers[int(batch.speaker[i])]})
    return out


def sub_batch(batch, x0, x1):
    """Documents [x0, x1) as their own PackedBatch (used for sharding and bounded CPU samples)."""
    a, b = int(batch.doc_off[x0]), int(batch.doc_off[x1])
    t0, t1 = int(batch.seg_off[a]), int(batch.seg_off[b])
    return PackedBatch(batch.text[t0:t1], batch.seg_off[a:b + 1] - t0, batch.start[a:b], batch.end[a:b],
                       batch.speaker[a:b], batch.doc_off[x0:x1 + 1] - a, batch.speakers, None)
