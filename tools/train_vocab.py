#!/usr/bin/env python
"""Train the deterministic `trbpe-v1` rank table (cl100k_base is not available offline).

Corpus = spans (cl100k pre-tokenisation pattern) of
  (a) the example transcript after the reference-equivalent formatting ("[ts] SPEAKER_xx: [ts] text ..."),
  (b) a synthetic lexicon: pseudo-words from a letter-trigram model of the example's words with Zipfian
      counts, capitalised / punctuated variants, numbers, timestamps, a few non-ASCII words,
so that the table has cl100k's size class (100 000 mergeable ranks + 5 specials) and an L2-resident,
not shared-memory-resident, pair table.  Output: llm_map_reduce_summarizer_b200/data/trbpe_v1.tiktoken.gz
(same line format as OpenAI's .tiktoken files: base64(token) SP rank) and data/lexicon.json (the
example's word-frequency table used by the synthetic transcript generator).

Deterministic: ties are broken by (count desc, left bytes, right bytes).
"""
import base64, collections, gzip, heapq, json, os, sys, time
import numpy as np
import regex

PAT = (r"(?i:'s|'t|'re|'ve|'m|'ll|'d)|[^\r\n\p{L}\p{N}]?\p{L}+|\p{N}{1,3}|"
       r" ?[^\s\p{L}\p{N}]+[\r\n]*|\s*[\r\n]+|\s+(?!\S)|\s+")
HERE = os.path.dirname(os.path.abspath(__file__))
DATA = os.path.join(HERE, "..", "llm_map_reduce_summarizer_b200", "data")


def build_corpus(example_path, n_pseudo=400000, seed=20261106):
    segs = json.load(open(example_path))["segments"]
    words = collections.Counter()
    for s in segs:
        for w in s["text"].split():
            words[w] += 1
    # lexicon for the synthetic generator: word -> count, order of first appearance kept out (sorted)
    lex = sorted(words.items(), key=lambda kv: (-kv[1], kv[0]))
    json.dump({"source": "transcript-example.json word frequencies (text.split())",
               "words": [w for w, _ in lex], "counts": [c for _, c in lex]},
              open(os.path.join(DATA, "lexicon.json"), "w"), separators=(",", ":"))
    rng = np.random.Generator(np.random.PCG64(seed))
    spans = collections.Counter()
    pat = regex.compile(PAT)

    def add(text, cnt):
        for m in pat.findall(text):
            spans[m] += cnt

    # (a) real words in context, heavily weighted
    for w, c in words.items():
        add(" " + w, c * 200)
        add(w, c * 20)
        add(" " + w.capitalize(), c * 10)
        add(" " + w.upper(), c)
    for mm in range(100):
        add("[%02d:%02d] SPEAKER_%02d:" % (mm, (mm * 7) % 60, mm % 16), 5000)
        add(" %d" % mm, 300)
        add("%02d" % mm, 50000)
    for n in range(1000):
        add("%d" % n, 50)
        add("%03d" % n, 20)
    add("--- TRANSCRIPT CHUNK INFORMATION ---\nTime Range: Speakers: Position: Chunk (approximately % through the transcript)\n--- TRANSCRIPT CHUNK CONTENT ---", 1000)
    for w in ["café", "naïve", "über", "señor", "Zürich", "résumé", "你好", "世界", "こんにちは", "привет", "مرحبا", "déjà", "façade", "jalapeño"]:
        add(" " + w, 400)
        add(w, 40)
    # (b) pseudo-words: letter trigram model over the real alphabetic words
    alpha = [w.lower().strip(".,?!'\"-") for w in words if w.isascii()]
    alpha = [w for w in alpha if w.isalpha() and len(w) >= 3]
    tri = collections.defaultdict(collections.Counter)
    for w in alpha:
        p = "^^" + w + "$"
        for i in range(len(p) - 2):
            tri[p[i:i + 2]][p[i + 2]] += words.get(w, 1)
    tri = {k: (list(v.keys()), np.array(list(v.values()), dtype=np.float64) / sum(v.values()))
           for k, v in sorted(tri.items())}
    made = set()
    attempts = 0
    while len(made) < n_pseudo and attempts < n_pseudo * 20:
        attempts += 1
        ctx, out = "^^", []
        while len(out) < 18:
            ks, ps = tri[ctx]
            ch = ks[int(rng.choice(len(ks), p=ps))]
            if ch == "$":
                break
            out.append(ch)
            ctx = ctx[1] + ch
        w = "".join(out)
        if len(w) >= 3 and w not in made:
            made.add(w)
    made = sorted(made)
    perm = rng.permutation(len(made))
    for rank, j in enumerate(perm):
        w = made[int(j)]
        c = max(1, int(60000 / (rank + 10) ** 0.9))
        add(" " + w, c * 4)
        add(w, c)
        if rank % 3 == 0:
            add(" " + w.capitalize(), c)
        if rank % 7 == 0:
            add(" " + w + "ing", c)
            add(" " + w + "ed", c)
        if rank % 11 == 0:
            add(" " + w + "'s", c)
    return spans


def train(spans, n_merges, log=sys.stderr):
    # words as lists of token ids; ids 0..255 = bytes
    vocab = [bytes([b]) for b in range(256)]
    items = sorted(spans.items())
    seqs = [list(w.encode("utf-8")) for w, _ in items]
    cnts = [c for _, c in items]
    pair_cnt = collections.defaultdict(int)
    pair_where = collections.defaultdict(set)
    for wi, s in enumerate(seqs):
        c = cnts[wi]
        for a, b in zip(s, s[1:]):
            pair_cnt[(a, b)] += c
            pair_where[(a, b)].add(wi)

    def key(p):
        return (-pair_cnt[p], vocab[p[0]], vocab[p[1]])

    heap = [(key(p), p) for p in pair_cnt]
    heapq.heapify(heap)
    t0 = time.time()
    merges = 0
    while merges < n_merges and heap:
        k, p = heapq.heappop(heap)
        c = pair_cnt.get(p, 0)
        if c <= 0:
            continue
        if k[0] != -c:
            heapq.heappush(heap, (key(p), p))
            continue
        newtok = vocab[p[0]] + vocab[p[1]]
        nid = len(vocab)
        vocab.append(newtok)
        merges += 1
        touched = set()
        for wi in list(pair_where[p]):
            s = seqs[wi]
            cw = cnts[wi]
            i = 0
            out = []
            changed = False
            n = len(s)
            while i < n:
                if i + 1 < n and s[i] == p[0] and s[i + 1] == p[1]:
                    out.append(nid)
                    i += 2
                    changed = True
                else:
                    out.append(s[i])
                    i += 1
            if not changed:
                continue
            for a, b in zip(s, s[1:]):
                pair_cnt[(a, b)] -= cw
                touched.add((a, b))
            for a, b in zip(out, out[1:]):
                pair_cnt[(a, b)] += cw
                pair_where[(a, b)].add(wi)
                touched.add((a, b))
            seqs[wi] = out
        pair_cnt.pop(p, None)
        pair_where.pop(p, None)
        touched.discard(p)
        for q in touched:
            if pair_cnt.get(q, 0) > 0:
                heapq.heappush(heap, (key(q), q))
        if merges % 10000 == 0:
            print(f"  {merges} merges, {time.time() - t0:.1f}s, last count {c}", file=log)
    return vocab


def main():
    example = sys.argv[1] if len(sys.argv) > 1 else "/root/reference/transcript-example.json"
    n_total = int(sys.argv[2]) if len(sys.argv) > 2 else 100000
    spans = build_corpus(example)
    print(len(spans), "distinct spans", file=sys.stderr)
    vocab = train(spans, n_total - 256)
    # tiktoken requires distinct byte strings; BPE training guarantees it only if no duplicate concat
    seen = {}
    out = []
    for tok in vocab:
        if tok in seen:
            continue
        seen[tok] = len(out)
        out.append(tok)
    path = os.path.join(DATA, "trbpe_v1.tiktoken.gz")
    with gzip.GzipFile(path, "wb", mtime=0) as f:
        for r, tok in enumerate(out):
            f.write(base64.b64encode(tok) + b" " + str(r).encode() + b"\n")
    print(len(out), "ranks written to", path, file=sys.stderr)


if __name__ == "__main__":
    main()
