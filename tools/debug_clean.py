#!/usr/bin/env python
"""Debug aid: random segments through stage P on the GPU, cleaned text and token counts against the oracle."""
import os, random, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import numpy as np
from llm_map_reduce_summarizer_b200 import engine as E
from oracle import preprocess as OP, tokenizer as TK

rng = random.Random(int(sys.argv[1]) if len(sys.argv) > 1 else 1)
alpha = ["a", "b", "the", "the", "is", "Z", "é", "1", " ", " ", " ", "  ", "\t", " ", " ", "　", ".", ",", "!", "?", "-",
         "你好", "_", "x0", "​", "word", "end.Next", "café", " \t ", "́", "<|", "'s"]
eng = E.Engine(0)
enc = TK.encoding()
bad = 0
for trial in range(int(sys.argv[2]) if len(sys.argv) > 2 else 30):
    nseg = rng.choice([1, 3, 40, 200, 700])
    segs = []
    for i in range(nseg):
        L = rng.choice([0, 1, 5, 30, 60, 200, 900]) if rng.random() < 0.3 else rng.randint(0, 80)
        segs.append({"start": float(i), "end": float(i + 1), "text": "".join(rng.choice(alpha) for _ in range(L)), "speaker": "A"})
    b = E.pack_documents([segs])
    eng.preprocess(b, merge_same_speaker=False, want_merged_text=False)
    P = eng.fetch_processed()
    clean, coff, k2r, ctok = P["clean"].tobytes(), P["coff"], P["k2r"], P["ctok"]
    want = [(i, OP.clean_text(s["text"])) for i, s in enumerate(segs)]
    want = [(i, t) for i, t in want if t]
    if len(want) != len(k2r) or [i for i, _ in want] != k2r.tolist():
        print("trial", trial, "kept mismatch", len(want), len(k2r)); bad += 1; continue
    for j, (i, t) in enumerate(want):
        got = clean[int(coff[j]):int(coff[j + 1])]
        if got != t.encode():
            print("trial", trial, "seg", i, "nseg", nseg, "\n raw ", segs[i]["text"].encode()[:200], "\n want", t.encode()[:200], "\n got ", got[:200]); bad += 1; break
        tk = len(enc.encode_ordinary(" " + t))
        if int(ctok[j]) != tk:
            print("trial", trial, "seg", i, "tokens", int(ctok[j]), "want", tk, t.encode()[:120]); bad += 1; break
print("bad", bad)
