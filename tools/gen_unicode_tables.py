#!/usr/bin/env python
"""Generate the per-code-point property table used by both the CUDA kernels and the oracle.

One byte per code point (0 .. 0x10FFFF), stored range-encoded in
llm_map_reduce_summarizer_b200/data/unicode_props.json:

  bit0  PY_SPACE   str.isspace()           (preprocessor.py:38,80 split()/strip())
  bit1  PY_WORD    re `\\w`                  (preprocessor.py:84)
  bit2-3 RS class  0=other 1=\\p{L} 2=\\p{N} 3=\\s   as seen by the Rust regex engine inside
                   tiktoken's CoreBPE (derived by probing CoreBPE itself)
  bit4  PY_UPPER   str.isupper()           (Punkt first_upper)
  bit5  PY_LOWER   str.islower()           (Punkt first_lower)
  bit6  PY_DIGIT   re `\\d`                  (Punkt ##number## / is_initial)

Run here (needs tiktoken); the output is committed so every box uses identical tables.
"""
import json, os, re, sys
import tiktoken

PAT_051 = (r"(?i:'s|'t|'re|'ve|'m|'ll|'d)|[^\r\n\p{L}\p{N}]?\p{L}+|\p{N}{1,3}|"
           r" ?[^\s\p{L}\p{N}]+[\r\n]*|\s*[\r\n]+|\s+(?!\S)|\s+")


def rust_classes():
    ranks = {bytes([b]): b for b in range(256)}
    r = 256
    for lead in (b"a", b"1", b"!"):
        for x in range(256):
            key = lead + bytes([x])
            if key not in ranks:
                ranks[key] = r
                r += 1
    enc = tiktoken.Encoding("probe", pat_str=PAT_051, mergeable_ranks=ranks, special_tokens={})
    cps = [c for c in range(0x110000) if not (0xD800 <= c <= 0xDFFF)]
    cls = bytearray(0x110000)
    B = 1 << 16
    for i in range(0, len(cps), B):
        blk = cps[i:i + B]
        chars = [chr(c) for c in blk]
        ks = [len(ch.encode("utf-8")) for ch in chars]
        la = enc.encode_ordinary_batch(["a" + ch for ch in chars])
        l1 = enc.encode_ordinary_batch(["1" + ch for ch in chars])
        lb = enc.encode_ordinary_batch(["!" + ch for ch in chars])
        for c, k, a, n, b in zip(blk, ks, la, l1, lb):
            if len(a) == k:
                cls[c] = 1
            elif len(n) == k:
                cls[c] = 2
            elif len(b) == k:
                cls[c] = 0
            else:
                cls[c] = 3
    cls[0x0A] = 3
    cls[0x0D] = 3
    return cls


def main():
    cls = rust_classes()
    w = re.compile(r"\w")
    d = re.compile(r"\d")
    props = bytearray(0x110000)
    for c in range(0x110000):
        if 0xD800 <= c <= 0xDFFF:
            continue
        ch = chr(c)
        v = 0
        if ch.isspace():
            v |= 1
        if w.match(ch):
            v |= 2
        v |= cls[c] << 2
        if ch.isupper():
            v |= 16
        if ch.islower():
            v |= 32
        if d.match(ch):
            v |= 64
        props[c] = v
    ranges = []
    s = 0
    for c in range(1, 0x110001):
        if c == 0x110000 or props[c] != props[s]:
            if props[s]:
                ranges.append([s, c - 1, props[s]])
            s = c
    out = os.path.join(os.path.dirname(__file__), "..", "llm_map_reduce_summarizer_b200", "data",
                       "unicode_props.json")
    import unicodedata
    with open(out, "w") as f:
        json.dump({"unidata_version": unicodedata.unidata_version,
                   "tiktoken": tiktoken.__version__, "ranges": ranges}, f, separators=(",", ":"))
    n = {k: sum(1 for c in range(0x110000) if props[c] & m) for k, m in
         [("py_space", 1), ("py_word", 2), ("upper", 16), ("lower", 32), ("digit", 64)]}
    n["rs_L"] = sum(1 for c in range(0x110000) if (props[c] >> 2) & 3 == 1)
    n["rs_N"] = sum(1 for c in range(0x110000) if (props[c] >> 2) & 3 == 2)
    n["rs_S"] = sum(1 for c in range(0x110000) if (props[c] >> 2) & 3 == 3)
    print(len(ranges), "ranges", n)


if __name__ == "__main__":
    main()
