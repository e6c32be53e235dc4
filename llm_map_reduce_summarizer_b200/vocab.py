"""Rank-table resolution (replaces tiktoken.get_encoding at /root/reference/big_chunkeroosky.py:43).

`cl100k_base` itself cannot be downloaded here (no network).  If the official file is present in
$TIKTOKEN_CACHE_DIR (name 9b5ad71b2ce5302211f9c61530b329a4922fc6a4, sha256 223921b7...b2a7) it is used;
otherwise the repo's deterministic table `trbpe_v1` (tools/train_vocab.py, 100 000 ranks, cl100k
pattern and special tokens).  The kernels are table driven, so either works unchanged.
"""
import base64, gzip, hashlib, json, os
import numpy as np

_DATA = os.path.join(os.path.dirname(os.path.abspath(__file__)), "data")
CL100K_CACHE_KEY = "9b5ad71b2ce5302211f9c61530b329a4922fc6a4"
CL100K_SHA256 = "223921b76ee99bde995b7ff738513eef100fb51d18c93597a113bcffe865b2a7"
SPECIALS = ["<|endoftext|>", "<|fim_prefix|>", "<|fim_middle|>", "<|fim_suffix|>", "<|endofprompt|>"]


def _parse(raw):
    toks, ranks = [], []
    for line in raw.splitlines():
        if line.strip():
            t, r = line.split()
            toks.append(base64.b64decode(t))
            ranks.append(int(r))
    return toks, ranks


def _cl100k_file():
    cache = os.environ.get("TIKTOKEN_CACHE_DIR") or os.environ.get("DATA_GYM_CACHE_DIR")
    if cache:
        p = os.path.join(cache, CL100K_CACHE_KEY)
        if os.path.exists(p):
            raw = open(p, "rb").read()
            if hashlib.sha256(raw).hexdigest() == CL100K_SHA256:
                return raw
    return None


def resolve_table_name(name="cl100k_base"):
    """Name of the table `load_rank_table(name)` would return, without parsing it."""
    if name != "cl100k_base":
        raise ValueError(
            f"Unknown encoding {name!r}: this build carries the cl100k pattern and special tokens only "
            "(the reference never passes anything else, big_chunkeroosky.py:27,43); pass rank_table= to use another table")
    return "cl100k_base" if _cl100k_file() is not None else "trbpe_v1"


_warned = False


def load_rank_table(name="cl100k_base"):
    """Returns (table_name, tokens: list[bytes], ranks: list[int]).

    Any name but "cl100k_base" raises ValueError, as tiktoken.get_encoding does for a name it does not know
    (the pre-tokenisation pattern and the special tokens here are cl100k's).  Without the official file the
    repo's own table `trbpe_v1` is used and a warning says so ONCE per process: token counts, and with them
    every chunk boundary, then differ from a reference run on the real cl100k_base."""
    global _warned
    if resolve_table_name(name) == "cl100k_base":
        t, r = _parse(_cl100k_file())
        return "cl100k_base", t, r
    if not _warned and os.environ.get("TRCHUNK_ALLOW_FALLBACK_VOCAB") != "1":
        import warnings
        warnings.warn("cl100k_base is not available offline ($TIKTOKEN_CACHE_DIR/%s missing or sha256 mismatch): using the "
                      "repo's rank table trbpe_v1; token counts and chunk boundaries are those of trbpe_v1, not of a "
                      "reference run on cl100k_base (set TRCHUNK_ALLOW_FALLBACK_VOCAB=1 to silence)" % CL100K_CACHE_KEY,
                      RuntimeWarning, stacklevel=2)
    _warned = True
    with gzip.open(os.path.join(_DATA, "trbpe_v1.tiktoken.gz"), "rb") as f:
        t, r = _parse(f.read())
    return "trbpe_v1", t, r


def load_props():
    d = json.load(open(os.path.join(_DATA, "unicode_props.json")))
    props = np.zeros(0x110000, dtype=np.uint8)
    for a, b, v in d["ranges"]:
        props[a:b + 1] = v
    return props


def pack_strings(strs):
    """list[bytes] -> (uint8 buffer, int64 offsets)."""
    off = np.zeros(len(strs) + 1, dtype=np.int64)
    if strs:
        off[1:] = np.cumsum(np.fromiter(map(len, strs), dtype=np.int64, count=len(strs)))
    buf = np.frombuffer(b"".join(strs), dtype=np.uint8).copy() if strs else np.zeros(0, dtype=np.uint8)
    return buf, off
