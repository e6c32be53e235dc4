"""Token-count oracle (TEST INFRASTRUCTURE).

Two independent checkers for the tiktoken boundary (/root/reference/big_chunkeroosky.py:43,83,298,370,510):

 * `encoding()`    - the real Rust CoreBPE from the installed tiktoken wheel, fed the repo's rank table
                     and the pattern string of the version the reference pins (tiktoken==0.5.1,
                     /root/reference/requirements.txt:2).  This is the pin.
 * `split_spans()` / `bpe_count()` - a pure-Python restatement of the published algorithm
                     (regex alternation semantics SURVEY.md A.5, byte-pair merge
                     tiktoken/_educational.py:83-116), used to cross-check small cases and to document
                     exactly what the CUDA kernels implement.
"""
import base64, gzip, json, os

PAT_051 = (r"(?i:'s|'t|'re|'ve|'m|'ll|'d)|[^\r\n\p{L}\p{N}]?\p{L}+|\p{N}{1,3}|"
           r" ?[^\s\p{L}\p{N}]+[\r\n]*|\s*[\r\n]+|\s+(?!\S)|\s+")
SPECIALS = ["<|endoftext|>", "<|fim_prefix|>", "<|fim_middle|>", "<|fim_suffix|>", "<|endofprompt|>"]
_DATA = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "llm_map_reduce_summarizer_b200", "data")


def load_ranks(path=None):
    path = path or os.path.join(_DATA, "trbpe_v1.tiktoken.gz")
    op = gzip.open if path.endswith(".gz") else open
    ranks = {}
    with op(path, "rb") as f:
        for line in f:
            if line.strip():
                tok, r = line.split()
                ranks[base64.b64decode(tok)] = int(r)
    return ranks


_ENC = {}


def encoding(ranks=None, name="trbpe_v1"):
    import tiktoken
    key = name if ranks is None else id(ranks)
    if key not in _ENC:
        rk = ranks if ranks is not None else load_ranks()
        n = max(rk.values()) + 1
        _ENC[key] = tiktoken.Encoding(name, pat_str=PAT_051, mergeable_ranks=rk,
                                      special_tokens={s: n + i for i, s in enumerate(SPECIALS)})
    return _ENC[key]


def load_props():
    d = json.load(open(os.path.join(_DATA, "unicode_props.json")))
    props = bytearray(0x110000)
    for a, b, v in d["ranges"]:
        props[a:b + 1] = bytes([v]) * (b - a + 1)
    return props


_PROPS = None


def _cls(ch):
    """0=other 1=L 2=N 3=whitespace (Rust regex tables)."""
    global _PROPS
    if _PROPS is None:
        _PROPS = load_props()
    return (_PROPS[ord(ch)] >> 2) & 3


def split_spans(s):
    """Leftmost-first alternation of the 0.5.1 cl100k pattern, restated as a hand matcher."""
    out = []
    n = len(s)
    p = 0
    while p < n:
        c = s[p]
        k = _cls(c)
        e = None
        # 1: contraction, only at a match start
        if c == "'" and p + 1 < n:
            a = s[p + 1]
            b = s[p + 2] if p + 2 < n else ""
            if a in "sStTmMdDſ":
                e = p + 2
            elif (a in "rR" and b in ("e", "E")) or (a in "vV" and b in ("e", "E")) or (a in "lL" and b in ("l", "L")):
                e = p + 3
        # 2: optional one char not in \r\n L N, then L+
        if e is None:
            q = p
            if k != 1 and k != 2 and c not in "\r\n" and p + 1 < n and _cls(s[p + 1]) == 1:
                q = p + 1
            if q < n and _cls(s[q]) == 1:
                while q < n and _cls(s[q]) == 1:
                    q += 1
                e = q
        # 3: 1-3 N
        if e is None and k == 2:
            q = p
            while q < n and q - p < 3 and _cls(s[q]) == 2:
                q += 1
            e = q
        # 4: optional space, 1+ other, any \r\n
        if e is None:
            q = p
            if c == " " and p + 1 < n and _cls(s[p + 1]) == 0:
                q = p + 1
            if _cls(s[q]) == 0:
                while q < n and _cls(s[q]) == 0:
                    q += 1
                while q < n and s[q] in "\r\n":
                    q += 1
                e = q
        if e is None:
            # whitespace run [p, r)
            r = p
            while r < n and _cls(s[r]) == 3:
                r += 1
            last_nl = -1
            for i in range(p, r):
                if s[i] in "\r\n":
                    last_nl = i
            if last_nl >= 0:                 # 5: \s*[\r\n]+
                e = last_nl + 1
            elif r == n:                     # 6: \s+(?!\S) at end of string
                e = r
            elif r - p >= 2:                 # 6: leave the last space for the next match
                e = r - 1
            else:                            # 7
                e = r
        out.append(s[p:e])
        p = e
    return out


def bpe_count(ranks, piece: bytes) -> int:
    if piece in ranks:
        return 1
    parts = [bytes([b]) for b in piece]
    while True:
        best, bi = None, -1
        for i in range(len(parts) - 1):
            r = ranks.get(parts[i] + parts[i + 1])
            if r is not None and (best is None or r < best):
                best, bi = r, i
        if best is None:
            return len(parts)
        parts[bi:bi + 2] = [parts[bi] + parts[bi + 1]]


def count_python(ranks, text: str) -> int:
    return sum(bpe_count(ranks, sp.encode("utf-8")) for sp in split_spans(text))
