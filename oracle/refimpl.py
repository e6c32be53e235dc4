"""The UNMODIFIED reference as the timed CPU arm (TEST / BENCH INFRASTRUCTURE).

`stage()` is the build recipe for `oracle/_ref/`: it copies the two files of the path,
`/root/reference/preprocessor.py` and `/root/reference/big_chunkeroosky.py` (main.py:139-149 calls into
exactly these), byte for byte into `oracle/_ref/`.  That directory is git-ignored (build output, never in
history) but travels to the GPU box with the snapshot like the built `.so` files, so that
`bench.py --impl reference` and the `cpu_baseline` leg can time the reference's own code on the box's host
cores.  `/root/reference` itself is only read here, in the build container, by `__graft_entry__.build()`.

`load()` imports the staged copies the way `oracle/refharness.py` imports the originals: the `nltk` shim
(restated untrained Punkt) on `sys.path`, `tiktoken.get_encoding` resolving to the real Rust CoreBPE fed
the repo's rank table.  Nothing in the product package imports this module.
"""
import contextlib, hashlib, io, os, shutil, sys

HERE = os.path.dirname(os.path.abspath(__file__))
REF_SRC = "/root/reference"
REF_DIR = os.path.join(HERE, "_ref")
FILES = ("preprocessor.py", "big_chunkeroosky.py")
_mods = None


def stage(force=False):
    """Copy the reference's two path files into oracle/_ref/ (no-op when /root/reference is absent)."""
    if not os.path.isdir(REF_SRC):
        return False
    os.makedirs(REF_DIR, exist_ok=True)
    for f in FILES:
        src, dst = os.path.join(REF_SRC, f), os.path.join(REF_DIR, f)
        if force or not os.path.exists(dst) or open(src, "rb").read() != open(dst, "rb").read():
            shutil.copyfile(src, dst)
    with open(os.path.join(REF_DIR, "SHA256SUMS"), "w") as out:
        for f in FILES:
            out.write("%s  %s\n" % (hashlib.sha256(open(os.path.join(REF_DIR, f), "rb").read()).hexdigest(), f))
    return True


def available():
    return all(os.path.exists(os.path.join(REF_DIR, f)) for f in FILES)


def load():
    """(preprocessor module, big_chunkeroosky module) from oracle/_ref, unmodified."""
    global _mods
    if _mods is not None:
        return _mods
    if not available():
        raise RuntimeError("oracle/_ref is not staged (run __graft_entry__.build() where /root/reference exists)")
    root = os.path.dirname(HERE)
    if root not in sys.path:
        sys.path.insert(0, root)
    sys.path.insert(0, os.path.join(HERE, "nltk_shim"))
    import tiktoken
    from . import tokenizer as tk
    enc = tk.encoding()
    tiktoken.get_encoding = lambda name: enc
    saved = {k: sys.modules.pop(k, None) for k in ("preprocessor", "big_chunkeroosky")}
    sys.path.insert(0, REF_DIR)
    try:
        with contextlib.redirect_stdout(io.StringIO()):
            import preprocessor as rp
            import big_chunkeroosky as rc
    finally:
        sys.path.remove(REF_DIR)
        for k, v in saved.items():
            sys.modules.pop(k, None)
            if v is not None:
                sys.modules[k] = v
    _mods = (rp, rc)
    return _mods


def run_docs(docs, max_tokens):
    """main.py:139-149 on each document: preprocess_transcript + chunk_transcript + postprocess_chunks with the
    reference's defaults, stdout discarded.  Returns (tokens, chunks)."""
    rp, rc = load()
    tok = nch = 0
    chunkers = {}
    with open(os.devnull, "w") as devnull, contextlib.redirect_stdout(devnull):
        for segs, mt in zip(docs, max_tokens):
            ch = chunkers.get(mt)
            if ch is None:
                ch = chunkers[mt] = rc.BigChunkeroosky(max_tokens_per_chunk=int(mt))
            proc = rp.preprocess_transcript(segs)
            chunks = ch.postprocess_chunks(ch.chunk_transcript(proc))
            tok += sum(c["token_count"] for c in chunks)
            nch += len(chunks)
    return tok, nch
