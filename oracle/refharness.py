"""Run the UNMODIFIED reference modules in the build container (TEST INFRASTRUCTURE; needs
/root/reference, so it is never used by -m gpu tests, smoke() or bench.py).

SURVEY.md Appendix C.1: nltk shim on sys.path + tiktoken.get_encoding patched to the repo rank table.
"""
import contextlib, io, os, sys

_REF = "/root/reference"
_mods = None


def available():
    return os.path.isdir(_REF)


def load():
    global _mods
    if _mods is not None:
        return _mods
    here = os.path.dirname(os.path.abspath(__file__))
    sys.path.insert(0, os.path.join(here, "nltk_shim"))
    import tiktoken
    from . import tokenizer as tk
    enc = tk.encoding()
    tiktoken.get_encoding = lambda name: enc
    saved = {k: sys.modules.pop(k, None) for k in ("preprocessor", "big_chunkeroosky")}
    sys.path.insert(0, _REF)
    try:
        import preprocessor as rp
        import big_chunkeroosky as rc
    finally:
        sys.path.remove(_REF)
        for k, v in saved.items():
            sys.modules.pop(k, None)
            if v is not None:
                sys.modules[k] = v
    _mods = (rp, rc)
    return _mods


def run(segments, max_tokens=4000, merge=True, max_dur=120, preserve=True, add_context=True,
        preprocess=True):
    rp, rc = load()
    with contextlib.redirect_stdout(io.StringIO()):
        proc = rp.preprocess_transcript(segments, merge, None, max_dur, preserve) if preprocess else segments
        ch = rc.BigChunkeroosky(max_tokens_per_chunk=max_tokens)
        chunks = ch.postprocess_chunks(ch.chunk_transcript(proc, add_context))
    return proc, chunks
