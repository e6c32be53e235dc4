"""Minimal `nltk` stand-in so the UNMODIFIED reference big_chunkeroosky.py imports offline.

TEST INFRASTRUCTURE: put oracle/nltk_shim on sys.path before importing the reference
(/root/reference/big_chunkeroosky.py:9,15-18,44).  Only the three names the reference touches exist.
"""
import types as _types
from oracle import punkt as _punkt

data = _types.SimpleNamespace(find=lambda *a, **k: True)
tokenize = _types.SimpleNamespace(PunktSentenceTokenizer=_punkt.PunktSentenceTokenizer)


def download(*a, **k):
    return True
