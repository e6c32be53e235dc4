"""trchunk: B200-native (sm_100a) preprocess + chunk path of llm-map-reduce-summarizer.

Drop-in modules (same entry points and dict formats as the reference, /root/reference/main.py:27-28):
    from llm_map_reduce_summarizer_b200 import preprocessor, big_chunkeroosky
Batch / multi-document API used by bench.py:
    from llm_map_reduce_summarizer_b200.engine import Engine, pack_documents
The compute path is libtrchunk.so (hand-written CUDA kernels behind the C ABI in include/trchunk.h).
There is no CPU fallback: importing works anywhere, but creating an Engine without a B200 raises.
"""
__all__ = ["preprocessor", "big_chunkeroosky", "engine", "vocab", "synth"]
