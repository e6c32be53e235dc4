"""Drop-in for the reference's `big_chunkeroosky` module (/root/reference/big_chunkeroosky.py).

Same constructor, attributes, methods and chunk dict format; the work runs in the CUDA
kernels of libtrchunk.so (stage C, include/trchunk.h: trc_chunk).
"""
from typing import Any, Dict, List

from . import engine as _eng
from . import preprocessor as _pre


class BigChunkeroosky:
    def __init__(self, max_tokens_per_chunk: int = 4000, overlap_tokens: int = 200,
                 tokenizer_name: str = "cl100k_base", context_tokens: int = 150):
        """big_chunkeroosky.py:23-44."""
        self.max_tokens_per_chunk = max_tokens_per_chunk
        self.overlap_tokens = overlap_tokens          # stored and never used, as in the reference (:40)
        self.context_tokens = context_tokens
        self.effective_max_tokens = max_tokens_per_chunk - context_tokens
        self.tokenizer_name = tokenizer_name
        self._engine = _eng.default_engine(tokenizer_name)

    def chunk_transcript(self, processed_segments: List[Dict[Any, Any]], add_context: bool = True) -> List[Dict[Any, Any]]:
        """big_chunkeroosky.py:46-145."""
        if not processed_segments:
            return []
        print(f"Chunkeroosky: Processing {len(processed_segments)} segments")                 # :67
        h = getattr(processed_segments, "_trc", None)
        ok = (h is not None and len(processed_segments) == len(h["ids"])
              and all(id(d) == i for d, i in zip(processed_segments, h["ids"])))
        if not ok:
            raise NotImplementedError(
                "chunk_transcript accepts the list returned by this package's preprocess_transcript "
                "(the reference's only call pattern, main.py:139-149); arbitrary hand-built segment lists "
                "are not packed for the device yet")
        eng = h["engine"]
        if eng.batch_serial != h["serial"]:              # another batch replaced it: run stage P again
            h["serial"] = eng.preprocess(h["batch"], want_merged_text=True, **h["params"])
        eng.chunk([self.max_tokens_per_chunk], self.context_tokens, add_context)
        R = eng.fetch_chunks()
        chunks = _eng.build_chunks(eng, h["batch"], h["P"], R, processed_segments, 0)
        for c in chunks:
            c["total_chunks"] = None                    # filled by postprocess_chunks, as in the reference (:170)
        print(f"Chunkeroosky: Created {len(chunks)} chunks from {len(processed_segments)} segments")  # :143
        return chunks

    def postprocess_chunks(self, chunks: List[Dict[Any, Any]]) -> List[Dict[Any, Any]]:
        """big_chunkeroosky.py:544-567 (the clause speaker fix-up never fires: clause segments carry no
        'parent_segment_start', :482-489, :564)."""
        total = len(chunks)
        for c in chunks:
            c["total_chunks"] = total
        return chunks

    def _format_time(self, seconds: float) -> str:
        """big_chunkeroosky.py:234-242."""
        return _pre.format_timestamp(seconds)
