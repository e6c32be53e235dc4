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
        self.rank_table = self._engine.vocab_name     # "cl100k_base", or "trbpe_v1" when the official file is absent

    def chunk_transcript(self, processed_segments: List[Dict[Any, Any]], add_context: bool = True) -> List[Dict[Any, Any]]:
        """big_chunkeroosky.py:46-145."""
        if not processed_segments:
            return []
        print(f"Chunkeroosky: Processing {len(processed_segments)} segments")                 # :67
        h = _pre.resident_handle(processed_segments)
        # the resident batch is only reused when the dicts are untouched and were counted with THIS chunker's table
        if h is None or h["engine"].vocab_name != self._engine.vocab_name:
            return self._chunk_foreign(processed_segments, add_context)
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

    def _chunk_foreign(self, processed_segments, add_context):
        """A list that did not come from this package's preprocess_transcript (hand-built, reloaded from
        JSON, edited ...): the dicts are packed as they are (texts verbatim, pieces from
        'segment_timestamps') and chunked on the device.  The device verifies that every combined
        segment's own start/end/text is what its pieces imply (preprocessor.py:186-213); anything else is
        outside the supported domain and raises instead of being approximated."""
        eng = self._engine
        batch = _eng.pack_processed(processed_segments)
        last = None
        for preserve in (True, False):           # combined text with "[ts] " markers, or plain joins (:190-200)
            eng.preprocess(batch, merge_same_speaker=True, max_segment_duration=None, preserve_timestamps=preserve,
                           want_merged_text=True)
            eng.chunk([self.max_tokens_per_chunk], self.context_tokens, add_context)
            R = eng.fetch_chunks()
            if int(R["doc_err"][0]) != -1:
                P = eng.fetch_processed()
                chunks = _eng.build_chunks(eng, batch, P, R, processed_segments, 0)
                for c in chunks:
                    c["total_chunks"] = None
                print(f"Chunkeroosky: Created {len(chunks)} chunks from {len(processed_segments)} segments")  # :143
                return chunks
            last = R
        raise NotImplementedError(
            "processed segments whose own start/end/text differ from what their segment_timestamps imply, "
            "or an over-budget segment whose text is not whitespace-normalised, are outside the device path")

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
