"""Batched token counting on the device: `len(tokenizer.encode(text))` for many strings at once.

The reference counts tokens in two more places outside the chunker (result_aggregator.py:357-392,
`_total_tokens` / `_calculate_batch_size`); both only need the LENGTH of the encoding, which is what the
K2 kernels produce.  Same semantics as tiktoken's default `encode`: text that
contains a special-token literal raises ValueError.
"""
from typing import Iterable, List

from . import engine as _eng


class TokenCounter:
    def __init__(self, tokenizer_name: str = "cl100k_base"):
        self._engine = _eng.default_engine(tokenizer_name)

    def count_batch(self, texts: Iterable[str]) -> List[int]:
        texts = list(texts)
        cnt, sp = self._engine.count_tokens(texts, leading_space=False)
        for s in sp:
            if s >= 0:
                raise _eng.special_token_error(self._engine, int(s))
        return [int(c) for c in cnt]

    def count(self, text: str) -> int:
        return self.count_batch([text])[0]

    def total(self, texts: Iterable[str]) -> int:
        """result_aggregator.py:382-392 `_total_tokens`."""
        return sum(self.count_batch(texts))
