"""CPU oracle for the preprocess+chunk hot path.  TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import
this package.  The product (llm_map_reduce_summarizer_b200) never does: it fails loudly when the CUDA
library or a GPU is missing.

Parity status: preprocess (a1-a5) and the chunker control flow (a6-a10, a12-a14) are PINNED against
the real reference executed in the build container (tools/make_golden.py -> tests/golden/*.json).
The tiktoken boundary (a8) is pinned against the real Rust CoreBPE (tiktoken wheel, same rank table).
The Punkt boundary (a11, nltk==3.8.1, not installable offline) is a restatement from the published
algorithm: "parity unpinned" for that one function (see oracle/punkt.py).
"""
