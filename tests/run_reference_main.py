"""Run the UNMODIFIED reference main.py (mock LLM) with either its own modules or this package's drop-ins.
TEST INFRASTRUCTURE (needs /root/reference).  usage: run_reference_main.py {reference|dropin} OUT_DIR"""
import os, sys, types

ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
sys.path.insert(0, ROOT)
mode, out = sys.argv[1], sys.argv[2]
import tiktoken
from oracle import tokenizer as _tk
tiktoken.get_encoding = lambda name: _tk.encoding()      # no network: repo rank table (also result_aggregator.py:50)
if mode == "dropin":
    from llm_map_reduce_summarizer_b200 import build, engine as E, preprocessor, big_chunkeroosky
    E.set_default_engine(E.Engine(0, build.build_hostsim()))      # host test build: this container has no GPU
    sys.modules["preprocessor"] = preprocessor
    sys.modules["big_chunkeroosky"] = big_chunkeroosky
else:
    sys.path.insert(0, os.path.join(ROOT, "oracle", "nltk_shim"))
try:
    import dotenv  # noqa: F401
except ImportError:
    m = types.ModuleType("dotenv"); m.load_dotenv = lambda *a, **k: False; sys.modules["dotenv"] = m
sys.path.insert(0, "/root/reference")
os.chdir("/root/reference")
for k in ("OPENAI_API_KEY", "ANTHROPIC_API_KEY"):
    os.environ.pop(k, None)
sys.argv = ["main.py", "--input", "transcript-example.json", "--output", os.path.join(out, "summary.md"),
            "--save-chunks", os.path.join(out, "chunks.json"), "--quiet", "--max-tokens-per-chunk", "2000"]
import main
main.main()
