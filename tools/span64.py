import sys, time, random
import os; sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from llm_map_reduce_summarizer_b200 import engine as E
eng = E.Engine(0)
rng = random.Random(9)
b64 = "ABCDEFGHIJKLMNOPQRSTUVWXYZabcdefghijklmnopqrstuvwxyz"
one = ["".join(rng.choice(b64) for _ in range(65536))]
eng.count_tokens(one)
ts = []
for _ in range(5):
    t0 = time.perf_counter(); c, _s = eng.count_tokens(one); ts.append(time.perf_counter() - t0)
print("64 KB span of letters through trc_count_tokens: %.2f ms (min of 5), %d tokens" % (min(ts) * 1e3, int(c[0])))
