#!/usr/bin/env python
"""Per-source-line totals of one kernel from an ncu report: joins `ncu --page source --csv` (SASS rows) with the
line table of the cubin (`nvdisasm -g`).   python tools/ncu_lines.py <report.ncu-rep> <mangled kernel name> [top]"""
import csv, io, os, re, subprocess, sys, tempfile, collections
rep, kern = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
root = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")
so = os.path.join(root, "llm_map_reduce_summarizer_b200", "libtrchunk.so")
tmp = tempfile.mkdtemp()
subprocess.run(["cuobjdump", "-xelf", "all", so], cwd=tmp, check=True, capture_output=True)
cub = [f for f in os.listdir(tmp) if f.endswith(".cubin")][0]
sass = subprocess.run(["nvdisasm", "-g", os.path.join(tmp, cub)], capture_output=True, text=True).stdout.splitlines()
start = next(i for i, l in enumerate(sass) if l.startswith(".text." + kern + ":"))
line_of, cur = {}, None
for l in sass[start + 1:]:
    if l.startswith("//-----") or l.startswith("\t.section"):
        break
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m:
        cur = (os.path.basename(m.group(1)), int(m.group(2)))
        continue
    m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(\S.*);", l)
    if m and cur:
        line_of[int(m.group(1), 16)] = cur
m0 = re.match(r"_Z(\d+)", kern)                         # reports with several kernels: only the rows of this one
base_name = kern[len(m0.group(0)):len(m0.group(0)) + int(m0.group(1))] if m0 else kern
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "-k", "regex:^" + base_name + "($|<|\\()"], capture_output=True, text=True).stdout
if os.environ.get("KERNEL_ID"):                          # several launches of the kernel in the report: pick one by its ID
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-id", ":::" + os.environ["KERNEL_ID"]], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr = rows[1]
ix = {h: i for i, h in enumerate(hdr)}
base = None
agg = collections.defaultdict(lambda: [0, 0, 0])
tot = [0, 0, 0]
for r in rows[2:]:
    if len(r) < len(hdr) or not r[0].startswith("0x"):
        continue
    a = int(r[0], 16)
    if base is None:
        base = a
    key = line_of.get(a - base, ("?", 0))
    v = [int(float(r[ix["# Samples"]] or 0)), int(float(r[ix["Instructions Executed"]] or 0)), int(float(r[ix["Thread Instructions Executed"]] or 0))]
    for k in range(3):
        agg[key][k] += v[k]; tot[k] += v[k]
print("total samples %d, warp instructions %d, thread instructions %d (%.1f lanes)" % (tot[0], tot[1], tot[2], tot[2] / max(tot[1], 1)))
src_cache = {}
def src(f, n):
    p = os.path.join(root, "llm_map_reduce_summarizer_b200", "csrc", f)
    if p not in src_cache:
        src_cache[p] = open(p).read().splitlines() if os.path.exists(p) else []
    L = src_cache[p]
    return L[n - 1].strip()[:90] if 0 < n <= len(L) else ""
sk = 1 if os.environ.get("SORT") == "instr" else 0      # SORT=instr: by warp instructions instead of samples
for key, v in sorted(agg.items(), key=lambda kv: -kv[1][sk])[:top]:
    print("%5.1f%% samples %5.1f%% instr  %4.1f lanes  %s:%d  %s" % (100.0 * v[0] / max(tot[0], 1), 100.0 * v[1] / max(tot[1], 1), v[2] / max(v[1], 1), key[0], key[1], src(*key)))
