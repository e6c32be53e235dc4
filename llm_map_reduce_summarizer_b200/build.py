"""Build libtrchunk.so (nvcc, sm_100a) in-tree.  `python -m llm_map_reduce_summarizer_b200.build`"""
import os, subprocess, sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libtrchunk.so")
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "--fmad=false", "-std=c++17",
              "--extended-lambda", "--expt-relaxed-constexpr", "-Xcompiler", "-fPIC", "-shared"]


def sources():
    return sorted(os.path.join(CSRC, f) for f in os.listdir(CSRC)) + [os.path.join(HERE, "..", "include", "trchunk.h")]


def stale(target):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(s) > t for s in sources())


def build(force=False, verbose=False):
    if not force and not stale(LIB):
        return LIB
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    cmd = [nvcc] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + [os.path.join(CSRC, "trchunk.cu"), "-o", LIB]
    subprocess.run(cmd, check=True)
    return LIB


def build_hostsim(force=False):
    """TEST INFRASTRUCTURE: the same sources compiled by g++ for the CPU test-suite (tests/hostsim/)."""
    out_dir = os.path.join(HERE, "..", "tests", "hostsim")
    os.makedirs(out_dir, exist_ok=True)
    out = os.path.join(out_dir, "libtrchunk_hostsim.so")
    if not force and not stale(out):
        return out
    cmd = ["g++", "-O2", "-g", "-std=c++17", "-ffp-contract=off", "-fPIC", "-shared", "-DTRC_HOSTSIM", "-x", "c++",
           os.path.join(CSRC, "trchunk.cu"), "-o", out]
    subprocess.run(cmd, check=True)
    return out


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
    if "--hostsim" in sys.argv:
        print(build_hostsim(force=True))
