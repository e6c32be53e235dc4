"""bench.py's reference arm (the one leg that runs without a GPU) prints one JSON line with the contract's keys."""
import json, os, subprocess, sys

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
ROOT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")


def test_reference_arm_line():
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0",
                          "--scale", "0.01", "--cpu-sample-mb", "0.3"], check=True, capture_output=True, text=True, timeout=300).stdout
    line = json.loads(out.strip().splitlines()[-1])
    assert line["impl"] == "reference" and line["metric"] == "transcript tokens/sec preprocess+chunk" and line["unit"] == "tokens/s"
    for k in ("value", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling", "vs_baseline", "dtype", "data", "config",
              "cpu_baseline", "e2e"):
        assert k in line, k
    from oracle import refimpl
    kind = "reference" if refimpl.available() else "port"
    assert line["value"] > 0 and line["cpu_baseline"]["kind"] == kind and line["cpu_baseline"]["cores"] >= 1
    if kind == "reference":
        assert line["cpu_baseline"]["oracle_port"]["value"] > 0 and line["cpu_baseline"]["single_process"]["cores"] == 1
    for k in ("workload", "scale", "docs_per_gpu", "segments_per_gpu", "raw_text_bytes_per_gpu", "rank_table", "l2", "merged_text"):
        assert k in line["config"], k
    assert line["e2e"]["value"] == line["value"] and line["e2e"]["h2d_bytes_per_step"] == 0 and line["e2e"]["d2h_bytes_per_step"] == 0
    assert line["config"]["workload"].startswith("config 3")


import pytest


@pytest.mark.gpu
def test_device_arm_line():
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--steps", "3", "--warmup", "3", "--scale", "0.05",
                          "--cpu-sample-mb", "0.3", "--e2e-repeats", "1", "--no-extra-configs"], check=True, capture_output=True, text=True, timeout=600).stdout
    line = json.loads(out.strip().splitlines()[-1])
    for k in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling", "vs_baseline", "dtype",
              "data", "config", "clocks", "e2e", "gpu_launches", "roofline", "cpu_baseline"):
        assert k in line, k
    assert line["n_gpus"] == 1 and line["steps"] == 3 and line["dtype"] == "u8" and line["scaling"] == "weak" and line["vs_baseline"] is None
    assert line["value"] > 0 and line["gpu_launches"] > 0
    r = line["roofline"]
    assert r["bound"] == "hbm" and r["unit"] == "GB/s" and abs(r["frac"] - r["achieved"] / r["peak"]) < 1e-9 and r["kernel"].startswith("k_")
    e = line["e2e"]
    assert e["value"] > 0 and e["h2d_bytes_per_step"] > 0 and e["d2h_bytes_per_step"] > 0 and e["value"] < line["value"]
    c = line["cpu_baseline"]
    assert c["kind"] in ("reference", "port") and c["cores"] == 1 and c["value"] > 0
    # the reference arm prints the very same config object (the driver compares them)
    ref = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0",
                          "--scale", "0.05", "--cpu-sample-mb", "0.3", "--no-cpu-baseline"], check=True, capture_output=True, text=True,
                         timeout=300).stdout
    assert json.loads(ref.strip().splitlines()[-1])["config"] == line["config"]


def test_affinity_helper_never_fails_without_a_gpu():
    """bench.pin_near_gpu reports what it did as a string and leaves the process where it is when NVML or the GPU's sysfs
    entry is missing (this container)."""
    import bench
    before = os.sched_getaffinity(0)
    note = bench.pin_near_gpu(0)
    assert isinstance(note, str) and note
    assert os.sched_getaffinity(0) <= before
