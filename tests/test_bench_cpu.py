"""bench.py on CPU: the reference arm's JSON contract (the driver parses this line) and the
bookkeeping that decides which kernel the roofline record is about."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.path.join(ROOT, "oracle", "_ref", "libosqp_ref.so")


@pytest.mark.skipif(not os.path.exists(REF), reason="oracle/_ref not built")
def test_reference_arm_prints_the_contract_line():
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--cpu-n", "300",
                        "--steps", "1", "--warmup", "0"], cwd=ROOT, stdout=subprocess.PIPE, stderr=subprocess.PIPE,
                       text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    line = json.loads(r.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference" and line["metric"] == "admm_iters_per_s"
    assert line["higher_is_better"] is True and line["unit"] == "ADMM iterations/s"
    assert line["n_gpus"] == 1 and line["steps"] == 1
    assert line["value"] > 0 and line["e2e"]["value"] > 0
    assert line["e2e"]["h2d_bytes_per_step"] == 0 and line["e2e"]["d2h_bytes_per_step"] == 0
    cb = line["cpu_baseline"]
    assert cb["kind"] == "reference" and cb["cores"] == 1 and cb["value"] == line["value"]
    # value = iterations / solve time, e2e = iterations / (setup + solve): the first cannot be the smaller
    assert line["value"] >= line["e2e"]["value"]


def test_reference_arm_under_torchrun_only_rank0_prints(tmp_path):
    """N > 1: rank 0 alone runs and prints the reference line, the other ranks exit 0 without work."""
    if not os.path.exists(REF):
        pytest.skip("oracle/_ref not built")
    env = dict(os.environ, RANK="1", WORLD_SIZE="2", LOCAL_RANK="1")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2",
                        "--cpu-n", "300", "--steps", "1", "--warmup", "0"], cwd=ROOT, env=env,
                       stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    assert r.stdout.strip() == ""


def test_traffic_table_is_read_from_the_committed_capture():
    sys.path.insert(0, ROOT)
    import bench
    t = bench.load_traffic()
    assert "spmv_M([P|A'][p;w])" in t
    rec = t["spmv_M([P|A'][p;w])"]
    assert rec["dram_bytes_per_launch"] > 1e8 and ".ncu-rep" in rec["source"]
