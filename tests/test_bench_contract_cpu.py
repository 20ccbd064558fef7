"""bench.py on a machine without a GPU: the reference arm prints the contract line, the GPU arm refuses to run.

The reference arm times oracle/grid4d_torch.py (the op-by-op port of models/hash_encoding.py) on the host cores;
it is the one place outside tests/ and smoke() that may execute oracle/ code.
"""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(*args, timeout=400):
    env = dict(os.environ, CUDA_VISIBLE_DEVICES="")          # same behaviour on a GPU box
    return subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), *args], cwd=ROOT, env=env,
                          capture_output=True, text=True, timeout=timeout)


def test_reference_arm_prints_the_contract_line():
    r = _run("--impl", "reference", "--steps", "1", "--warmup", "1")
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [ln for ln in r.stdout.splitlines() if ln.startswith("{")]
    assert len(lines) == 1                                     # ONE JSON line
    d = json.loads(lines[0])
    assert d["impl"] == "reference"
    assert d["metric"] == "grid4d_hash_encode_fwd_bwd_mpoints_per_s" and d["unit"] == "Mpoints/s"
    assert d["higher_is_better"] is True and d["n_gpus"] == 1 and d["steps"] == 1 and d["warmup"] == 1
    assert d["value"] > 0 and d["ms_per_step"] > 0
    assert d["config"]["workload"] == "c2"
    cb = d["cpu_baseline"]
    assert cb["kind"] in ("port", "reference") and cb["cores"] >= 1 and cb["sample"] and cb["value"] == d["value"]
    e2e = d["e2e"]
    assert e2e["value"] == d["value"] and e2e["unit"] == d["unit"]
    assert e2e["h2d_bytes_per_step"] == 0 and e2e["d2h_bytes_per_step"] == 0
    assert d["gpu_launches"] == 0
    # same config keys as the GPU arm's line (the driver compares them), plus what the bounded sample was
    assert {"workload", "description", "points_per_gpu", "coords", "levels", "table_size", "interpolation", "l2_policy"} <= set(d["config"])
    import bench
    assert d["config"] == bench.describe_config("c2", dict(bench.WORKLOADS["c2"]), 2 ** 22, 1)      # word for word the GPU arm's


def test_reference_arm_at_the_reference_geometry_runs_the_reference_or_its_port():
    """--workload c1r: the reference's own Grid4DEncoder geometry.  With oracle/_ref built (this container: build()
    copies the unmodified modules) the arm times the reference classes themselves, kind "reference"; else the port."""
    r = _run("--impl", "reference", "--workload", "c1r", "--steps", "1", "--warmup", "1")
    assert r.returncode == 0, r.stderr[-2000:]
    d = json.loads([ln for ln in r.stdout.splitlines() if ln.startswith("{")][0])
    have_ref = os.path.exists(os.path.join(ROOT, "oracle", "_ref", "models", "hash_encoding.py"))
    assert d["cpu_baseline"]["kind"] == ("reference" if have_ref else "port")
    assert d["config"]["workload"] == "c1r" and d["value"] > 0


def test_reference_arm_under_torchrun_only_rank0_prints():
    env = dict(os.environ, CUDA_VISIBLE_DEVICES="", RANK="1", LOCAL_RANK="1", WORLD_SIZE="2",
               MASTER_ADDR="127.0.0.1", MASTER_PORT="29533")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2",
                        "--steps", "1", "--warmup", "1"], cwd=ROOT, env=env, capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stderr[-2000:]
    assert not [ln for ln in r.stdout.splitlines() if ln.startswith("{")]     # the other ranks exit 0 without work


def test_gpu_arm_has_no_cpu_fallback():
    r = _run("--steps", "1", "--warmup", "1", "--no-cpu-baseline", timeout=200)
    assert r.returncode != 0
    assert not [ln for ln in r.stdout.splitlines() if ln.startswith("{")]
    assert "no CPU fallback" in r.stderr
