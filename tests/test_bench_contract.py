"""bench.py's contract, checked where it can run without a GPU.

The reference arm (`--impl reference`) is the CPU oracle port, so it runs here; the product arm needs a B200 and must
refuse to run without one (no CPU fallback).  The committed full line in profiles/ is checked for the contract keys.
"""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

BASE_KEYS = {"metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling",
             "vs_baseline", "dtype", "data", "config", "e2e", "gpu_launches"}


def run_bench(*flags, env=None):
    e = dict(os.environ)
    e.update(env or {})
    return subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), *flags], capture_output=True, text=True,
                          timeout=600, env=e, cwd=ROOT)


def test_reference_arm_prints_one_contract_line():
    # torchrun exports OMP_NUM_THREADS=1 to its workers: the reference arm must still use the whole box
    out = run_bench("--impl", "reference", "--steps", "1", "--warmup", "0", env={"OMP_NUM_THREADS": "1"})
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [l for l in out.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and BASE_KEYS <= set(d)
    base = json.load(open(os.path.join(ROOT, "BASELINE.json")))
    assert base["metric"].startswith(d["metric"]) and d["unit"] == "requests/s"
    assert d["value"] > 0 and d["gpu_launches"] == 0
    assert d["cpu_baseline"]["kind"] in ("port", "reference") and d["cpu_baseline"]["value"] == d["value"]
    assert d["cpu_baseline"]["cores"] == len(os.sched_getaffinity(0))
    assert d["e2e"] == {"value": d["value"], "unit": d["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert "workload" in d["config"] and "model" not in d["config"]


def test_product_arm_refuses_to_run_without_a_gpu():
    import torch
    if torch.cuda.is_available():
        import pytest
        pytest.skip("a GPU is present")
    out = run_bench("--steps", "1", "--warmup", "1", "--no-cpu-baseline")
    assert out.returncode != 0
    assert not [l for l in out.stdout.splitlines() if l.startswith("{")], "no number may be printed without a GPU"


def test_committed_full_line_has_the_contract_keys():
    d = json.loads(open(os.path.join(ROOT, "profiles", "r2_bench_n1.json")).read().strip().splitlines()[-1])
    assert BASE_KEYS | {"clocks", "roofline", "cpu_baseline"} <= set(d)
    r = d["roofline"]
    assert r["bound"] == "hbm" and abs(r["frac"] - r["achieved"] / r["peak"]) < 1e-9 and r["unit"] == "GB/s"
    assert abs(r["bytes_per_launch"] - r["bytes_per_step"] * r["steps_per_launch"]) < 1
    assert d["e2e"]["h2d_bytes_per_step"] > 0 and d["e2e"]["d2h_bytes_per_step"] > 0
    assert 0 < d["e2e"]["frac_of_h2d_peak"] <= 1.05 and set(d["e2e"]["other_transports"]) == {"f64", "bf16"}
    assert 1 <= d["gpu_launches"] <= d["steps"]
    assert d["cpu_baseline"]["cores"] >= 1 and d["cpu_baseline"]["sample"]
    # round 2: parity of the timed output, BASELINE configs[0] literally, the other configs, the host-types leg
    assert d["parity_sample"]["tolerance_used"] <= 1.0 and d["parity_sample"]["winner_agreement"] == 1.0
    assert d["cpu_baseline_cfg1"]["threads"] == 1 and d["cpu_baseline_cfg1"]["reps"] >= 1000 and d["gpu_cfg1"]["reps"] >= 1000
    for k in ("cfg3_strong", "cfg4_fp32", "cfg4_tf32"):
        assert d["extra"][k]["parity_sample"]["tolerance_used"] <= 1.0, k
    assert all(rt["check_passed"] for rt in d["extra"]["cfg5_pool"]["routes"])
    assert set(d["e2e_from_reference_types"]) == {"cfg2_shape", "cfg3_shape", "cfg4_shape"}
    assert d["host_device_paths_agree"] is True


def test_committed_multi_gpu_line_carries_both_pool_routes():
    d = json.loads(open(os.path.join(ROOT, "profiles", "r2_bench_8gpu.json")).read().strip().splitlines()[-1])
    assert d["n_gpus"] == 8 and d["scaling"] == "weak" and len(d["per_rank_ms_per_step"]) == 8
    routes = {rt["route"]: rt for rt in d["extra"]["cfg5_pool"]["routes"]}
    assert set(routes) == {"nccl_allgather", "fused_p2p_gather"} and all(rt["check_passed"] for rt in routes.values())
    assert d["extra"]["cfg5_pool"]["rows_total"] == 262144
    assert d["extra"]["cfg3_strong"]["requests_total"] == 65536
