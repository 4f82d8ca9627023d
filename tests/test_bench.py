"""bench.py without a GPU: the reference arm (the reference's own CPU path, timed), the roofline arithmetic of every bound, and
the contract keys of the JSON line."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BENCH = os.path.join(ROOT, "bench.py")
sys.path.insert(0, ROOT)


def _line(out):
    return json.loads(out.strip().splitlines()[-1])


def test_reference_arm_prints_the_contract_line():
    r = subprocess.run([sys.executable, BENCH, "--impl", "reference", "--steps", "1", "--warmup", "0", "--ref-n", "2048"],
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
    d = _line(r.stdout)
    assert d["impl"] == "reference" and d["unit"] == "Gint/s" and d["higher_is_better"] is True and d["value"] > 0
    assert d["cpu_baseline"]["kind"] in ("reference", "port") and d["cpu_baseline"]["cores"] >= 1
    assert d["cpu_baseline"]["value"] == d["value"] == d["e2e"]["value"]
    assert d["e2e"]["h2d_bytes_per_step"] == 0 and d["e2e"]["d2h_bytes_per_step"] == 0 and d["gpu_launches"] == 0
    assert "sample" in d["cpu_baseline"] and d["config"]["sample_n"] == 2048


def test_reference_arm_under_torchrun_only_rank_0_works():
    env = dict(os.environ, MASTER_ADDR="127.0.0.1")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
                        "--master-port", "29537", BENCH, "--impl", "reference", "--gpus", "2", "--steps", "1", "--warmup", "0",
                        "--ref-n", "1024"], capture_output=True, text=True, env=env, timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [l for l in r.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1 and json.loads(lines[0])["n_gpus"] == 2


def test_product_arm_refuses_to_run_without_a_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    r = subprocess.run([sys.executable, BENCH, "--steps", "1", "--warmup", "0", "--configs", "none"], capture_output=True, text=True, timeout=300)
    assert r.returncode != 0 and "no CPU path" in (r.stderr + r.stdout)


@pytest.mark.parametrize("spec,bound,peak_t", [
    ({"n": 1048576, "init": "spiral", "force": "PAIRWISE", "prec": "fp32"}, "fp32", 148 * 128 / 13 * 1.965e9),
    ({"n": 262144, "init": "galaxy", "force": "ADAPTIVE_MUTUAL", "prec": "fp32"}, "mufu", 148 * 16 / 3 * 1.965e9),
    ({"n": 262144, "init": "spiral", "force": "PAIRWISE", "prec": "fp64"}, "fp64", None),
    ({"n": 262144, "init": "galaxy", "force": "ADAPTIVE_MUTUAL", "prec": "fp64"}, "fp64", None)])
def test_roofline_of_every_bound(spec, bound, peak_t):
    import bench
    rates = bench.pipe_rates()
    inter = float(spec["n"]) ** 2
    r = bench.roofline_for(spec, 148, 1.965e9, "test", rates, inter, 100.0, 100.5, 2)
    assert r["bound"] == bound and r["unit"] == "TFLOP/s" and r["traffic"] is None
    ach = inter / 0.1
    assert r["achieved_Tint_per_s"] == pytest.approx(ach / 1e12)
    if peak_t is None:
        dp = bench.DP_PAIRWISE if spec["force"] == "PAIRWISE" else bench.DP_ADAPTIVE
        peak_t = 148 * rates["dfma"] * 32 / dp * 1.965e9
    assert r["frac"] == pytest.approx(ach / peak_t) and r["achieved"] / r["peak"] == pytest.approx(r["frac"])
    assert r["share_of_step"] == pytest.approx(100.0 / 100.5)
    if bound == "fp32":
        assert r["executed_mix"]["frac"] == pytest.approx(ach / (148 * rates["ffma2"] * 64 / 11 * 1.965e9))


def test_generators_of_the_bench_configs_have_the_named_sizes():
    import bench
    for spec in bench.EXTRA_CONFIGS:
        if spec["n"] > 300000:
            continue
        pm, v = bench.generate(spec)
        assert pm.shape == (spec["n"], 4) and v.shape == (spec["n"], 4) and (pm[:, 3] >= 0).all()
        assert str(spec["n"]) in bench.describe(spec)
