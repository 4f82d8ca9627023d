"""Force-evaluation throughput across the BASELINE.json configurations (GPU box): ms per evaluation and G-int/s.
Usage: python tools/perf_survey.py [quick]"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import nbtest  # noqa: E402

lib = nbtest.Lib()
CASES = [
    ("C2 random 65536 PAIRWISE fp32", lambda: nbtest.gen_random(65536, seed=42), nbtest.PAIRWISE, nbtest.FP32, 10),
    ("C2 random 65536 PAIRWISE fp64", lambda: nbtest.gen_random(65536, seed=42), nbtest.PAIRWISE, nbtest.FP64, 5),
    ("C4 galaxy 262144 ADAPTIVE fp32", lambda: nbtest.gen_galaxy(87381, 2, seed=42), nbtest.ADAPTIVE, nbtest.FP32, 5),
    ("C4 galaxy 262144 ADAPTIVE fp64", lambda: nbtest.gen_galaxy(87381, 2, seed=42), nbtest.ADAPTIVE, nbtest.FP64, 2),
    ("   galaxy 262144 PAIRWISE fp32", lambda: nbtest.gen_galaxy(87381, 2, seed=42), nbtest.PAIRWISE, nbtest.FP32, 5),
    ("   galaxy 262144 PAIRWISE fp64", lambda: nbtest.gen_galaxy(87381, 2, seed=42), nbtest.PAIRWISE, nbtest.FP64, 2),
    ("   spiral 262144 PAIRWISE fp32", lambda: nbtest.gen_spiral(262144, seed=42), nbtest.PAIRWISE, nbtest.FP32, 5),
    ("C3 spiral 1048576 PAIRWISE fp32", lambda: nbtest.gen_spiral(1048576, seed=42), nbtest.PAIRWISE, nbtest.FP32, 3),
    ("   spiral 1048576 ADAPTIVE fp32", lambda: nbtest.gen_spiral(1048576, seed=42), nbtest.ADAPTIVE, nbtest.FP32, 3),
    ("   spiral 1048576 PAIRWISE_AVX2_FP32 (Plummer) fp32", lambda: nbtest.gen_spiral(1048576, seed=42), nbtest.AVX2, nbtest.FP32, 3),
]
if len(sys.argv) > 1 and sys.argv[1] == "quick":
    CASES = CASES[:3]
out = []
for name, gen, force, prec, steps in CASES:
    pm, v = gen()
    n = pm.shape[0]
    with lib.context(n, prec) as ctx:
        ctx.set_methods(nbtest.VV, force, 1)
        ctx.upload(pm, v)
        ctx.compute_accel()
        ctx.step(0.01, 2)
        ctx.sync()
        ctx.force_timing(reset=True)
        ctx.timer_start()
        ctx.step(0.01, steps)
        total = ctx.timer_stop()
        fk, cnt = ctx.force_timing(reset=True)
    rec = {"case": name, "n": n, "force_ms": fk, "step_ms": total / steps, "Gint_per_s": n * n / (fk * 1e-3) / 1e9,
           "step_Gint_per_s": n * n / (total / steps * 1e-3) / 1e9}
    out.append(rec)
    print(f"{name:34s} force {fk:9.3f} ms  step {total / steps:9.3f} ms  {rec['Gint_per_s']:8.1f} Gint/s (kernel)  {rec['step_Gint_per_s']:8.1f} Gint/s (step)", flush=True)
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(out, open(os.path.join(ROOT, "gpurun_out", "perf_survey.json"), "w"), indent=1)
