"""Sweep of the far/near opening parameter (far_kappa_milli) and of the far kernel's j-split count at the smaller
BASELINE configs: force-evaluation time (CUDA events inside the library, mean of 8) and max relative acceleration error
against the library's own FP64 mode. Usage (GPU box): python tools/kappa_sweep.py [out.json] [json-spec]
where json-spec overrides {"cases": [[gen, n], ...], "kappas": [...], "splits": [...]}."""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import nbtest  # noqa: E402

CASES = [("random", 65536), ("random", 131072), ("galaxy", 262144), ("spiral", 262144)]
KAPPAS = [150, 250, 400, 600, 1000]
SPLITS = [0, 4, 9, 14, 18, 23, 28, 32]
if len(sys.argv) > 2:
    spec = json.loads(sys.argv[2])
    CASES = [tuple(c) for c in spec.get("cases", CASES)]
    KAPPAS = spec.get("kappas", KAPPAS)
    SPLITS = [0] + spec.get("splits", SPLITS[1:])
lib = nbtest.Lib()
rows = []


def timed(ctx, reps=8):
    ctx.compute_accel(); ctx.sync(); ctx.force_timing(True)
    for _ in range(reps):
        ctx.compute_accel()
    ctx.sync()
    return ctx.force_timing(True)[0]


for gen, n in CASES:
    pm, v = {"spiral": lambda: nbtest.gen_spiral(n, seed=42), "galaxy": lambda: nbtest.gen_galaxy((n - 1) // 3, 2, seed=42),
             "random": lambda: nbtest.gen_random(n, seed=42)}[gen]()
    with lib.context(n, nbtest.FP64) as ctx:
        ctx.set_methods(nbtest.VV, nbtest.PAIRWISE, 1); ctx.upload(pm, v); ctx.compute_accel()
        ref = ctx.accel()
    with lib.context(n, nbtest.FP32) as ctx:
        ctx.set_methods(nbtest.VV, nbtest.PAIRWISE, 1); ctx.upload(pm, v)
        for kap in KAPPAS:
            ctx.set_option("far_kappa_milli", kap)
            ms = timed(ctx)
            err = nbtest.rel_err(ctx.accel(), ref)
            rows.append({"gen": gen, "n": n, "kappa_milli": kap, "j_splits": 0, "force_ms": ms, "Gint_s": n * n / ms / 1e6,
                         "err_max": float(err.max()), "err_median": float(np.median(err))})
            print(rows[-1], flush=True)
        ctx.set_option("far_kappa_milli", 250)
        for S in SPLITS[1:]:
            ctx.set_option("j_splits", S)
            ms = timed(ctx)
            rows.append({"gen": gen, "n": n, "kappa_milli": 250, "j_splits": S, "force_ms": ms, "Gint_s": n * n / ms / 1e6})
            print(rows[-1], flush=True)
        ctx.set_option("j_splits", 0)
out = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "gpurun_out", "kappa_sweep.json")
os.makedirs(os.path.dirname(out), exist_ok=True)
json.dump(rows, open(out, "w"), indent=0)
