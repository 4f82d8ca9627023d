"""Accuracy diagnosis of the FP32 force kernel: per-particle relative error against the CPU oracle next to the
cancellation ratio C_i = sum_j |a_ij| / |a_i| (how much bigger the individual pair terms are than their vector sum).
Usage (GPU box): python tools/diag_accuracy.py [spiral|galaxy|random] [N] [pairwise|adaptive]"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import nbtest  # noqa: E402

gen = sys.argv[1] if len(sys.argv) > 1 else "spiral"
n = int(sys.argv[2]) if len(sys.argv) > 2 else 30000
force = nbtest.ADAPTIVE if (len(sys.argv) > 3 and sys.argv[3] == "adaptive") else nbtest.PAIRWISE
pm, v = {"spiral": lambda: nbtest.gen_spiral(n, seed=12), "galaxy": lambda: nbtest.gen_galaxy((n - 1) // 3, 2, seed=12),
         "random": lambda: nbtest.gen_random(n, seed=12)}[gen]()
n = pm.shape[0]
lib, oracle = nbtest.Lib(), nbtest.Oracle()
ref = oracle.forces(pm, force)


def gpu(prec=nbtest.FP32, **opts):
    with lib.context(n, prec) as ctx:
        for k, val in opts.items():
            ctx.set_option(k, val)
        ctx.set_methods(nbtest.VV, force, 1)
        ctx.upload(pm, v)
        ctx.compute_accel()
        return ctx.accel()


def cancel_ratio(i):
    d = pm[:, :3] - pm[i, :3]
    r2 = (d * d).sum(1); r2[i] = np.inf
    terms = nbtest.G * pm[:, 3:4] * d / (r2 ** 1.5)[:, None]
    return np.linalg.norm(terms, axis=1).sum() / np.linalg.norm(terms.sum(0))


for name, opts in (("v2 default", {}), ("v2 kappa=0 (always two-float)", {"far_kappa_milli": 0}), ("v1 global two-float", {"force_variant": 1}),
                   ("v2 kappa=0.25", {"far_kappa_milli": 250}), ("v2 kappa=4", {"far_kappa_milli": 4000})):
    a = gpu(**opts)
    err = nbtest.rel_err(a, ref)
    w = np.argsort(err)[-3:][::-1]
    print(f"{name:32s} max {err.max():.3e} p99.9 {np.quantile(err, 0.999):.3e} median {np.median(err):.3e}  worst:",
          [(int(i), f"{err[i]:.2e}", f"C={cancel_ratio(i):.0f}") for i in w])
c_all = np.array([cancel_ratio(i) for i in np.random.default_rng(0).choice(n, 200, replace=False)])
print(f"cancellation ratio of 200 random particles: median {np.median(c_all):.1f} max {c_all.max():.1f}")
