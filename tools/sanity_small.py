"""Small all-paths run for compute-sanitizer (memcheck / racecheck, one tool per gpurun call):
every kernel variant, force law, precision and the integrators on a few thousand particles."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import nbtest  # noqa: E402

lib, oracle = nbtest.Lib(), nbtest.Oracle()
pm, v = nbtest.gen_spiral(2500, seed=3)
worst = 0.0
for prec in (nbtest.FP32, nbtest.FP64):
    for force in (nbtest.PAIRWISE, nbtest.AVX2, nbtest.ADAPTIVE):
        ref = oracle.forces(pm, force)
        for variant in ((1, 2, 3) if prec == nbtest.FP32 else (0,)):
            with lib.context(2500, prec) as ctx:
                ctx.set_option("force_variant", variant)
                ctx.set_option("resort_interval", 2)
                ctx.set_methods(nbtest.VV, force, 1)
                ctx.upload(pm, v)
                ctx.compute_accel()
                err = nbtest.rel_err(ctx.accel(), ref).max()
                worst = max(worst, err)
                ctx.step(0.04, 3)
                ctx.set_methods(nbtest.EULER, force, 1)
                ctx.step(0.04, 2)
                e = ctx.energy(); m = ctx.momentum(); p, w = ctx.download()
                assert np.isfinite(p).all() and np.isfinite(e).all()
            print(f"prec={prec} force={force} variant={variant}: accel err {err:.2e}", flush=True)
print("sanity_small ok, worst accel err", worst)
