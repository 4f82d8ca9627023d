"""FILE_CSV loop timing: nb_run_simulation at N = 65,536 (RANDOM, PAIRWISE, Velocity Verlet, FP32 mode), 50 steps with a
snapshot after every step (51 snapshots incl. the initial one), wall clock of the whole call; and the energy alone."""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests"))
import nbtest
lib = nbtest.Lib()
n = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
pm, v = nbtest.gen_random(n, seed=42)
for rep in range(3):
    t0 = time.perf_counter()
    out = lib.run_simulation(pm, v, nbtest.VV, nbtest.PAIRWISE, 0.01, 50, 1, precision=nbtest.FP32)
    dt = time.perf_counter() - t0
    print(f"nb_run_simulation N={n} 50 steps, 51 snapshots: {dt * 1e3:.1f} ms (E0={out[3][0, 1]:.12e}, E50={out[3][50, 1]:.12e})")
with lib.context(n, nbtest.FP32) as ctx:
    ctx.upload(pm, v); ctx.compute_accel(); ctx.energy()
    t0 = time.perf_counter()
    for _ in range(10):
        e = ctx.energy()
    print(f"nb_energy N={n}: {(time.perf_counter() - t0) * 100:.3f} ms per call, E={e[0]:.12e}")
    t0 = time.perf_counter(); ctx.step(0.01, 20); ctx.sync()
    print(f"nb_step N={n}: {(time.perf_counter() - t0) * 50:.3f} ms per step")
