"""Where the wall time of one nb_run_benchmark-style call goes (host clock around each C-ABI call, device idle between).
Usage (GPU box): python tools/e2e_breakdown.py [N]"""
import ctypes as C
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import nbtest  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1048576
lib = nbtest.Lib()
pm0, v0 = nbtest.gen_spiral(n, seed=42)
pm = torch.from_numpy(pm0).pin_memory().numpy(); v = torch.from_numpy(v0).pin_memory().numpy()
lib.check(lib.l.nb_run_benchmark(n, nbtest._p(pm.copy()), nbtest._p(v.copy()), nbtest.VV, nbtest.PAIRWISE, 0.04, 0, 256, 0, 1, 0))
for rep in range(2):
    T = {}
    def tick(name, t0):
        T[name] = (time.perf_counter() - t0) * 1e3
    h = C.c_void_p()
    t0 = time.perf_counter(); lib.check(lib.l.nb_create(C.byref(h), n, 0, 0)); tick("create", t0)
    t0 = time.perf_counter(); lib.check(lib.l.nb_set_methods(h, nbtest.VV, nbtest.PAIRWISE, 1)); tick("set_methods", t0)
    t0 = time.perf_counter(); lib.check(lib.l.nb_upload(h, nbtest._p(pm), nbtest._p(v))); tick("upload (H2D + sort + pack, sync)", t0)
    t0 = time.perf_counter(); lib.check(lib.l.nb_compute_accel(h)); tick("compute_accel enqueue (incl. partial-slot cudaMalloc)", t0)
    t0 = time.perf_counter(); lib.check(lib.l.nb_sync(h)); tick("  ... wait", t0)
    t0 = time.perf_counter(); lib.check(lib.l.nb_step(h, 0.04, 1)); tick("step enqueue", t0)
    t0 = time.perf_counter(); lib.check(lib.l.nb_sync(h)); tick("  ... wait", t0) if False else T.__setitem__("  ... wait (step)", (time.perf_counter() - t0) * 1e3)
    t0 = time.perf_counter(); lib.check(lib.l.nb_download(h, nbtest._p(pm), nbtest._p(v))); tick("download (gather + D2H, sync)", t0)
    t0 = time.perf_counter(); lib.check(lib.l.nb_destroy(h)); tick("destroy", t0)
    pm[:] = pm0; v[:] = v0
    print(f"--- repetition {rep}: total {sum(T.values()):.2f} ms")
    for k, val in T.items():
        print(f"{k:58s} {val:9.3f} ms")
