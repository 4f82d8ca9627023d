"""nb_force_noalloc — the one entry point that never allocates — against the context-based nb_launch_force.

Collected last on purpose (file name): if it fails, a run that stops at the first failure has already executed every other
test. It has to pass both alone in a fresh process (first call = set-up of the built-in scratch block) and after the other
GPU tests (scratch block never touched before, contexts created and destroyed earlier)."""
import ctypes as C

import numpy as np
import pytest

import nbtest
from nbtest import ADAPTIVE, FP32, PAIRWISE

pytestmark = pytest.mark.gpu

MAX_N = 32768          # NB_NOALLOC_MAX_N in include/nbody_b200.h
TOL = 1e-5             # the FP32 tolerance used by tests/test_gpu_parity.py


def _bind(lib):
    fn = lib.l.nb_force_noalloc
    fn.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_long, C.c_double, C.c_void_p]
    fn.restype = C.c_int
    return fn


def _launch_force(lib, ctx, force, d_pos, d_acc, n, stream):
    lib.check(lib.l.nb_launch_force(ctx.h, force, C.c_void_p(d_pos.data_ptr()), C.c_void_p(d_acc.data_ptr()), -1.0, n,
                                    C.c_void_p(stream.cuda_stream)))


def test_noalloc_matches_existing_entry_point_twice_in_a_row(lib, oracle):
    import torch
    fn = _bind(lib)
    stream = torch.cuda.Stream()
    # two calls in a row on the same scratch block: a big system first, then a smaller, different one, so that the
    # second call runs on top of the first one's left-overs (tiles and partial sums beyond its own size stay stale)
    cases = [(PAIRWISE, nbtest.gen_spiral(20000, seed=21)[0]), (ADAPTIVE, nbtest.gen_galaxy(1500, 2, seed=22)[0]),
             (PAIRWISE, nbtest.gen_random(333, seed=23)[0])]
    for force, pm in cases:
        n = pm.shape[0]
        avg = float(oracle.average_mass(pm))
        d_pos = torch.from_numpy(pm).cuda()
        d_new = torch.full_like(d_pos, float("nan"))
        d_old = torch.zeros_like(d_pos)
        torch.cuda.synchronize()                      # the fills above ran on torch's default stream
        with torch.cuda.stream(stream):
            rc = fn(force, C.c_void_p(d_pos.data_ptr()), C.c_void_p(d_new.data_ptr()), n, avg, C.c_void_p(stream.cuda_stream))
            assert rc == 0, lib.err()
            stream.synchronize()
            with lib.context(n, FP32) as ctx:
                _launch_force(lib, ctx, force, d_pos, d_old, n, stream)
                stream.synchronize()
        a_new, a_old = d_new.cpu().numpy(), d_old.cpu().numpy()
        assert np.isfinite(a_new).all() and np.all(a_new[:, 3] == 0.0)
        assert nbtest.rel_err(a_new, a_old).max() <= TOL, (force, n, nbtest.rel_err(a_new, a_old).max())
        assert nbtest.rel_err(a_new, oracle.forces(pm, force)).max() <= TOL
        np.testing.assert_array_equal(d_pos.cpu().numpy(), pm)            # inputs untouched


def test_noalloc_same_input_twice_is_bit_identical(lib):
    import torch
    fn = _bind(lib)
    pm = nbtest.gen_spiral(MAX_N, seed=5)[0]                              # exactly at the capacity limit
    d_pos = torch.from_numpy(pm).cuda()
    outs = []
    for _ in range(2):
        d_acc = torch.zeros_like(d_pos)
        assert fn(PAIRWISE, C.c_void_p(d_pos.data_ptr()), C.c_void_p(d_acc.data_ptr()), MAX_N, 0.0, None) == 0, lib.err()
        torch.cuda.synchronize()
        outs.append(d_acc.cpu().numpy())
    assert np.array_equal(outs[0], outs[1])


def test_noalloc_rejects_what_it_cannot_do(lib):
    import torch
    fn = _bind(lib)
    d = torch.zeros((8, 4), dtype=torch.float64, device="cuda")
    p = C.c_void_p(d.data_ptr())
    assert fn(PAIRWISE, p, p, MAX_N + 1, 0.0, None) == 1
    assert "capacity" in lib.err() and str(MAX_N) in lib.err()
    assert fn(PAIRWISE, None, p, 8, 0.0, None) == 1
    assert fn(ADAPTIVE, p, p, 8, 0.0, None) == 1 and "avg_mass" in lib.err()
    assert fn(nbtest.BARNES_HUT, p, p, 8, 0.0, None) == 4
    assert fn(PAIRWISE, p, p, 0, 0.0, None) == 1
