"""nb_force_noalloc — the one entry point that never allocates — against the context-based nb_launch_force.

Collected last on purpose (file name): if it fails, a run that stops at the first failure has already executed every other
test. It has to pass both alone in a fresh process (first call = set-up of the built-in scratch block) and after the other
GPU tests (scratch block never touched before, contexts created and destroyed earlier)."""
import ctypes as C

import numpy as np
import pytest

import nbtest
from nbtest import ADAPTIVE, AVX2, FP32, PAIRWISE

pytestmark = pytest.mark.gpu

MAX_N = 32768          # NB_NOALLOC_MAX_N in include/nbody_b200.h
TOL = 1e-5             # the FP32 tolerance used by tests/test_gpu_parity.py
CLUSTERED_TOL = 1e-5   # the same bound on strongly clustered input (observed 3.5e-6; the header quotes the observed values)


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
    far = nbtest.gen_spiral(12000, seed=24)[0]
    far[:, :3] += [5.0e4, -3.0e4, 1.0e4]              # a system far from the origin: two-float coordinates must absorb the offset
    cases = [(PAIRWISE, nbtest.gen_spiral(20000, seed=21)[0]), (ADAPTIVE, nbtest.gen_galaxy(1500, 2, seed=22)[0]),
             (PAIRWISE, nbtest.gen_random(333, seed=23)[0]), (AVX2, nbtest.gen_spiral(6000, seed=25)[0]), (PAIRWISE, far)]
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
        err = nbtest.rel_err(a_new, oracle.forces(pm, force)).max()
        print(f"nb_force_noalloc force={force} n={n}: max rel err vs oracle {err:.2e}")
        assert err <= TOL
        np.testing.assert_array_equal(d_pos.cpu().numpy(), pm)            # inputs untouched


def test_noalloc_clustered_input(lib, oracle):
    """Galaxy + planets (9000 stars with two close companions each, one 4e6 Msun centre): the input class where a particle's
    two nearest neighbours pull in opposite directions and single FP32 pair terms lose digits in the sum. The bound asserted
    here is the one include/nbody_b200.h states for this entry point."""
    import torch
    fn = _bind(lib)
    for force in (PAIRWISE, ADAPTIVE):
        pm = nbtest.gen_galaxy(9000, 2, seed=26)[0]
        n = pm.shape[0]
        assert n <= MAX_N
        d_pos = torch.from_numpy(pm).cuda()
        d_acc = torch.zeros_like(d_pos)
        torch.cuda.synchronize()
        assert fn(force, C.c_void_p(d_pos.data_ptr()), C.c_void_p(d_acc.data_ptr()), n, float(oracle.average_mass(pm)), None) == 0, lib.err()
        torch.cuda.synchronize()
        err = nbtest.rel_err(d_acc.cpu().numpy(), oracle.forces(pm, force))
        print(f"nb_force_noalloc galaxy+planets n={n} force={force}: max rel err {err.max():.2e}, median {np.median(err):.2e}")
        assert err.max() <= CLUSTERED_TOL


def test_noalloc_really_allocates_nothing(lib):
    """The contract itself: once every kernel of the path has been loaded (first call per force method), a call must leave
    the device's free-memory figure exactly where it was — no cudaMalloc, no pool growth, nothing freed either."""
    import torch
    fn = _bind(lib)
    pm = nbtest.gen_spiral(9000, seed=27)[0]
    d_pos = torch.from_numpy(pm).cuda()
    d_acc = torch.zeros_like(d_pos)
    p_in, p_out = C.c_void_p(d_pos.data_ptr()), C.c_void_p(d_acc.data_ptr())
    for force in (PAIRWISE, AVX2, ADAPTIVE):
        assert fn(force, p_in, p_out, 9000, 1.0, None) == 0, lib.err()    # loads the kernels of this method
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    for rep in range(3):
        for force in (PAIRWISE, AVX2, ADAPTIVE):
            assert fn(force, p_in, p_out, 9000 - 1000 * rep, 1.0, None) == 0, lib.err()
            torch.cuda.synchronize()
            assert torch.cuda.mem_get_info()[0] == free0


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
