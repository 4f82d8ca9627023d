"""CPU tests of the C-ABI boundary: the library loads without a GPU, exports every symbol include/nbody_b200.h declares,
its host-only entry points work, and device entry points fail loudly (no CPU fallback) when there is no GPU."""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest

import nbtest

HEADER = os.path.join(nbtest.ROOT, "include", "nbody_b200.h")


def declared_functions():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(nb_[a-z0-9_]+)\s*\(", src)))


def test_header_declares_the_boundary():
    names = declared_functions()
    for must in ("nb_create", "nb_destroy", "nb_upload", "nb_download", "nb_step", "nb_compute_accel", "nb_get_accel",
                 "nb_energy", "nb_momentum", "nb_run_benchmark", "nb_run_simulation", "nb_launch_force",
                 "nb_comm_init", "nb_comm_unique_id", "nb_partition_bounds", "nb_last_error"):
        assert must in names


def test_library_exports_every_declared_symbol(lib):
    for name in declared_functions():
        assert hasattr(lib.l, name), f"{name} declared in include/nbody_b200.h but not exported"
    out = subprocess.run(["nm", "-D", "--defined-only", nbtest.LIB_PATH], capture_output=True, text=True).stdout
    exported = set(re.findall(r" T (nb_[a-z0-9_]+)", out))
    assert set(declared_functions()) <= exported


def test_library_is_sm100a_only(lib):
    assert lib.l.nb_built_sm() == 100
    assert lib.l.nb_abi_version() >= 1
    out = subprocess.run(["cuobjdump", "--list-elf", nbtest.LIB_PATH], capture_output=True, text=True).stdout
    archs = set(re.findall(r"sm_(\d+a?)", out))
    assert archs == {"100a"}, archs


def test_product_does_not_link_the_oracle():
    out = subprocess.run(["ldd", nbtest.LIB_PATH], capture_output=True, text=True).stdout
    assert "nboracle" not in out and "nbref" not in out
    syms = subprocess.run(["nm", "-D", nbtest.LIB_PATH], capture_output=True, text=True).stdout
    assert "nbo_" not in syms and "nbref_" not in syms


def test_packed_kernels_use_blackwell_instructions():
    sass = subprocess.run(["cuobjdump", "-sass", nbtest.LIB_PATH], capture_output=True, text=True).stdout
    for mnem in ("FFMA2", "FADD2", "FMUL2", "MUFU.RSQ", "UBLKCP", "SYNCS"):
        assert mnem in sass, f"{mnem} not found in SASS"


@pytest.mark.parametrize("n,world", [(1000, 1), (65536, 2), (65536, 8), (1048576, 8), (16777216, 8), (262144, 4),
                                     (25, 2), (1025, 2), (3000, 8)])
def test_partition_bounds(lib, n, world):
    covered = 0
    prev_end = 0
    sizes = []
    for r in range(world):
        b, c = lib.partition_bounds(n, r, world)
        assert b == min(prev_end, n) or c == 0
        assert b % 1024 == 0 or c == 0
        prev_end = b + c
        covered += c
        sizes.append(c)
    assert covered == n
    # balanced up to one alignment unit
    assert max(sizes) - min(s for s in sizes if s > 0 or world == 1) <= 1024 or n < 1024 * world * 2


def test_partition_bounds_rejects_bad_arguments(lib):
    b, c = C.c_long(0), C.c_long(0)
    assert lib.l.nb_partition_bounds(0, 0, 1, C.byref(b), C.byref(c)) == 1
    assert lib.l.nb_partition_bounds(10, 2, 2, C.byref(b), C.byref(c)) == 1
    assert "nb_partition_bounds" in lib.err()


def test_no_gpu_fails_loudly(lib):
    if lib.l.nb_device_count() > 0:
        pytest.skip("a GPU is present")
    h = C.c_void_p(None)
    rc = lib.l.nb_create(C.byref(h), 1000, 0, nbtest.FP32)
    assert rc == 5 and not h.value                      # NB_ERR_NO_DEVICE, never a CPU fallback
    assert "no CUDA device" in lib.err()
    pm, v = nbtest.gen_random(10)
    with pytest.raises(nbtest.NbError) as ei:
        lib.run_benchmark(pm, v, nbtest.VV, nbtest.PAIRWISE, 0.01, 1)
    assert ei.value.code == 5


def test_argument_validation_needs_no_gpu(lib):
    h = C.c_void_p(None)
    assert lib.l.nb_create(C.byref(h), 0, 0, nbtest.FP32) == 1
    assert lib.l.nb_create(C.byref(h), 100, 0, 7) == 1
    assert lib.l.nb_create(None, 100, 0, 0) == 1
    assert lib.l.nb_set_methods(None, 0, 0, 1) == 1
    assert lib.l.nb_step(None, 0.1, 1) == 1
    assert lib.l.nb_destroy(None) == 0


def test_whole_run_entry_points_validate_before_touching_a_device(lib):
    pm = np.zeros((4, 4)); v = np.zeros((4, 4)); pm[:, 3] = 1.0; pm[:, 0] = np.arange(4)
    fr = (C.c_long * 2)()
    for w, h in ((0, 10), (10, -1), (20000, 10)):
        assert lib.l.nb_run_visualization(4, nbtest._p(pm), nbtest._p(v), 1, 0, 0.01, 3, 1, w, h, 0.0, None, 1, None, fr, 0, 1, -1) == 1
        assert "image" in lib.err()
    assert lib.l.nb_run_visualization(4, nbtest._p(pm), nbtest._p(v), 1, 0, 0.01, -3, 1, 64, 64, 0.0, None, 1, None, fr, 0, 1, -1) == 1
    # NULL output arrays of the FILE_CSV driver
    assert lib.l.nb_run_simulation(4, nbtest._p(pm), nbtest._p(v), 1, 0, 0.01, 3, 1, None, None, None, 4, 1, 1, -1) == 1


def test_partition_bounds_properties(lib):
    """Any (n, world): the slices are disjoint, ordered, tile-aligned, cover [0, n) exactly and all but the tail have the
    same padded size (what the all-gather needs)."""
    from hypothesis import given, settings, strategies as st

    @settings(max_examples=300, deadline=None)
    @given(n=st.integers(1, 1 << 30), world=st.integers(1, 64))
    def check(n, world):
        per = None
        end = 0
        for r in range(world):
            b, c = lib.partition_bounds(n, r, world)
            assert 0 <= c and b + c <= n
            if c:
                assert b == end and b % 1024 == 0
                end = b + c
                if b + c < n:                                   # a full slice
                    per = per or c
                    assert c == per and c % 1024 == 0
                elif per:
                    assert c <= per
        assert end == n
    check()
