"""Headless point-splat (nb_render): the CUDA evaluation of the reference viewer's shader arithmetic against the plain-C
restatement in oracle/splat_oracle.c. CPU tests exercise the restatement alone."""
import ctypes as C

import numpy as np
import pytest

import nbtest

_fp = C.POINTER(C.c_float)


def oracle_render(oracle, pm, v, W, H, cam=(0.2, 0.9, 5.0, 0.0, 0.0), scale=-1.0, point_size=30.0, color_mode=1,
                  color=(1, 1, 1, 0.9), clear=(0.02, 0.02, 0.05, 1.0), max_velocity=None):
    l = oracle.l
    l.nbo_splat_ranges.argtypes = [C.c_long, nbtest._dp, nbtest._dp, _fp]
    l.nbo_splat.argtypes = [C.c_long, nbtest._dp, nbtest._dp, C.c_int, C.c_int] + [C.c_float] * 7 + [C.c_int, _fp, _fp] + \
        [C.c_float] * 3 + [C.POINTER(C.c_ubyte)]
    r4 = (C.c_float * 4)()
    l.nbo_splat_ranges(pm.shape[0], nbtest._p(pm), nbtest._p(v), r4)
    mn, mx, sp, md = [float(x) for x in r4]
    if not mn < mx:                                   # ParticleRenderer.cpp:255-262
        mn = mn if mn < 3e38 else 1.0
        mx = np.float32(mn) * np.float32(10.0)
    if scale <= 0:
        scale = float(np.float32(2.0) / np.float32(md)) if md > 0 else 1.0
    vmax = max(sp if max_velocity is None else max_velocity, 0.001)
    img = np.zeros((H, W, 4), dtype=np.uint8)
    col = (C.c_float * 4)(*color); clr = (C.c_float * 4)(*clear)
    rc = l.nbo_splat(pm.shape[0], nbtest._p(pm), nbtest._p(v), W, H, *cam, scale, point_size, color_mode, col, clr,
                     mn, mx, vmax, img.ctypes.data_as(C.POINTER(C.c_ubyte)))
    assert rc == 0
    return img, (mn, mx, vmax, scale)


def gpu_render(lib, ctx, W, H, cam=(0.2, 0.9, 5.0, 0.0, 0.0), scale=-1.0, point_size=30.0, color_mode=1):
    lib.l.nb_render.argtypes = [C.c_void_p, C.c_int, C.c_int] + [C.c_float] * 7 + [C.c_int, _fp, _fp, C.POINTER(C.c_ubyte), _fp]
    img = np.zeros((H, W, 4), dtype=np.uint8)
    info = (C.c_float * 4)()
    lib.check(lib.l.nb_render(ctx.h, W, H, *cam, scale, point_size, color_mode, None, None,
                              img.ctypes.data_as(C.POINTER(C.c_ubyte)), info))
    return img, [float(x) for x in info]


def test_oracle_splat_basic_properties(oracle):
    pm, v = nbtest.gen_spiral(2000, seed=4)
    img, (mn, mx, vmax, scale) = oracle_render(oracle, pm, v, 320, 180)
    clear = np.array([round(0.02 * 255), round(0.02 * 255), round(0.05 * 255), 255], dtype=np.uint8)
    assert img.shape == (180, 320, 4)
    bg = np.all(img == clear, axis=2)
    assert 0.3 < bg.mean() < 0.999                      # particles drawn, background left
    assert mx == pytest.approx(mn * 10.0)               # equal masses: artificial x10 range
    # a single heavy particle at the origin lands at the image centre (camera looks down -z at the origin)
    one = np.array([[0.0, 0.0, 0.0, 1.0], [1.0, 0.0, 0.0, 1e-3]]); vv = np.zeros((2, 4)); vv[0, 0] = 1.0
    img, _ = oracle_render(oracle, one, vv, 201, 101, scale=1.0)
    ys, xs = np.nonzero(~np.all(img == clear, axis=2))
    assert abs(xs.mean() - 100) < 12 and abs(ys.mean() - 50) < 6
    assert img[50, 100, 0] > 200 and img[50, 100, 2] < 60       # fastest particle: red end of the ramp
    # particles behind the camera or outside the frustum are culled
    far = np.array([[0.0, 0.0, 50.0, 1.0], [1e4, 0.0, 0.0, 2.0]])
    img, _ = oracle_render(oracle, far, np.zeros((2, 4)), 64, 64, scale=1.0)
    assert np.all(img == clear)


@pytest.mark.gpu
@pytest.mark.parametrize("gen,n,size", [("spiral", 20000, (640, 360)), ("random", 3000, (320, 200)), ("galaxy", 3001, (400, 300))])
def test_cuda_splat_matches_restatement(lib, oracle, gen, n, size):
    pm, v = {"spiral": lambda: nbtest.gen_spiral(n, seed=6), "random": lambda: nbtest.gen_random(n, seed=6),
             "galaxy": lambda: nbtest.gen_galaxy((n - 1) // 3, 2, seed=6)}[gen]()
    W, H = size
    with lib.context(pm.shape[0], nbtest.FP32) as ctx:
        ctx.upload(pm, v)
        for cam, mode in (((0.2, 0.9, 5.0, 0.0, 0.0), 1), ((-0.7, 2.4, 3.0, 0.3, -0.2), 0)):
            img, info = gpu_render(lib, ctx, W, H, cam=cam, color_mode=mode)
            ref, (mn, mx, vmax, scale) = oracle_render(oracle, pm, v, W, H, cam=cam, color_mode=mode, max_velocity=info[2])
            assert info[0] == pytest.approx(mn) and info[1] == pytest.approx(mx) and info[3] == pytest.approx(scale, rel=1e-6)
            diff = np.abs(img.astype(int) - ref.astype(int))
            # identical arithmetic up to libm / FMA-contraction differences: a pixel at a sprite edge may flip
            assert (diff.max(axis=2) == 0).mean() >= 0.999, (gen, (diff.max(axis=2) > 0).mean())
            assert np.quantile(diff, 0.9999) <= 2
            assert (img[:, :, :3].astype(int).sum(axis=2) > 30).mean() > 0.001     # something was drawn


@pytest.mark.gpu
def test_cuda_splat_order_and_speed_average(lib, oracle):
    """Blending is order dependent (particle index order, LEQUAL depth): two overlapping sprites in both orders; the
    colour-ramp maximum speed is the 0.9/0.1 moving average over successive calls."""
    pm = np.array([[0.0, 0.0, 0.0, 1.0], [0.0, 0.0, 0.5, 1.0], [3.0, 3.0, 0.0, 1e-3]])
    v = np.zeros((3, 4)); v[0, 0] = 1.0; v[1, 1] = 0.2
    imgs = []
    for order in ([0, 1, 2], [1, 0, 2]):
        with lib.context(3, nbtest.FP64) as ctx:
            ctx.upload(pm[order], v[order])
            img, info = gpu_render(lib, ctx, 200, 120, scale=0.2)
            ref, _ = oracle_render(oracle, pm[order], v[order], 200, 120, scale=0.2, max_velocity=info[2])
            assert np.abs(img.astype(int) - ref.astype(int)).max() <= 1
            imgs.append(img)
            img2, info2 = gpu_render(lib, ctx, 200, 120, scale=0.2)
            assert info2[2] == pytest.approx(info[2])            # same speeds: the average does not move
    assert not np.array_equal(imgs[0], imgs[1])
