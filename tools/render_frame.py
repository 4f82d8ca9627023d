"""Renders the current state of a generated system with nb_render and writes a binary PPM (no graphics stack needed).
Usage (GPU box): python tools/render_frame.py out.ppm [spiral|galaxy|random] [N] [steps]"""
import ctypes as C
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import nbtest  # noqa: E402

out = sys.argv[1] if len(sys.argv) > 1 else "frame.ppm"
gen = sys.argv[2] if len(sys.argv) > 2 else "spiral"
n = int(sys.argv[3]) if len(sys.argv) > 3 else 100000
steps = int(sys.argv[4]) if len(sys.argv) > 4 else 0
pm, v = {"spiral": lambda: nbtest.gen_spiral(n, seed=42), "galaxy": lambda: nbtest.gen_galaxy((n - 1) // 3, 2, seed=42),
         "random": lambda: nbtest.gen_random(n, seed=42)}[gen]()
lib = nbtest.Lib()
W, H = 1280, 720                                           # the reference's window (VisualizationUtils.cpp:15-16)
_fp = C.POINTER(C.c_float)
lib.l.nb_render.argtypes = [C.c_void_p, C.c_int, C.c_int] + [C.c_float] * 7 + [C.c_int, _fp, _fp, C.POINTER(C.c_ubyte), _fp]
with lib.context(pm.shape[0]) as ctx:
    ctx.set_methods(nbtest.VV, nbtest.PAIRWISE, 1)
    ctx.upload(pm, v)
    if steps:
        ctx.step(0.04, steps)
    img = np.zeros((H, W, 4), dtype=np.uint8); info = (C.c_float * 4)()
    lib.check(lib.l.nb_render(ctx.h, W, H, 0.2, 0.9, 5.0, 0.0, 0.0, -1.0, 30.0, 1, None, None,
                              img.ctypes.data_as(C.POINTER(C.c_ubyte)), info))
with open(out, "wb") as f:
    f.write(b"P6\n%d %d\n255\n" % (W, H))
    f.write(img[:, :, :3].tobytes())
print("wrote", out, "mass range / max speed / scale:", list(info))
