"""How the production kernel classifies j-tiles (far / refined per warp / FP64 sub-tiles) for one force evaluation.
Usage (GPU box): python tools/tile_stats.py [spiral|galaxy|random] [N] [kappa_milli]"""
import ctypes as C
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import nbtest  # noqa: E402

gen = sys.argv[1] if len(sys.argv) > 1 else "spiral"
n = int(sys.argv[2]) if len(sys.argv) > 2 else 1048576
kap = int(sys.argv[3]) if len(sys.argv) > 3 else 250
pm, v = {"spiral": lambda: nbtest.gen_spiral(n, seed=42), "galaxy": lambda: nbtest.gen_galaxy((n - 1) // 3, 2, seed=42),
         "random": lambda: nbtest.gen_random(n, seed=42)}[gen]()
lib = nbtest.Lib()
with lib.context(pm.shape[0]) as ctx:
    ctx.set_option("far_kappa_milli", kap)
    ctx.set_methods(nbtest.VV, nbtest.PAIRWISE, 1)
    ctx.upload(pm, v)
    ctx.compute_accel(); ctx.sync()
    ctx.set_option("collect_stats", 1)
    ctx.compute_accel(); ctx.sync()
    out = (C.c_ulonglong * 5)()
    lib.check(lib.l.nb_tile_stats(ctx.h, out))
    ms = ctx.last_force_ms()
    import numpy as np
    cap = (pm.shape[0] // 128 + 8) * 64
    buf = np.zeros(cap, dtype=np.uint64); nn = C.c_long(0)
    lib.l.nb_tile_stats_warps.argtypes = [C.c_void_p, C.c_void_p, C.c_long, C.POINTER(C.c_long)]
    lib.check(lib.l.nb_tile_stats_warps(ctx.h, buf.ctypes.data_as(C.c_void_p), cap, C.byref(nn)))
    wc = buf[buf > 0].astype(np.int64)
far, near, wfar, w64, wmax = [int(x) for x in out]
tot = far + near
print("per-warp FP64 sub-tile counts (nonzero entries): n", len(wc), "mean", wc.mean() if len(wc) else 0, "quantiles 50/90/99/99.9/max",
      [int(np.quantile(wc, q)) for q in (0.5, 0.9, 0.99, 0.999, 1.0)] if len(wc) else [], "share of work in warps above 4x median:",
      float(wc[wc > 4 * np.median(wc)].sum()) / max(1, wc.sum()) if len(wc) else 0)
print(f"{gen} N={pm.shape[0]} kappa={kap/1000}: tiles per CTA-pass total {tot}: far {far} ({100*far/tot:.2f}%), refined {near} ({100*near/tot:.2f}%); "
      f"sub-tiles in refined tiles: warp-frame far {wfar}, FP64 {w64} ({100*w64/max(1,wfar+w64):.1f}%); "
      f"FP64 share of all pair work {100*w64*64/(8*tot*256):.3f}%; busiest warp {wmax} FP64 sub-tiles (mean {w64/max(1,tot/ (pm.shape[0]/256) )/8:.1f} per warp and split... ); force {ms:.2f} ms (with counters)")
