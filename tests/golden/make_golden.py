"""Regenerates tests/golden/ref_vectors.npz from the reference's OWN CPU code (oracle/_ref/libnbref.so, built from
/root/reference by `make -C oracle ref`). Run in the build container only:  python tests/golden/make_golden.py
The committed .npz is what travels to the GPU box, where /root/reference does not exist."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import nbtest  # noqa: E402

ref = nbtest.Ref.try_load()
assert ref is not None, "build oracle/_ref first: make -C oracle ref"
ref.set_threads(4)
out = {}

# KAT-1: three bodies (Sun, Earth-mass at 1 AU, Jupiter-mass at 5.2 AU)
pm3 = np.array([[0, 0, 0, 1.0], [1, 0, 0, 3e-6], [0, 5.2, 0, 1e-3]], dtype=np.float64)
out["kat1_pm"] = pm3
out["kat1_pairwise"] = ref.forces(pm3, nbtest.PAIRWISE)
out["kat1_adaptive"] = ref.forces(pm3, nbtest.ADAPTIVE)

# KAT-2: data_in/sistema.dat, dt = 0.04 yr (config.json defaults)
pm, v = nbtest.load_sistema()
rpm, rv = ref.init_file(os.path.join(nbtest.ROOT, "data_in", "sistema.dat"), 25)
# the reference is built with -ffast-math (x / c -> x * (1/c)), so the two loaders agree to 1 ulp, not bitwise
assert np.allclose(pm, rpm, rtol=1e-15, atol=0) and np.allclose(v, rv, rtol=1e-15, atol=0), "python loader != reference loader"
pm, v = rpm, rv
out["sis_pm"], out["sis_v"] = pm, v
out["sis_E0"] = ref.energy(pm, v)
for name, integ, force in [("vv_pw", nbtest.VV, nbtest.PAIRWISE), ("vv_ad", nbtest.VV, nbtest.ADAPTIVE),
                           ("eu_pw", nbtest.EULER, nbtest.PAIRWISE), ("eu_ad", nbtest.EULER, nbtest.ADAPTIVE)]:
    p2, v2 = ref.run(pm, v, integ, force, 0.04, 2500)
    out[f"sis_E2500_{name}"] = ref.energy(p2, v2)
    p3, v3 = ref.run(pm, v, integ, force, 0.04, 50)
    out[f"sis_pm50_{name}"], out[f"sis_v50_{name}"] = p3, v3
    out[f"sis_E50_{name}"] = ref.energy(p3, v3)

# Random box (RANDOM generator distributions), N = 96 and N = 333 (ragged: not a multiple of any tile size)
for n in (96, 333):
    pmr, vr = nbtest.gen_random(n, seed=7)
    out[f"rnd{n}_pm"], out[f"rnd{n}_v"] = pmr, vr
    out[f"rnd{n}_pairwise"] = ref.forces(pmr, nbtest.PAIRWISE)
    out[f"rnd{n}_adaptive"] = ref.forces(pmr, nbtest.ADAPTIVE)
    out[f"rnd{n}_E0"] = ref.energy(pmr, vr)
    p2, v2 = ref.run(pmr, vr, nbtest.VV, nbtest.PAIRWISE, 0.01, 20)
    out[f"rnd{n}_vv_pm"], out[f"rnd{n}_vv_v"], out[f"rnd{n}_vv_E"] = p2, v2, ref.energy(p2, v2)
    p2, v2 = ref.run(pmr, vr, nbtest.EULER, nbtest.ADAPTIVE, 0.01, 20)
    out[f"rnd{n}_eu_pm"], out[f"rnd{n}_eu_v"], out[f"rnd{n}_eu_E"] = p2, v2, ref.energy(p2, v2)

# Close pair below the cut-off (r^2 < 1e-10 must be skipped) and an exact duplicate
pmc = np.array([[0, 0, 0, 1.0], [3e-6, 0, 0, 2.0], [1, 1, 0, 0.5], [1, 1, 0, 0.25], [0, 2, 1, 1e-3]], dtype=np.float64)
out["cut_pm"] = pmc
out["cut_pairwise"] = ref.forces(pmc, nbtest.PAIRWISE)
out["cut_adaptive"] = ref.forces(pmc, nbtest.ADAPTIVE)

# Galaxy with planets (huge coordinates, tiny separations), N = 1 + 40*3
pmg, vg = nbtest.gen_galaxy(40, planets=2, seed=3)
out["gal_pm"], out["gal_v"] = pmg, vg
out["gal_pairwise"] = ref.forces(pmg, nbtest.PAIRWISE)
out["gal_adaptive"] = ref.forces(pmg, nbtest.ADAPTIVE)
out["gal_avg_mass"] = ref.average_mass(pmg)

np.savez_compressed(os.path.join(HERE, "ref_vectors.npz"), **out)
print("wrote", os.path.join(HERE, "ref_vectors.npz"), len(out), "arrays")
