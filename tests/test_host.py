"""The C++17 host drop-in (n-body-simulation_b200/host/): config.json surface, generators, BENCHMARK / FILE_CSV sinks, and the reference's own
System.cpp linked against n-body-simulation_b200/host/ref_shims.cpp. CPU tests need no GPU; the gpu-marked ones run the binaries on the box."""
import ctypes as C
import json
import os
import subprocess

import numpy as np
import pytest

import nbtest

PKG = os.path.join(nbtest.ROOT, "n-body-simulation_b200")
HOST_LIB = os.path.join(PKG, "libnbody_host.so")
HOST_BIN = os.path.join(PKG, "nbody_host")
REF_BIN = os.path.join(nbtest.ROOT, "oracle", "_ref", "ref_system_on_b200")

BASE = {
    "init": {
        "selected": "RANDOM",
        "FROM_FILE": {"filePath": os.path.join(nbtest.ROOT, "data_in", "sistema.dat"), "partialCount": 25},
        "RANDOM": {"nParticles": 600, "minMass": 1e-10, "maxMass": 1e-5, "L": 1e2, "maxVelocity": 1.0, "minDistance": 1e-3},
        "GALAXY": {"nStars": 100, "blackHoleMass": 4e6, "starMassMin": 0.1, "starMassMax": 10.0, "maxPosition": 330000,
                   "maxVelocity": 42.15, "addPlanets": True, "minPlanets": 2, "maxPlanets": 2, "planetMassMin": 1.7e-7,
                   "planetMassMax": 9.6e-4, "planetDistanceMax": 6.68, "planetVelocityMax": 6.32},
        "STELLAR_SYSTEM": {"nPlanets": 50, "starMass": 1.0, "planetMassMin": 1.7e-8, "planetMassMax": 9.6e-4,
                           "minPlanetDistance": 0.73, "maxPlanetDistance": 16.68, "addMoons": True, "minMoons": 1, "maxMoons": 3,
                           "moonMassMin": 5e-10, "moonMassMax": 5e-8, "moonDistanceMax": 0.00067, "moonVelocityMax": 0.211},
        "SPIRAL_GALAXY": {"nStars": 5000, "totalMass": 1e4, "nArms": 2, "galaxyRadius": 3.09398e3, "thickness": 6.18795e2,
                          "perturbation": 5.0},
    },
    "threads": 2, "years": 100.0, "dtYears": 4e-2, "output": {"dir": "", "file": "out.csv"},
    "integrationMethod": "VELOCITY_VERLET", "forceMethod": "PAIRWISE", "executionMode": "GPU", "outputMode": "BENCHMARK",
}


def cfg_text(**over):
    c = json.loads(json.dumps(BASE))
    for k, v in over.items():
        if k == "selected":
            c["init"]["selected"] = v
        else:
            c[k] = v
    # the reference's shipped file has comment lines after the closing brace (config.json:67-68)
    return json.dumps(c, indent=1) + "\n# trailing comment line, tolerated like the reference's `in >> j`\n"


@pytest.fixture(scope="module")
def host():
    l = C.CDLL(HOST_LIB)
    l.nbh_last_error.restype = C.c_char_p
    l.nbh_init_from_config.argtypes = [C.c_char_p, C.c_long, C.POINTER(C.c_long), nbtest._dp, nbtest._dp, C.c_char_p, C.c_long]
    l.nbh_parse_config.argtypes = [C.c_char_p, C.POINTER(C.c_long)]
    return l


def init_from(host, text, cap=400000):
    pm = np.zeros((cap, 4)); v = np.zeros((cap, 4)); n = C.c_long(0); meta = C.create_string_buffer(4096)
    rc = host.nbh_init_from_config(text.encode(), cap, C.byref(n), pm.ctypes.data_as(nbtest._dp), v.ctypes.data_as(nbtest._dp),
                                   meta, 4096)
    if rc != 0:
        raise RuntimeError(host.nbh_last_error().decode())
    return pm[:n.value].copy(), v[:n.value].copy(), meta.value.decode()


def test_config_surface(host):
    out = (C.c_long * 8)()
    assert host.nbh_parse_config(cfg_text().encode(), out) == 0, host.nbh_last_error()
    assert list(out) == [2500, 5, 1, 0, 1, 0, 2, 0]           # steps = years/dt (main.cpp:496), stepFreq = max(1, n/500)
    assert host.nbh_parse_config(cfg_text(forceMethod="ADAPTIVE_MUTUAL", integrationMethod="EULER", outputMode="FILE_CSV",
                                          executionMode="CPU", seed=7).encode(), out) == 0
    assert list(out)[2:6] == [0, 2, 0, 1] and out[7] == 1
    for key, bad, msg in (("forceMethod", "FAST", "Unknown force method: FAST"),
                          ("integrationMethod", "RK4", "Unknown integration method: RK4"),
                          ("executionMode", "TPU", "Unknown execution mode: TPU"),
                          ("outputMode", "PNG", "Unknown output mode: PNG")):
        assert host.nbh_parse_config(cfg_text(**{key: bad}).encode(), out) == 1
        assert host.nbh_last_error().decode() == msg
    assert host.nbh_parse_config(b'{"init": {"selected": "RANDOM"}}', out) == 1
    assert "missing key" in host.nbh_last_error().decode()
    assert host.nbh_parse_config(b'{"init": ', out) == 1


def test_generators_are_seeded_and_shaped(host):
    a = init_from(host, cfg_text(seed=5))
    b = init_from(host, cfg_text(seed=5))
    c = init_from(host, cfg_text(seed=6))
    assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1]) and not np.array_equal(a[0], c[0])
    pm, v, meta = a
    assert pm.shape == (600, 4) and "# Number of particles: 600" in meta and "# Total particle count: 600" in meta
    assert np.abs(pm[:, :3]).max() <= 50 and np.abs(v[:, :3]).max() <= 1 and np.all(v[:, 3] == 0)
    assert pm[:, 3].min() >= 1e-10 and pm[:, 3].max() <= 1e-5
    d = np.linalg.norm(pm[:, None, :3] - pm[None, :, :3], axis=2) + np.eye(600) * 1e9
    assert d.min() >= 1e-3                                      # minDistance honoured

    pm, v, meta = init_from(host, cfg_text(selected="GALAXY", seed=1))
    assert pm.shape[0] == 1 + 100 * 3 and pm[0, 3] == 4e6 and np.all(pm[0, :3] == 0)   # min = max = 2 planets per star
    stars, p1 = pm[1::3], pm[2::3]
    assert np.abs(p1[:, :3] - stars[:, :3]).max() <= 6.68 and p1[:, 3].max() <= 9.6e-4 and stars[:, 3].min() >= 0.1

    pm, v, meta = init_from(host, cfg_text(selected="SPIRAL_GALAXY", seed=1))
    assert pm.shape[0] == 5000 and np.allclose(pm[:, 3], 1e4 / 5000)
    assert np.linalg.norm(pm[:, :3], axis=1).max() <= np.hypot(3.09398e3, 6.18795e2 / 2) * 1.0001
    assert "# Initializer: Spiral Galaxy" in meta

    pm, v, meta = init_from(host, cfg_text(selected="STELLAR_SYSTEM", seed=1))
    assert pm[0, 3] == 1.0 and np.all(pm[:, 2] == 0)            # everything in the XY plane
    assert 1 + 50 * 2 <= pm.shape[0] <= 1 + 50 * 4
    # config passes moonDistanceMax / moonVelocityMax into the (min, max) moon-distance slots (ConfigParser.hpp:146-152)
    assert "# moonDistance range: [0.00067, 0.211] AU" in meta

    pm, v, meta = init_from(host, cfg_text(selected="FROM_FILE"))
    ref_pm, ref_v = nbtest.load_sistema()
    np.testing.assert_allclose(pm, ref_pm, rtol=1e-15); np.testing.assert_allclose(v, ref_v, rtol=1e-15)
    assert "# Total in file: 25" in meta


def test_binary_snapshot_roundtrip(host, tmp_path):
    """Checkpoint/resume: a .nbin snapshot written by the host is read back bit-exactly through FROM_FILE."""
    pm, v, _ = init_from(host, cfg_text(selected="GALAXY", seed=4))
    path = str(tmp_path / "state.nbin")
    host.nbh_save_snapshot.argtypes = [C.c_char_p, C.c_long, nbtest._dp, nbtest._dp]
    assert host.nbh_save_snapshot(path.encode(), pm.shape[0], pm.ctypes.data_as(nbtest._dp), v.ctypes.data_as(nbtest._dp)) == 0
    assert os.path.getsize(path) == 16 + pm.shape[0] * 64
    text = cfg_text(selected="FROM_FILE").replace(json.dumps(os.path.join(nbtest.ROOT, "data_in", "sistema.dat")), json.dumps(path))
    pm2, v2, meta = init_from(host, text.replace('"partialCount": 25', '"partialCount": 0'))
    assert np.array_equal(pm, pm2) and np.array_equal(v, v2) and "binary snapshot" in meta
    pm3, v3, _ = init_from(host, text.replace('"partialCount": 25', '"partialCount": 10'))
    assert np.array_equal(pm[:10], pm3) and np.array_equal(v[:10], v3)
    (tmp_path / "bad.nbin").write_bytes(b"garbage")
    with pytest.raises(RuntimeError, match="NBIN0001"):
        init_from(host, text.replace(json.dumps(path), json.dumps(str(tmp_path / "bad.nbin"))))


def test_host_generators_match_reference_distributions(host, ref):
    pm, v, _ = init_from(host, cfg_text(selected="SPIRAL_GALAXY", seed=3))
    rp, rv = ref.init_spiral(5000)
    for a, b in ((pm, rp), (v, rv)):
        ra, rb = np.linalg.norm(a[:, :3], axis=1), np.linalg.norm(b[:, :3], axis=1)
        assert abs(ra.mean() - rb.mean()) < 0.05 * rb.mean() and abs(ra.std() - rb.std()) < 0.08 * rb.std()
    pm, v, _ = init_from(host, cfg_text(seed=3))
    rp, rv = ref.init_random(600)
    assert abs(pm[:, 3].mean() - rp[:, 3].mean()) < 0.25 * rp[:, 3].mean()      # unseeded reference draw: sigma ~2.4 %
    assert abs(np.abs(v[:, :3]).mean() - np.abs(rv[:, :3]).mean()) < 0.1


def test_generator_metadata_text_matches_the_reference(host, ref):
    """The "# ..." metadata block each generator contributes to the CSV (core/Initializer.cpp) — compared as text with what
    the reference's own generators emit for the same parameters (particle values differ: the reference is unseeded)."""
    ref.l.nbref_last_metadata.restype = C.c_long
    ref.l.nbref_last_metadata.argtypes = [C.c_char_p, C.c_long]

    def ref_meta():
        buf = C.create_string_buffer(8192)
        ref.l.nbref_last_metadata(buf, 8192)
        return buf.value.decode()

    ref.init_random(600)
    assert init_from(host, cfg_text(seed=1))[2] == ref_meta()
    ref.init_spiral(5000)
    import re
    mask = lambda t: re.sub(r"(normal\): )\([^)]*\)", r"\1(...)", t)       # the random disc orientation is printed: mask its value
    ours, theirs = init_from(host, cfg_text(selected="SPIRAL_GALAXY", seed=1))[2], ref_meta()
    assert "(...)" in mask(ours) and mask(ours) == mask(theirs)
    ref.init_file(BASE["init"]["FROM_FILE"]["filePath"], 25)
    assert init_from(host, cfg_text(selected="FROM_FILE"))[2] == ref_meta()
    g = BASE["init"]["GALAXY"]
    pm = np.zeros((400, 4)); v = np.zeros((400, 4)); k = C.c_long(0)
    ref.l.nbref_init_galaxy.argtypes = [C.c_int] + [C.c_double] * 5 + [C.c_int] * 3 + [C.c_double] * 4 + [C.c_long, C.POINTER(C.c_long), nbtest._dp, nbtest._dp]
    assert ref.l.nbref_init_galaxy(g["nStars"], g["blackHoleMass"], g["starMassMin"], g["starMassMax"], g["maxPosition"], g["maxVelocity"],
                                   1, g["minPlanets"], g["maxPlanets"], g["planetMassMin"], g["planetMassMax"], g["planetDistanceMax"],
                                   g["planetVelocityMax"], 400, C.byref(k), nbtest._p(pm), nbtest._p(v)) == 0
    assert init_from(host, cfg_text(selected="GALAXY", seed=1))[2] == ref_meta()
    # STELLAR_SYSTEM, with the reference's config quirk: moonDistanceMax / moonVelocityMax land in the (moonDistanceMin,
    # moonDistanceMax) slots (core/ConfigParser.hpp:146-152)
    q = BASE["init"]["STELLAR_SYSTEM"]
    ref.l.nbref_init_stellar.argtypes = [C.c_double, C.c_int] + [C.c_double] * 4 + [C.c_int] * 3 + [C.c_double] * 4 + \
                                        [C.c_long, C.POINTER(C.c_long), nbtest._dp, nbtest._dp]
    assert ref.l.nbref_init_stellar(q["starMass"], q["nPlanets"], q["planetMassMin"], q["planetMassMax"], q["minPlanetDistance"],
                                    q["maxPlanetDistance"], 1, q["minMoons"], q["maxMoons"], q["moonMassMin"], q["moonMassMax"],
                                    q["moonDistanceMax"], q["moonVelocityMax"], 400, C.byref(k), nbtest._p(pm), nbtest._p(v)) == 0
    mask_n = lambda t: re.sub(r"(Total particle count: )\d+", r"\1N", t)      # 1..3 moons per planet: the count is random
    hp, hv, hmeta = init_from(host, cfg_text(selected="STELLAR_SYSTEM", seed=1))
    assert mask_n(hmeta) == mask_n(ref_meta())
    # same construction on both sides: the star at rest at the origin, planets on circular orbits in the z = 0 plane
    for P, V, cnt in ((hp, hv, hp.shape[0]), (pm, v, k.value)):
        assert 1 + 50 * 2 <= cnt <= 1 + 50 * 4 and P[0, 3] == 1.0 and not P[0, :3].any() and not V[0, :3].any()
        assert not P[:cnt, 2].any() and not V[:cnt, 2].any()
        r = np.linalg.norm(P[1:cnt, :2], axis=1); sp = np.linalg.norm(V[1:cnt, :2], axis=1)
        circular = np.isclose(sp, np.sqrt(39.47841760435743 * 1.0 / r), rtol=1e-12, atol=0)   # planets: v = sqrt(G M / r) exactly
        assert circular.sum() == 50
        r = r[circular]
        assert r.min() >= q["minPlanetDistance"] and r.max() <= q["maxPlanetDistance"]


def test_csv_sink_is_byte_identical_to_the_reference_writer(host, ref, tmp_path):
    """FILE_CSV text (metadata, header, fixed 8-digit fields, 15-digit energy): the host's writer against the reference's own
    flushCSVOutput (output/OutputUtils.cpp:8-59, compiled unmodified into oracle/_ref) on the same OutputData contents."""
    rng = np.random.default_rng(12)
    n, T = 37, 5
    pd = rng.normal(0, 1, (T, n, 6)) * 10.0 ** rng.integers(-9, 7, (T, n, 6))        # tiny, huge, negative values
    pd[0, 0] = [0.0, -0.0, 1e-9, -4.9999999949e-9, 123456789.123456789, -1.0]         # rounding edges of %.8f
    pd[0, 1] = [np.inf, -np.inf, np.nan, 5e-324, 1.7976931348623157e308, -2.5e-8]     # non-finite, denormal, largest, tie
    pd[1] *= 10.0 ** rng.integers(-300, 290, (n, 6)).astype(float)                     # the whole double range in one snapshot
    sd = np.stack([np.arange(T) * 0.04, rng.normal(-4e-3, 1e-3, T)], axis=1)
    st = np.stack([np.arange(n, dtype=float), 10.0 ** rng.uniform(-10, 6, n)], axis=1)
    meta = "# years: 100.000000\n# Random particles\n# dt (years): 0.040000\n"
    args = [C.c_char_p, C.c_char_p, C.c_long, C.c_long, nbtest._dp, nbtest._dp, nbtest._dp]
    host.nbh_write_csv.argtypes = args
    ref.l.nbref_write_csv.argtypes = args
    a, b = tmp_path / "host.csv", tmp_path / "ref.csv"
    pdc, sdc, stc = (np.ascontiguousarray(x) for x in (pd, sd, st))
    assert host.nbh_write_csv(str(a).encode(), meta.encode(), n, T, nbtest._p(pdc), nbtest._p(sdc), nbtest._p(stc)) == 0
    assert ref.l.nbref_write_csv(str(b).encode(), meta.encode(), n, T, nbtest._p(pdc), nbtest._p(sdc), nbtest._p(stc)) == 0
    ta, tb = a.read_bytes(), b.read_bytes()
    assert len(ta.splitlines()) == 3 + 1 + 1 + n * T
    assert ta == tb


def test_device_init_key_is_validated_before_any_device_is_touched(tmp_path):
    """"deviceInit": true is for SPIRAL_GALAXY + BENCHMARK + a seed; anything else is a configuration error (no GPU needed to say so)."""
    for over in (dict(selected="RANDOM", seed=1), dict(selected="SPIRAL_GALAXY"), dict(selected="SPIRAL_GALAXY", seed=1, outputMode="FILE_CSV")):
        c = json.loads(cfg_text(**over).split("\n#")[0])
        c["deviceInit"] = True
        (tmp_path / "config.json").write_text(json.dumps(c))
        r = subprocess.run([HOST_BIN], cwd=tmp_path, capture_output=True, text=True, timeout=120)
        assert r.returncode == 1 and "deviceInit" in r.stderr, r.stderr


def test_host_binary_fails_loudly_without_gpu(tmp_path, lib):
    if lib.l.nb_device_count() > 0:
        pytest.skip("a GPU is present")
    (tmp_path / "config.json").write_text(cfg_text(seed=1))
    r = subprocess.run([HOST_BIN], cwd=tmp_path, capture_output=True, text=True)
    assert r.returncode == 1 and "no CUDA device" in r.stderr
    (tmp_path / "config.json").write_text(cfg_text(executionMode="CPU"))
    r = subprocess.run([HOST_BIN], cwd=tmp_path, capture_output=True, text=True)
    assert r.returncode == 1 and "no CPU fallback" in r.stderr
    (tmp_path / "config.json").write_text(cfg_text(outputMode="VISUALIZATION"))
    r = subprocess.run([HOST_BIN], cwd=tmp_path, capture_output=True, text=True)
    assert r.returncode == 1 and "reference-only" in r.stderr
    out = tmp_path / "s.csv"
    r = subprocess.run([HOST_BIN, "--auto-benchmark", "--sizes", "256", "--runs", "1", "--out", str(out)], capture_output=True, text=True)
    assert r.returncode == 2 and 'false,"no CUDA device available (there is no CPU fallback)"' in out.read_text()


SWEEP_HEADER = ("NumParticles,Years,DtYears,ExecutionMode,ForceMethod,IntegrationMethod,NumThreads,BlockSize,InitType,"
                "NumRuns,TotalSteps,MeanExecutionTime,StdExecutionTime,MeanStepsPerSecond,StdStepsPerSecond,CVStepsPerSecond,"
                "MinStepsPerSecond,MaxStepsPerSecond,MeanParticleStepsPerSecond,StdParticleStepsPerSecond,Success,ErrorMessage")


def test_auto_benchmark_statistics_and_csv_schema(host):
    """The --auto-benchmark harness (reference core/main.cpp:52-80 statistics, :238-288 CSV, :300-304 adaptive years)."""
    host.nbh_sample_stats.argtypes = [nbtest._dp, C.c_long, nbtest._dp]
    host.nbh_sweep_years_for.restype = C.c_double
    host.nbh_sweep_csv_from_times.argtypes = [C.c_int, C.c_double, C.c_double, C.c_int, C.c_int, nbtest._dp, C.c_long, C.c_char_p, C.c_long]
    v = np.array([2.0, 4.0, 4.0, 4.0, 5.0, 5.0, 7.0, 9.0]); out = np.zeros(5)
    assert host.nbh_sample_stats(v.ctypes.data_as(nbtest._dp), len(v), out.ctypes.data_as(nbtest._dp)) == 0
    assert out[0] == v.mean() and abs(out[1] - v.std(ddof=1)) < 1e-15 and out[2] == 2.0 and out[3] == 9.0
    assert abs(out[4] - v.std(ddof=1) / v.mean()) < 1e-15
    one = np.array([3.5])
    host.nbh_sample_stats(one.ctypes.data_as(nbtest._dp), 1, out.ctypes.data_as(nbtest._dp))
    assert list(out) == [3.5, 0.0, 3.5, 3.5, 0.0]                       # a single run has no spread (main.cpp:72-75)
    assert [host.nbh_sweep_years_for(n) for n in (64, 999, 1000, 9999, 10000, 131072)] == [1.5, 1.5, 0.25, 0.25, 0.1, 0.1]
    secs = np.array([0.5, 0.25, 1.0]); buf = C.create_string_buffer(4096)
    assert host.nbh_sweep_csv_from_times(4096, 0.25, 0.01, 1, 1, secs.ctypes.data_as(nbtest._dp), 3, buf, 4096) == 0
    head, row = buf.value.decode().strip().split("\n")
    assert head == SWEEP_HEADER
    f = row.split(",")
    assert len(f) == 22 and f[:11] == ["4096", "0.25", "0.01", "GPU", "PAIRWISE_AVX2/FP32", "VELOCITY_VERLET", "1", "-1", "random", "3", "25"]
    sps = 25 / secs
    assert f[11] == "%.6f" % secs.mean() and f[12] == "%.6f" % secs.std(ddof=1)
    assert f[13] == "%.2f" % sps.mean() and f[14] == "%.2f" % sps.std(ddof=1) and f[15] == "%.4f" % (sps.std(ddof=1) / sps.mean())
    assert f[16] == "%.2f" % sps.min() and f[17] == "%.2f" % sps.max()
    assert f[18] == "%.3e" % (4096 * sps).mean() and f[19] == "%.3e" % (4096 * sps).std(ddof=1)
    assert f[20] == "true" and f[21] == '""'


def test_auto_benchmark_cli_rejects_bad_options():
    r = subprocess.run([HOST_BIN, "--help"], capture_output=True, text=True)
    assert r.returncode == 0 and "--auto-benchmark" in r.stdout and "config.json" in r.stdout
    r = subprocess.run([HOST_BIN, "--auto-benchmark", "--forces", "LEAPFROG"], capture_output=True, text=True)
    assert r.returncode == 1 and "unknown force method" in r.stderr
    r = subprocess.run([HOST_BIN, "--auto-benchmark", "--bogus"], capture_output=True, text=True)
    assert r.returncode == 1 and "unknown option" in r.stderr


@pytest.mark.gpu
def test_host_benchmark_mode(tmp_path):
    (tmp_path / "config.json").write_text(cfg_text(seed=1, years=0.4, dtYears=0.04))
    r = subprocess.run([HOST_BIN], cwd=tmp_path, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    assert "Simulation time:" in r.stdout and "(using GPU)" in r.stdout and "Force method: Pairwise" in r.stdout
    (tmp_path / "config.json").write_text(cfg_text(seed=1, forceMethod="BARNES_HUT"))
    r = subprocess.run([HOST_BIN], cwd=tmp_path, capture_output=True, text=True, timeout=300)
    assert r.returncode == 1 and "reference-only" in r.stderr


@pytest.mark.gpu
def test_auto_benchmark_sweep_on_gpu(tmp_path):
    out = tmp_path / "sweep.csv"
    r = subprocess.run([HOST_BIN, "--auto-benchmark", "--sizes", "256,4096", "--forces", "PAIRWISE,ADAPTIVE_MUTUAL", "--runs", "3",
                        "--years", "0.05", "--seed", "11", "--out", str(out)], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout + r.stderr
    lines = out.read_text().strip().split("\n")
    assert lines[0] == SWEEP_HEADER and len(lines) == 5
    for ln, (n, fm) in zip(lines[1:], [(256, "PAIRWISE"), (256, "ADAPTIVE_MUTUAL"), (4096, "PAIRWISE"), (4096, "ADAPTIVE_MUTUAL")]):
        f = ln.split(",")
        assert f[0] == str(n) and f[3] == "GPU" and f[4] == fm and f[9] == "3" and f[10] == "5" and f[20] == "true"
        assert float(f[13]) > 0 and float(f[16]) <= float(f[13]) <= float(f[17]) and float(f[18]) > 0
    # BARNES_HUT stays reference-only: the row records the failure (Success=false + message) and the exit code is 2
    r = subprocess.run([HOST_BIN, "--auto-benchmark", "--sizes", "256", "--forces", "BARNES_HUT", "--runs", "1", "--out", str(out)],
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 2
    f = out.read_text().strip().split("\n")[1].split(",")
    assert f[20] == "false" and "reference-only" in f[21]


@pytest.mark.gpu
def test_host_visualization_mode_writes_headless_frames(tmp_path, host):
    """outputMode VISUALIZATION + a "frames" block: the reference's render loop (VisualizationKernels.cu:52-300) without a
    window. 12 steps, one step per frame, every 4th frame written -> 12 rendered, 3 PPM files; the state advances exactly
    as in BENCHMARK mode."""
    frames = {"dir": "frames", "width": 320, "height": 180, "physicsStepsPerFrame": 1, "every": 4}
    text = cfg_text(seed=3, years=0.5, dtYears=0.04, outputMode="VISUALIZATION", frames=frames, snapshotOut="viz.nbin",
                    precision="FP64")
    (tmp_path / "config.json").write_text(text)
    r = subprocess.run([HOST_BIN], cwd=tmp_path, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    assert "Frames: 12 rendered, 3 written" in r.stdout
    files = sorted(os.listdir(tmp_path / "frames"))
    assert files == ["frame_000000.ppm", "frame_000001.ppm", "frame_000002.ppm"]
    imgs = []
    for f in files:
        raw = (tmp_path / "frames" / f).read_bytes()
        assert raw.startswith(b"P6\n320 180\n255\n") and len(raw) == len(b"P6\n320 180\n255\n") + 320 * 180 * 3
        imgs.append(np.frombuffer(raw[-320 * 180 * 3:], dtype=np.uint8).reshape(180, 320, 3))
    background = np.array([5, 5, 13], dtype=np.uint8)                    # clear colour (0.02, 0.02, 0.05) quantised
    lit = [(im != background).any(axis=2).sum() for im in imgs]
    assert min(lit) > 600                                               # 600 particles with 30 px sprites: thousands of pixels
    assert not np.array_equal(imgs[0], imgs[2])                         # the particles move between frames
    (tmp_path / "config.json").write_text(cfg_text(seed=3, years=0.5, dtYears=0.04, snapshotOut="bench.nbin", precision="FP64"))
    r = subprocess.run([HOST_BIN], cwd=tmp_path, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    assert (tmp_path / "viz.nbin").read_bytes() == (tmp_path / "bench.nbin").read_bytes()
    # without the "frames" block the mode stays a clean error
    (tmp_path / "config.json").write_text(cfg_text(seed=3, outputMode="VISUALIZATION"))
    r = subprocess.run([HOST_BIN], cwd=tmp_path, capture_output=True, text=True, timeout=300)
    assert r.returncode == 1 and "reference-only" in r.stderr and "frames" in r.stderr


@pytest.mark.gpu
def test_host_file_csv_mode(tmp_path, host, oracle):
    text = cfg_text(seed=2, years=1.0, dtYears=0.01, outputMode="FILE_CSV", forceMethod="ADAPTIVE_MUTUAL")
    text = text.replace('"nParticles": 600', '"nParticles": 40')
    (tmp_path / "config.json").write_text(text)
    r = subprocess.run([HOST_BIN], cwd=tmp_path, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    lines = (tmp_path / "out.csv").read_text().splitlines()
    hdr = lines.index("particle_id,time,x,y,z,vx,vy,vz,mass,energy")
    assert lines[0].startswith("# years: 1.0") and any(l == "# force method: Adaptive Mutual" for l in lines[:hdr])
    rows = [l.split(",") for l in lines[hdr + 1:]]
    assert len(rows) == 101 * 40                                # steps=100, stepFreq=1 -> 101 snapshots
    pm, v, _ = init_from(host, text)
    first = np.array(rows[:40], dtype=float)
    np.testing.assert_allclose(first[:, 2:5], pm[:, :3], atol=6e-9)
    np.testing.assert_allclose(first[:, 8], pm[:, 3], atol=6e-9)
    assert first[0, 0] == 0 and first[39, 0] == 39 and float(rows[40][1]) == pytest.approx(0.01)
    assert float(rows[0][9]) == pytest.approx(oracle.energy(pm, v), rel=1e-9)
    p_end, v_end = oracle.step(pm, v, nbtest.VV, nbtest.ADAPTIVE, 0.01, 100)
    last = np.array(rows[-40:], dtype=float)
    np.testing.assert_allclose(last[:, 2:5], p_end[:, :3], atol=2e-7)


@pytest.mark.gpu
def test_host_checkpoint_resume(tmp_path, host, oracle):
    """"snapshotOut" + FROM_FILE *.nbin: 10 steps, checkpoint, 10 more steps == 20 steps of the CPU integrator."""
    snap1, snap2 = str(tmp_path / "a.nbin"), str(tmp_path / "b.nbin")
    base = json.loads(cfg_text(seed=11, years=0.1, dtYears=0.01, precision="FP64").split("\n#")[0])
    base["init"]["RANDOM"]["nParticles"] = 200
    base["snapshotOut"] = snap1
    (tmp_path / "config.json").write_text(json.dumps(base))
    assert subprocess.run([HOST_BIN], cwd=tmp_path, capture_output=True, text=True, timeout=300).returncode == 0
    base["init"]["selected"] = "FROM_FILE"
    base["init"]["FROM_FILE"] = {"filePath": snap1, "partialCount": 0}
    base["snapshotOut"] = snap2
    (tmp_path / "config.json").write_text(json.dumps(base))
    assert subprocess.run([HOST_BIN], cwd=tmp_path, capture_output=True, text=True, timeout=300).returncode == 0
    raw = np.fromfile(snap2, dtype=np.float64, offset=16)
    pm2, v2 = raw[:800].reshape(200, 4), raw[800:].reshape(200, 4)
    start = json.loads(cfg_text(seed=11).split("\n#")[0]); start["init"]["RANDOM"]["nParticles"] = 200
    pm0, v0, _ = init_from(host, json.dumps(start))
    p_ref, v_ref = oracle.step(pm0, v0, nbtest.VV, nbtest.PAIRWISE, 0.01, 20)
    np.testing.assert_allclose(pm2[:, :3], p_ref[:, :3], atol=1e-9)
    np.testing.assert_allclose(v2[:, :3], v_ref[:, :3], atol=1e-9)


@pytest.mark.gpu
def test_host_device_side_initial_conditions(tmp_path, oracle):
    """"deviceInit": true — SPIRAL_GALAXY made on the GPU (nb_init_spiral), BENCHMARK run, final state in the snapshot file:
    equals the CPU integrator started from the numpy restatement of the same counter-based draws."""
    snap = str(tmp_path / "end.nbin")
    base = json.loads(cfg_text(selected="SPIRAL_GALAXY", seed=77, years=0.08, dtYears=0.02, precision="FP64").split("\n#")[0])
    sp = base["init"]["SPIRAL_GALAXY"]
    sp["nStars"] = 1500
    base["deviceInit"] = True
    base["snapshotOut"] = snap
    (tmp_path / "config.json").write_text(json.dumps(base))
    r = subprocess.run([HOST_BIN], cwd=tmp_path, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "on the device" in r.stdout
    raw = np.fromfile(snap, dtype=np.float64, offset=16)
    pm1, v1 = raw[:6000].reshape(1500, 4), raw[6000:].reshape(1500, 4)
    pm0, v0, _ = nbtest.gen_spiral_counter(1500, 77, sp["totalMass"], sp["nArms"], sp["galaxyRadius"], sp["thickness"], sp["perturbation"])
    p_ref, v_ref = oracle.step(pm0, v0, nbtest.VV, nbtest.PAIRWISE, 0.02, 4)
    np.testing.assert_allclose(pm1[:, :3], p_ref[:, :3], atol=1e-8)
    np.testing.assert_allclose(v1[:, :3], v_ref[:, :3], atol=1e-8)
    base["seed"] = None
    del base["seed"]
    (tmp_path / "config.json").write_text(json.dumps(base))
    r = subprocess.run([HOST_BIN], cwd=tmp_path, capture_output=True, text=True, timeout=300)
    assert r.returncode != 0 and "seed" in (r.stdout + r.stderr)


@pytest.mark.gpu
@pytest.mark.parametrize("prec,tol", [("0", 2e-5), ("1", 1e-9)], ids=["fp32", "fp64"])
def test_reference_system_cpp_runs_on_the_b200_library(prec, tol):
    """The reference's own System.cpp/Forces.cpp, linked with n-body-simulation_b200/host/ref_shims.cpp: System(GPU).runBenchmark goes through
    runBenchmarkOnGPU -> nb_run_benchmark and lands on the same state as the reference's CPU/OpenMP run."""
    if not os.path.exists(REF_BIN):
        pytest.skip("oracle/_ref/ref_system_on_b200 not built (reference tree absent at build time)")
    env = dict(os.environ, NB_PRECISION=prec, OMP_NUM_THREADS="8")
    r = subprocess.run([REF_BIN, "1500", "10", "0.01", "0", "1"], capture_output=True, text=True, env=env, timeout=300)
    assert r.returncode == 0, r.stderr
    d = json.loads(r.stdout.strip().splitlines()[-1])
    assert d["n"] == 1500 and d["max_abs_dpos"] <= tol * 50 and d["max_abs_dvel"] <= tol * 50
    assert abs(d["E_gpu"] - d["E_cpu"]) <= 1e-6 * abs(d["E_cpu"])
