"""Test helpers: ctypes bindings for the product C ABI, the C oracle and (optionally) the compiled reference,
plus seeded numpy generators with the distributions of the reference's Initializer (core/Initializer.cpp)."""
import ctypes as C
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB_PATH = os.environ.get("NB_LIB_PATH", os.path.join(ROOT, "n-body-simulation_b200", "libnbody_b200.so"))
ORACLE_PATH = os.path.join(ROOT, "oracle", "libnboracle.so")
REF_PATH = os.path.join(ROOT, "oracle", "_ref", "libnbref.so")

EULER, VV = 0, 1
PAIRWISE, AVX2, ADAPTIVE, BARNES_HUT = 0, 1, 2, 3
FP32, FP64 = 0, 1
G = 39.47841760435743

_dp = C.POINTER(C.c_double)


def _p(a):
    return a.ctypes.data_as(_dp)


def as_state(pos_mass, vel):
    pm = np.ascontiguousarray(pos_mass, dtype=np.float64).reshape(-1, 4).copy()
    v = np.ascontiguousarray(vel, dtype=np.float64).reshape(-1, 4).copy()
    return pm, v


class Oracle:
    """oracle/libnboracle.so — the plain-C restatement (the checker)."""

    def __init__(self):
        self.l = C.CDLL(ORACLE_PATH)
        self.l.nbo_average_mass.restype = C.c_double
        self.l.nbo_energy.restype = C.c_double
        self.l.nbo_average_mass.argtypes = [C.c_long, _dp]
        self.l.nbo_forces.argtypes = [C.c_long, _dp, C.c_int, _dp]
        self.l.nbo_forces_subset.argtypes = [C.c_long, _dp, C.c_int, C.c_long, C.POINTER(C.c_long), _dp]
        self.l.nbo_step.argtypes = [C.c_long, _dp, _dp, C.c_int, C.c_int, C.c_double, C.c_int]
        self.l.nbo_energy.argtypes = [C.c_long, _dp, _dp]
        self.l.nbo_momentum.argtypes = [C.c_long, _dp, _dp, _dp]
        self.l.nbo_com_velocity.argtypes = [C.c_long, _dp, _dp, _dp]
        self.l.nbo_set_threads.argtypes = [C.c_int]

    def set_threads(self, t):
        self.l.nbo_set_threads(int(t))

    def max_threads(self):
        return int(self.l.nbo_max_threads())

    def forces(self, pm, method):
        n = pm.shape[0]
        out = np.zeros((n, 3))
        assert self.l.nbo_forces(n, _p(pm), method, _p(out)) == 0
        return out

    def forces_subset(self, pm, method, idx):
        idx = np.ascontiguousarray(idx, dtype=np.int64)
        out = np.zeros((len(idx), 3))
        assert self.l.nbo_forces_subset(pm.shape[0], _p(pm), method, len(idx),
                                        idx.ctypes.data_as(C.POINTER(C.c_long)), _p(out)) == 0
        return out

    def step(self, pm, v, integ, method, dt, steps):
        pm, v = pm.copy(), v.copy()
        assert self.l.nbo_step(pm.shape[0], _p(pm), _p(v), integ, method, dt, steps) == 0
        return pm, v

    def energy(self, pm, v):
        return float(self.l.nbo_energy(pm.shape[0], _p(pm), _p(v)))

    def momentum(self, pm, v):
        out = np.zeros(3)
        self.l.nbo_momentum(pm.shape[0], _p(pm), _p(v), _p(out))
        return out

    def average_mass(self, pm):
        return float(self.l.nbo_average_mass(pm.shape[0], _p(pm)))


class Ref:
    """oracle/_ref/libnbref.so — the reference's own CPU sources compiled unmodified (optional)."""

    @staticmethod
    def try_load():
        if not os.path.exists(REF_PATH):
            return None
        try:
            return Ref()
        except OSError:
            return None

    def __init__(self):
        self.l = C.CDLL(REF_PATH)
        self.l.nbref_energy.restype = C.c_double
        self.l.nbref_average_mass.restype = C.c_double
        self.l.nbref_forces.argtypes = [C.c_long, _dp, C.c_int, _dp]
        self.l.nbref_run.argtypes = [C.c_long, _dp, _dp, C.c_int, C.c_int, C.c_double, C.c_int]
        self.l.nbref_energy.argtypes = [C.c_long, _dp, _dp]
        self.l.nbref_average_mass.argtypes = [C.c_long, _dp]
        self.l.nbref_set_threads.argtypes = [C.c_int]
        lp = C.POINTER(C.c_long)
        self.l.nbref_init_random.argtypes = [C.c_int] + [C.c_double] * 5 + [C.c_long, lp, _dp, _dp]
        self.l.nbref_init_spiral.argtypes = [C.c_int, C.c_double, C.c_int, C.c_double, C.c_double, C.c_double,
                                             C.c_long, lp, _dp, _dp]
        self.l.nbref_init_galaxy.argtypes = [C.c_int] + [C.c_double] * 5 + [C.c_int] * 3 + [C.c_double] * 4 + \
                                            [C.c_long, lp, _dp, _dp]
        self.l.nbref_init_file.argtypes = [C.c_char_p, C.c_int, C.c_long, lp, _dp, _dp]

    def set_threads(self, t):
        self.l.nbref_set_threads(int(t))

    def max_threads(self):
        return int(self.l.nbref_max_threads())

    def forces(self, pm, method):
        out = np.zeros((pm.shape[0], 3))
        assert self.l.nbref_forces(pm.shape[0], _p(pm), method, _p(out)) == 0
        return out

    def run(self, pm, v, integ, method, dt, steps):
        pm, v = pm.copy(), v.copy()
        assert self.l.nbref_run(pm.shape[0], _p(pm), _p(v), integ, method, dt, steps) == 0
        return pm, v

    def energy(self, pm, v):
        return float(self.l.nbref_energy(pm.shape[0], _p(pm), _p(v)))

    def average_mass(self, pm):
        return float(self.l.nbref_average_mass(pm.shape[0], _p(pm)))

    def _init(self, fn, cap, *args):
        pm = np.zeros((cap, 4)); v = np.zeros((cap, 4)); n = C.c_long(0)
        assert fn(*args, cap, C.byref(n), _p(pm), _p(v)) == 0
        return pm[:n.value].copy(), v[:n.value].copy()

    def init_random(self, n, min_mass=1e-10, max_mass=1e-5, L=100.0, max_vel=1.0, min_dist=1e-3):
        return self._init(self.l.nbref_init_random, n, n, min_mass, max_mass, L, max_vel, min_dist)

    def init_spiral(self, n, total_mass=1e4, arms=2, radius=3.09398e3, thickness=6.18795e2, perturbation=5.0):
        return self._init(self.l.nbref_init_spiral, n, n, total_mass, arms, radius, thickness, perturbation)

    def init_file(self, path, count):
        return self._init(self.l.nbref_init_file, count, path.encode(), count)


class NbError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"nb error {code}: {msg}")
        self.code = code


class Lib:
    """n-body-simulation_b200/libnbody_b200.so — the product, through its C ABI only."""

    def __init__(self):
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(f"{LIB_PATH} is missing: run `make lib` (there is no CPU fallback)")
        l = self.l = C.CDLL(LIB_PATH)
        vp = C.c_void_p
        l.nb_last_error.restype = C.c_char_p
        l.nb_launch_count.restype = C.c_long
        l.nb_create.argtypes = [C.POINTER(vp), C.c_long, C.c_int, C.c_int]
        l.nb_destroy.argtypes = [vp]
        l.nb_set_methods.argtypes = [vp, C.c_int, C.c_int, C.c_int]
        l.nb_partition_bounds.argtypes = [C.c_long, C.c_int, C.c_int, C.POINTER(C.c_long), C.POINTER(C.c_long)]
        l.nb_comm_unique_id.argtypes = [vp]
        l.nb_comm_init.argtypes = [vp, vp, C.c_int, C.c_int]
        l.nb_upload.argtypes = [vp, _dp, _dp]
        l.nb_download.argtypes = [vp, _dp, _dp]
        l.nb_compute_accel.argtypes = [vp]
        l.nb_get_accel.argtypes = [vp, _dp]
        l.nb_step.argtypes = [vp, C.c_double, C.c_int]
        l.nb_sync.argtypes = [vp]
        l.nb_energy.argtypes = [vp, _dp]
        l.nb_momentum.argtypes = [vp, _dp]
        l.nb_average_mass.argtypes = [vp, _dp]
        l.nb_snapshot.argtypes = [vp, C.c_double, _dp, _dp, _dp]
        l.nb_launch_count.argtypes = [vp]
        l.nb_last_force_ms.argtypes = [vp, C.POINTER(C.c_float)]
        l.nb_set_option.argtypes = [vp, C.c_char_p, C.c_long]
        l.nb_force_timing.argtypes = [vp, C.c_int, C.POINTER(C.c_float), C.POINTER(C.c_int)]
        l.nb_timer_start.argtypes = [vp]
        if os.environ.get("NB_LIB_PATH") is None or hasattr(l, "nb_download_slice"):      # an older build may be loaded for A/B timings
            l.nb_near_timing.argtypes = [vp, C.POINTER(C.c_float), C.POINTER(C.c_int)]
            l.nb_get_order.argtypes = [vp, C.POINTER(C.c_uint)]
            l.nb_download_slice.argtypes = [vp, _dp, _dp, C.POINTER(C.c_uint), C.c_long, C.POINTER(C.c_long)]
            l.nb_init_spiral.argtypes = [vp, C.c_ulonglong, C.c_double, C.c_int, C.c_double, C.c_double, C.c_double, _dp]
        l.nb_timer_stop.argtypes = [vp, C.POINTER(C.c_float)]
        l.nb_run_benchmark.argtypes = [C.c_long, _dp, _dp, C.c_int, C.c_int, C.c_double, C.c_int, C.c_int,
                                       C.c_int, C.c_int, C.c_int]
        l.nb_run_simulation.argtypes = [C.c_long, _dp, _dp, C.c_int, C.c_int, C.c_double, C.c_int, C.c_int,
                                        _dp, _dp, _dp, C.c_long, C.c_int, C.c_int, C.c_int]
        l.nb_launch_force.argtypes = [vp, C.c_int, vp, vp, C.c_double, C.c_long, vp]
        l.nb_run_visualization.argtypes = [C.c_long, _dp, _dp, C.c_int, C.c_int, C.c_double, C.c_int, C.c_int, C.c_int, C.c_int,
                                           C.c_float, C.c_char_p, C.c_int, vp, C.POINTER(C.c_long), C.c_int, C.c_int, C.c_int]

    def err(self):
        return self.l.nb_last_error().decode()

    def check(self, rc):
        if rc != 0:
            raise NbError(rc, self.err())

    def partition_bounds(self, n, rank, world):
        b, c = C.c_long(0), C.c_long(0)
        self.check(self.l.nb_partition_bounds(n, rank, world, C.byref(b), C.byref(c)))
        return b.value, c.value

    def context(self, n, precision=FP32, device=0):
        return Context(self, n, precision, device)

    def run_benchmark(self, pm, v, integ, force, dt, steps, precision=FP32, subtract_com=1, device=0, block_size=256):
        pm, v = pm.copy(), v.copy()
        self.check(self.l.nb_run_benchmark(pm.shape[0], _p(pm), _p(v), integ, force, dt, steps, block_size,
                                           precision, subtract_com, device))
        return pm, v

    def run_simulation(self, pm, v, integ, force, dt, steps, step_freq, precision=FP64, subtract_com=1, device=0):
        pm, v = pm.copy(), v.copy()
        n = pm.shape[0]
        T = steps // step_freq + 1
        pdata = np.zeros((T, n, 6)); sdata = np.zeros((T, 2)); stat = np.zeros((n, 2))
        self.check(self.l.nb_run_simulation(n, _p(pm), _p(v), integ, force, dt, steps, step_freq,
                                            _p(pdata), _p(sdata), _p(stat), T, precision, subtract_com, device))
        return pm, v, pdata, sdata, stat


class Context:
    def __init__(self, lib, n, precision, device):
        self.lib, self.l, self.n = lib, lib.l, n
        self.h = C.c_void_p(None)
        lib.check(self.l.nb_create(C.byref(self.h), n, device, precision))

    def close(self):
        if self.h:
            self.l.nb_destroy(self.h)
            self.h = C.c_void_p(None)

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def set_methods(self, integ, force, subtract_com=1):
        self.lib.check(self.l.nb_set_methods(self.h, integ, force, subtract_com))

    def set_option(self, key, value):
        self.lib.check(self.l.nb_set_option(self.h, key.encode(), value))

    def upload(self, pm, v):
        self.lib.check(self.l.nb_upload(self.h, _p(pm), _p(v)))

    def download(self):
        pm = np.zeros((self.n, 4)); v = np.zeros((self.n, 4))
        self.lib.check(self.l.nb_download(self.h, _p(pm), _p(v)))
        return pm, v

    def download_slice(self, pm=None, v=None, index=None):
        """Own rows only (multi-GPU egress without the gather): returns views of length count."""
        cap = self.n if pm is None else pm.shape[0]
        pm = np.zeros((cap, 4)) if pm is None else pm
        v = np.zeros((cap, 4)) if v is None else v
        index = np.zeros(cap, dtype=np.uint32) if index is None else index
        cnt = C.c_long(0)
        self.lib.check(self.l.nb_download_slice(self.h, _p(pm), _p(v), index.ctypes.data_as(C.POINTER(C.c_uint)), cap, C.byref(cnt)))
        return pm[:cnt.value], v[:cnt.value], index[:cnt.value]

    def init_spiral(self, seed, total_mass=1e4, arms=2, radius=3.09398e3, thickness=6.18795e2, perturbation=5.0):
        nrm = np.zeros(3)
        self.lib.check(self.l.nb_init_spiral(self.h, seed, total_mass, arms, radius, thickness, perturbation, _p(nrm)))
        return nrm

    def compute_accel(self):
        self.lib.check(self.l.nb_compute_accel(self.h))

    def accel(self):
        a = np.zeros((self.n, 4))
        self.lib.check(self.l.nb_get_accel(self.h, _p(a)))
        return a

    def step(self, dt, nsteps):
        self.lib.check(self.l.nb_step(self.h, dt, nsteps))

    def sync(self):
        self.lib.check(self.l.nb_sync(self.h))

    def energy(self):
        e = np.zeros(3)
        self.lib.check(self.l.nb_energy(self.h, _p(e)))
        return e

    def momentum(self):
        p = np.zeros(4)
        self.lib.check(self.l.nb_momentum(self.h, _p(p)))
        return p

    def average_mass(self):
        m = C.c_double(0)
        self.lib.check(self.l.nb_average_mass(self.h, C.byref(m)))
        return m.value

    def last_force_ms(self):
        ms = C.c_float(0)
        self.lib.check(self.l.nb_last_force_ms(self.h, C.byref(ms)))
        return ms.value

    def force_timing(self, reset=True):
        ms, cnt = C.c_float(0), C.c_int(0)
        self.lib.check(self.l.nb_force_timing(self.h, 1 if reset else 0, C.byref(ms), C.byref(cnt)))
        return ms.value, cnt.value

    def near_timing(self):
        ms, cnt = C.c_float(0), C.c_int(0)
        self.lib.check(self.l.nb_near_timing(self.h, C.byref(ms), C.byref(cnt)))
        return ms.value, cnt.value

    def order(self):
        o = np.zeros(self.n, dtype=np.uint32)
        self.lib.check(self.l.nb_get_order(self.h, o.ctypes.data_as(C.POINTER(C.c_uint))))
        return o

    def timer_start(self):
        self.lib.check(self.l.nb_timer_start(self.h))

    def timer_stop(self):
        ms = C.c_float(0)
        self.lib.check(self.l.nb_timer_stop(self.h, C.byref(ms)))
        return ms.value

    def launches(self):
        return int(self.l.nb_launch_count(self.h))


# ---------------------------------------------------------------------------------------------------------------------
# Seeded generators with the reference's distributions (the reference seeds from std::random_device, so identical
# inputs need our own seed): core/Initializer.cpp:168-239 (RANDOM), :360-469 (SPIRAL_GALAXY), :71-154 (GALAXY).
def gen_random(n, seed=42, min_mass=1e-10, max_mass=1e-5, L=100.0, max_vel=1.0):
    r = np.random.default_rng(seed)
    pm = np.zeros((n, 4)); v = np.zeros((n, 4))
    pm[:, 3] = r.uniform(min_mass, max_mass, n)
    pm[:, :3] = r.uniform(-L / 2, L / 2, (n, 3))
    v[:, :3] = r.uniform(-max_vel, max_vel, (n, 3))
    return pm, v


def gen_spiral(n, seed=42, total_mass=1e4, arms=2, radius=3.09398e3, thickness=6.18795e2, perturbation=5.0):
    r = np.random.default_rng(seed)
    v0 = np.sqrt(G * total_mass / radius)
    th, ph = r.uniform(0, np.pi), r.uniform(0, 2 * np.pi)
    nrm = np.array([np.sin(th) * np.cos(ph), np.sin(th) * np.sin(ph), np.cos(th)])
    e1 = np.array([-nrm[1], nrm[0], 0.0]); e1 /= np.linalg.norm(e1)
    e2 = np.cross(nrm, e1)
    u = r.uniform(0, 1, n)
    rad = radius * np.sqrt(u)
    arm = np.arange(n) % arms
    phi = 2 * np.pi * arm / arms + rad / radius * 2 * np.pi + r.uniform(-np.pi / 8, np.pi / 8, n)
    xd, yd, zd = rad * np.cos(phi), rad * np.sin(phi), r.uniform(-thickness / 2, thickness / 2, n)
    pm = np.zeros((n, 4)); v = np.zeros((n, 4))
    pm[:, :3] = xd[:, None] * e1 + yd[:, None] * e2 + zd[:, None] * nrm
    pm[:, 3] = total_mass / n
    sp = v0 * np.sqrt(rad / radius)
    v[:, :3] = sp[:, None] * ((-np.sin(phi))[:, None] * e1 + np.cos(phi)[:, None] * e2) + \
        r.uniform(-perturbation, perturbation, (n, 3))
    return pm, v


def gen_galaxy(n_stars, planets=2, seed=42, bh_mass=4e6, m_min=0.1, m_max=10.0, max_pos=330000.0, max_vel=42.15,
               pm_min=1.7e-7, pm_max=9.6e-4, p_dist=6.68, p_vel=6.32):
    """Central black hole + stars, each with `planets` close companions: the input that breaks plain FP32."""
    r = np.random.default_rng(seed)
    n = 1 + n_stars * (1 + planets)
    pm = np.zeros((n, 4)); v = np.zeros((n, 4))
    pm[0, 3] = bh_mass
    k = 1
    for _ in range(n_stars):
        m = r.uniform(m_min, m_max); p = r.uniform(-max_pos, max_pos, 3); vel = r.uniform(-max_vel, max_vel, 3)
        pm[k, :3], pm[k, 3], v[k, :3] = p, m, vel
        k += 1
        for _ in range(planets):
            pm[k, :3] = p + r.uniform(-p_dist, p_dist, 3)
            pm[k, 3] = r.uniform(pm_min, pm_max)
            v[k, :3] = vel + r.uniform(-p_vel, p_vel, 3)
            k += 1
    return pm, v


def load_sistema(path=None, count=25):
    """core/Initializer.cpp:17-53: first token = body count, then mass[kg] x y z[km] vx vy vz[km/s] per body."""
    path = path or os.path.join(ROOT, "data_in", "sistema.dat")
    tok = open(path).read().split()
    vals = [float(t) for t in tok[1:1 + 7 * count]]
    a = np.array(vals).reshape(-1, 7)
    pm = np.zeros((len(a), 4)); v = np.zeros((len(a), 4))
    pm[:, 3] = a[:, 0] / 1.989e30
    pm[:, :3] = a[:, 1:4] / 1.496e8
    v[:, :3] = a[:, 4:7] * 0.2108
    return pm, v


def rel_err(a, b):
    """Per-particle ||a-b|| / ||b|| (rows); rows with ||b|| == 0 are compared absolutely."""
    a = np.asarray(a)[:, :3]; b = np.asarray(b)[:, :3]
    nb = np.linalg.norm(b, axis=1)
    d = np.linalg.norm(a - b, axis=1)
    return np.where(nb > 0, d / np.where(nb > 0, nb, 1.0), d)


def _mix64(z):
    z = (z + np.uint64(0x9E3779B97F4A7C15))
    z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
    z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
    return z ^ (z >> np.uint64(31))


def _u01(seed, i, k):
    with np.errstate(over="ignore"):
        s = _mix64(np.uint64(seed))
        h = _mix64(s ^ _mix64(np.asarray(i, dtype=np.uint64) * np.uint64(8) + np.uint64(k)))
    return (h >> np.uint64(11)).astype(np.float64) * (1.0 / 9007199254740992.0)


def gen_spiral_counter(n, seed, total_mass=1e4, arms=2, radius=3.09398e3, thickness=6.18795e2, perturbation=5.0):
    """numpy restatement of nb_init_spiral (csrc/nb_integrate.cu: counter-based draws, the reference's formulas of
    core/Initializer.cpp:360-469): the checker of the device-side generator."""
    sysid = np.uint64(0x7FFFFFFFFFFFFFFF)
    th, ph = float(_u01(seed, sysid, 0)) * np.pi, float(_u01(seed, sysid, 1)) * 2 * np.pi
    nrm = np.array([np.sin(th) * np.cos(ph), np.sin(th) * np.sin(ph), np.cos(th)])
    e1 = np.array([-nrm[1], nrm[0], 0.0]) / np.hypot(nrm[0], nrm[1])
    e2 = np.cross(nrm, e1)
    i = np.arange(n, dtype=np.uint64)
    r = radius * np.sqrt(_u01(seed, i, 0))
    phi = 2 * np.pi * (np.arange(n) % arms) / arms + r / radius * 2 * np.pi + (_u01(seed, i, 1) * 2 - 1) * (np.pi / 8)
    xd, yd, zd = r * np.cos(phi), r * np.sin(phi), (_u01(seed, i, 2) - 0.5) * thickness
    sp = np.sqrt(G * total_mass / radius) * np.sqrt(r / radius)
    pm = np.zeros((n, 4)); v = np.zeros((n, 4))
    pm[:, :3] = xd[:, None] * e1 + yd[:, None] * e2 + zd[:, None] * nrm
    pm[:, 3] = total_mass / n
    pert = np.stack([_u01(seed, i, 3 + c) * 2 - 1 for c in range(3)], axis=1) * perturbation
    v[:, :3] = sp[:, None] * ((-np.sin(phi))[:, None] * e1 + np.cos(phi)[:, None] * e2) + pert
    return pm, v, nrm
