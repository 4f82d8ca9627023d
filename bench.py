#!/usr/bin/env python
"""bench.py — headline benchmark of the B200 all-pairs gravity + Velocity-Verlet step.

    python bench.py --gpus N --steps K --warmup W            # our arm (CUDA, through the C ABI)
    python bench.py --impl reference --gpus N --steps K ...   # the reference's own CPU implementation (host cores)
    torchrun ... bench.py --gpus N ...                        # N > 1: one rank per GPU, NCCL

Headline workload (BASELINE.json configs[2], the N=1M point of the metric): SPIRAL_GALAXY N = 1,048,576 (config.json
parameters), PAIRWISE + VELOCITY_VERLET, FP32 pair math with FP64 accumulation. One "step" = one integrator step =
one all-pairs force evaluation (N^2 pair interactions, self pairs counted). Prints ONE JSON line on rank 0.

The same line carries, under "configs", the other BASELINE.json configurations at the same GPU count, each with its own
clock record, roofline (against the bound of ITS kernel) and CPU baseline: C2 (RANDOM 65,536 PAIRWISE), C4 (GALAXY 262,144
ADAPTIVE_MUTUAL, FP32 pair math + FP64 accumulation, and FP64 mode), C5 (SPIRAL 16,777,216 PAIRWISE, one timed step).
Every multi-GPU record also carries "parity": sampled accelerations of the stepped state against the long-double CPU
arbiter (oracle/), and the multi-GPU positions against a single-GPU run of the same steps; the process exits non-zero
when the FP32 tolerance of 1e-5 is exceeded.
"""
import argparse
import ctypes as C
import json
import os
import statistics
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(ROOT, "tests"))
import nbtest  # noqa: E402  (ctypes bindings + seeded generators; the oracle classes are used only as checker / CPU baseline)

# Algorithmic work per interaction and what the kernels execute (tools/sass_loop_mix.py)
SLOTS_PAIRWISE = 13          # FP32-pipe issue slots (+ 1 MUFU) per PAIRWISE interaction, algorithmic
PACKED_PER_PAIR = 11         # packed FFMA2/FMUL2/FADD2 instructions the far loop executes per PAIR of interactions
SLOTS_ADAPTIVE = 19          # 13 + the softening arithmetic; the bound is the MUFU (rsqrt, lg2, ex2 = 3 per interaction)
MUFU_ADAPTIVE = 3
DP_PAIRWISE = 16.0           # FP64 kernel: DFMA+DMUL+DADD per interaction, algorithmic (the SASS loop has 16.5)
DP_ADAPTIVE = 26.0           # + q^(2/3) (6) and the softened s^2 (4)
TOL_FP32, TOL_FP64 = 1e-5, 1e-12


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return json.load(open(p)), "MEASURED_PEAKS.json"
    return {"sm_max_mhz": 1965.0, "hbm_gbs": 6650.0}, "fallback (B200_PROFILING.md)"


def pipe_rates():
    """Issue rates measured by tools/pipe_bench on a B200 (warp instructions / clk / SM), committed under profiles/."""
    for name in ("r02_pipe_bench_b200.json", "r01_pipe_bench_b200.json"):
        p = os.path.join(ROOT, "profiles", name)
        if os.path.exists(p):
            r = json.load(open(p))["rates_warp_instr_per_clk_per_sm"]
            return {"ffma": r["ffma"], "ffma2": r["ffma2"], "mufu": r["mufu_rsq"], "dfma": r["dfma"], "source": "profiles/" + name}
    return {"ffma": 4.0, "ffma2": 2.0, "mufu": 0.5, "dfma": 2.0, "source": "nominal"}


class ClockSampler:
    """Samples nvidia-smi clocks / throttle reasons while the timed region runs."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index, period_ms=100):
        self.index, self.rows, self.proc, self.period = index, [], None, period_ms

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", str(self.period)],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def wait_first(self, timeout=3.0):
        """nvidia-smi needs a few hundred ms before its first line: short timed regions wait for it."""
        t0 = time.perf_counter()
        while self.proc and not self.rows and time.perf_counter() - t0 < timeout:
            time.sleep(0.02)

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, mx, reasons, power = [], [], set(), []
        for r in self.rows:
            try:
                sm.append(float(r[1])); mx.append(float(r[2])); power.append(float(r[3]))
            except (ValueError, IndexError):
                continue
            for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        # median over the samples taken under load (upper half of the power readings)
        thr = statistics.median(power)
        loaded = [s for s, p in zip(sm, power) if p >= thr] or sm
        return {"sm_mhz": statistics.median(loaded), "sm_max_mhz": max(mx), "power_w_max": max(power),
                "samples": len(sm), "reasons": sorted(reasons)}


def ncu_traffic_bytes(spec, world):
    """dram__bytes_read.sum + dram__bytes_write.sum per force evaluation from the committed `ncu --set full` capture
    (profiles/), valid for the headline workload only; None otherwise."""
    if not (spec["n"] == 1048576 and spec["force"] == "PAIRWISE" and spec["prec"] == "fp32" and spec["init"] == "spiral" and world == 1):
        return None, None
    for name in ("r02_ncu_force_far_near_1M.txt", "r01_ncu_force_far_near_splits64.txt"):
        path = os.path.join(ROOT, "profiles", name)
        try:
            tot = 0.0
            for line in open(path):
                if line.startswith("dram__bytes_read.sum [Mbyte]") or line.startswith("dram__bytes_write.sum [Mbyte]"):
                    v = float(line.split("=")[1])
                    if v == v:                      # ncu reports nan for kernels whose counters it could not collect
                        tot += v * 1e6
            if tot:
                return tot, "profiles/" + name
        except OSError:
            continue
    return None, None


def reference_cuda_bar(n):
    """Second same-box bar: the reference's own FP64 CUDA kernels compiled unmodified for sm_100a
    (oracle/_ref/ref_cuda_bench, built in the container where /root/reference exists), driven through the reference's
    runBenchmarkOnGPU on a spiral galaxy of the workload's size (2 steps + initial force). None if the binary is absent."""
    exe = os.path.join(ROOT, "oracle", "_ref", "ref_cuda_bench")
    if not os.path.exists(exe):
        return None
    try:
        r = subprocess.run([exe, str(n), "2", "0", "256"], capture_output=True, text=True, timeout=300)
        d = json.loads(r.stdout.strip().splitlines()[-1])
        return {"value": d["Gint_per_s"], "unit": "Gint/s", "seconds": d["seconds"], "impl": d["impl"],
                "note": "runBenchmarkOnGPU(BLOCK_SIZE=256) incl. its per-call cudaMalloc/H2D/D2H, FP64, no COM removal"}
    except Exception as e:  # noqa: BLE001  (a baseline that cannot run must not fail the bench)
        return {"unavailable": str(e)[:200]}


def host_cores(eng):
    """Threads for the CPU arm: every core this process may run on. omp_get_max_threads() is not used alone because
    torchrun exports OMP_NUM_THREADS=1 to its workers, which would silently turn the OpenMP reference into a serial run."""
    try:
        return max(len(os.sched_getaffinity(0)), 1)
    except AttributeError:
        return max(os.cpu_count() or 1, eng.max_threads())


def pinned(n_rows):
    import torch
    t = torch.empty((n_rows, 4), dtype=torch.float64).pin_memory()
    return t, t.numpy()


def generate(spec):
    """Seeded numpy restatements of the reference generators (same distributions; the reference itself cannot be seeded)."""
    n, init = spec["n"], spec["init"]
    if init == "spiral":
        return nbtest.gen_spiral(n, seed=42)
    if init == "random":
        return nbtest.gen_random(n, seed=42)
    if init == "galaxy":
        pm, v = nbtest.gen_galaxy((n - 1) // 3, 2, seed=42)     # 1 black hole + stars with two planets each
        assert pm.shape[0] == n, (pm.shape, n)
        return pm, v
    raise SystemExit(f"unknown init {init}")


def describe(spec):
    gen = {"spiral": "SPIRAL_GALAXY", "random": "RANDOM", "galaxy": "GALAXY (black hole + stars with 2 planets each)"}[spec["init"]]
    arith = "FP32 pair math (two-float positions) + FP64 near field + FP64 accumulation" if spec["prec"] == "fp32" else "FP64"
    return f"{gen} N={spec['n']} (config.json parameters, seed 42), {spec['force']} + VELOCITY_VERLET, {arith}"


def cpu_baseline(spec, steps=1, kind_pref="reference"):
    """The reference's own OpenMP CPU path (oracle/_ref = System::runBenchmark in CPU mode, compiled from the reference
    sources: pairwiseAcceleration or pairwiseAccelerationAdaptiveMutual as the config says) on a bounded sample of the
    workload; falls back to the C restatement (kind "port")."""
    ref = nbtest.Ref.try_load() if kind_pref == "reference" else None
    eng = ref if ref is not None else nbtest.Oracle()
    kind = "reference" if ref is not None else "port"
    cores = host_cores(eng)
    eng.set_threads(cores)
    force = nbtest.PAIRWISE if spec["force"] == "PAIRWISE" else nbtest.ADAPTIVE
    # sample size: ~10 s of CPU work. One VV step = 2 force evaluations = 2 n^2 interactions at ~0.27 (PAIRWISE) /
    # ~0.1 (ADAPTIVE_MUTUAL: cbrt + sqrt per pair) Gint/s/core.
    rate = (0.27e9 if force == nbtest.PAIRWISE else 0.1e9) * cores
    n_s = 16384
    while n_s < 131072 and 2.0 * (2 * n_s) ** 2 / rate < 12.0:
        n_s *= 2
    n_s = min(n_s, spec["n"])
    sample = dict(spec, n=n_s if spec["init"] != "galaxy" else 3 * ((n_s - 1) // 3) + 1)
    pm, v = generate(sample)
    n_s = pm.shape[0]
    t0 = time.perf_counter()
    if ref is not None:
        eng.run(pm, v, nbtest.VV, force, 0.04, steps)
    else:
        eng.step(pm, v, nbtest.VV, force, 0.04, steps)
    dt = time.perf_counter() - t0
    gints = 2.0 * steps * float(n_s) ** 2 / dt / 1e9
    fn = "pairwiseAcceleration" if force == nbtest.PAIRWISE else "pairwiseAccelerationAdaptiveMutual"
    return {"value": gints, "unit": "Gint/s", "cores": cores, "kind": kind, "seconds": dt,
            "sample": f"{spec['init']} N={n_s} (same generator/parameters as the workload"
                      f"{', full size' if n_s == spec['n'] else ''}), {steps} Velocity-Verlet step(s) = {2 * steps} all-pairs force "
                      f"evaluations ({fn}) via System::runBenchmark CPU/OpenMP; all-pairs Gint/s is N-independent"}


def run_reference_arm(args, rank, world):
    if rank != 0:
        return
    K, W = args.steps, args.warmup
    ref = nbtest.Ref.try_load()
    eng = ref if ref is not None else nbtest.Oracle()
    kind = "reference" if ref is not None else "port"
    cores = host_cores(eng); eng.set_threads(cores)
    n_s = args.ref_n
    pm, v = nbtest.gen_spiral(n_s, seed=42)
    run = (lambda: eng.run(pm, v, nbtest.VV, nbtest.PAIRWISE, 0.04, 1)) if ref is not None else \
          (lambda: eng.step(pm, v, nbtest.VV, nbtest.PAIRWISE, 0.04, 1))
    for _ in range(W):
        run()
    t0 = time.perf_counter()
    for _ in range(K):
        run()
    dt = time.perf_counter() - t0
    gints = 2.0 * K * float(n_s) ** 2 / dt / 1e9
    line = {"impl": "reference", "metric": "pair interactions/sec", "value": gints, "unit": "Gint/s", "n_gpus": args.gpus,
            "steps": K, "warmup": W, "ms_per_step": dt / K * 1e3, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": f"SPIRAL_GALAXY N={args.n} PAIRWISE + VELOCITY_VERLET (CPU arm timed on a bounded sample "
                                   f"N={n_s}; one CPU step = 2 force evaluations)", "sample_n": n_s},
            "cpu_baseline": {"value": gints, "unit": "Gint/s", "cores": cores, "kind": kind,
                             "sample": f"SPIRAL_GALAXY N={n_s}, {K} VV steps, System::runBenchmark CPU/OpenMP"},
            "e2e": {"value": gints, "unit": "Gint/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


def roofline_for(spec, n_sm, f_max, peak_src, rates, per_gpu_interactions, fk_ms, t_step_ms, world):
    """Roofline of the force evaluation of one configuration against the pipe that bounds ITS kernel."""
    ach = per_gpu_interactions / (fk_ms * 1e-3) if fk_ms > 0 else 0.0
    adaptive, fp64 = spec["force"] != "PAIRWISE", spec["prec"] == "fp64"
    r = {"kernel_ms": fk_ms, "share_of_step": fk_ms / t_step_ms if t_step_ms > 0 else None, "achieved_Tint_per_s": ach / 1e12}
    if fp64:
        dp = DP_ADAPTIVE if adaptive else DP_PAIRWISE
        peak = n_sm * (rates["dfma"] * 32.0 / dp) * f_max
        r.update({"bound": "fp64", "ops_per_interaction": dp, "unit": "TFLOP/s", "achieved": ach * dp * 2 / 1e12, "peak": peak * dp * 2 / 1e12,
                  "bound_note": f"FP64-pipe bound: {dp} DFMA/DMUL/DADD per interaction of force_fp64_kernel (algorithmic count, "
                                f"DESIGN.md 4) at the measured DFMA rate of {rates['dfma']:.3f} warp-instr/clk/SM",
                  "kernel": "force_fp64_kernel"})
    elif adaptive:
        peak = n_sm * (16.0 / MUFU_ADAPTIVE) * f_max
        r.update({"bound": "mufu", "ops_per_interaction": SLOTS_ADAPTIVE, "unit": "TFLOP/s", "achieved": ach * SLOTS_ADAPTIVE * 2 / 1e12,
                  "peak": peak * SLOTS_ADAPTIVE * 2 / 1e12,
                  "bound_note": "MUFU bound: 3 special-function operations per ADAPTIVE_MUTUAL interaction (rsqrt, lg2, ex2) at "
                                f"16 lanes/clk/SM (measured {rates['mufu'] * 32:.2f}); the {SLOTS_ADAPTIVE} FP32-pipe slots fit beside them",
                  "kernel": "force_fp32_r4_kernel<ADAPTIVE> (far) + force_near64_kernel<ADAPTIVE> (near)"})
    else:
        peak = n_sm * (128.0 / SLOTS_PAIRWISE) * f_max
        peak_mix = n_sm * (rates["ffma2"] * 64.0 / PACKED_PER_PAIR) * f_max
        r.update({"bound": "fp32", "ops_per_interaction": SLOTS_PAIRWISE, "unit": "TFLOP/s", "achieved": ach * SLOTS_PAIRWISE * 2 / 1e12,
                  "peak": peak * SLOTS_PAIRWISE * 2 / 1e12,
                  "bound_note": "compute-bound on FP32 issue (not hbm, not tensor): 13 FP32-pipe slots + 1 MUFU per PAIRWISE "
                                "interaction (SURVEY.md 8d); achieved/peak are interactions/s x 13 slots x 2 flop",
                  "executed_mix": {"packed_instr_per_pair": PACKED_PER_PAIR, "ffma2_rate_warp_instr_per_clk_per_sm": rates["ffma2"],
                                   "peak_Tint_per_s": peak_mix / 1e12, "frac": ach / peak_mix,
                                   "note": "against what the far loop actually issues: 11 packed FFMA2/FMUL2/FADD2 per pair of "
                                           "interactions at the measured FFMA2 issue rate (tools/pipe_bench)"},
                  "kernel": "force_fp32_r4_kernel (far, FP32) + force_near64_kernel (near, FP64, own stream); per-GPU time from "
                            "CUDA events on the context stream around both"})
    r["frac"] = ach / peak
    r["peak_Tint_per_s"] = peak / 1e12
    r["traffic"], traffic_src = ncu_traffic_bytes(spec, world)
    if r["traffic"]:
        r["traffic_note"] = (f"DRAM bytes per force evaluation from {traffic_src} (far kernel; ncu returned no DRAM counters for the near kernel at this grid size — its 32 partial slots add about 0.8 GB); algorithmic: packed tiles "
                             f"({40 * spec['n'] / 1e6:.0f} MB, re-read from L2 by every CTA) + result ({24 * spec['n'] / 1e6:.0f} MB); the rest is "
                             "the partial-sum slots of the fixed-order reduction, written once and read once")
    r["peak_source"] = (f"{n_sm} SMs x sm_max_mhz from {peak_src}; pipe rates (warp-instr/clk/SM) ffma {rates['ffma']:.3f} ffma2 "
                        f"{rates['ffma2']:.3f} mufu {rates['mufu']:.3f} dfma {rates['dfma']:.3f} from {rates['source']}; no FP32/FP64 entry "
                        "in MEASURED_PEAKS.json")
    r["hbm_check_GBps"] = per_gpu_interactions * 40.0 / 1024.0 / (fk_ms * 1e-3) / 1e9 if fk_ms > 0 else None
    return r


class Runner:
    def __init__(self, lib, rank, world, local_rank, opts=()):
        self.opts = [(kv.split("=")[0], int(kv.split("=")[1])) for kv in opts]
        import torch
        import torch.distributed as dist
        self.torch, self.dist = torch, dist
        self.lib, self.rank, self.world, self.local = lib, rank, world, local_rank
        self.oracle = None

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def comm_id(self):
        torch, dist, lib = self.torch, self.dist, self.lib
        idbuf = torch.zeros(128, dtype=torch.uint8)
        if self.rank == 0:
            raw = (C.c_char * 128)()
            lib.check(lib.l.nb_comm_unique_id(raw))
            idbuf = torch.frombuffer(bytearray(raw.raw), dtype=torch.uint8).clone()
        idbuf = idbuf.cuda()
        dist.broadcast(idbuf, 0)
        return (C.c_char * 128).from_buffer_copy(bytes(idbuf.cpu().numpy().tobytes()))

    def max_over_ranks(self, x):
        t = self.torch.tensor([x], dtype=self.torch.float64, device="cuda")
        if self.world > 1:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def gather(self, x):
        t = self.torch.tensor([x], dtype=self.torch.float64, device="cuda")
        if self.world == 1:
            return [float(x)]
        out = [self.torch.zeros_like(t) for _ in range(self.world)]
        self.dist.all_gather(out, t)
        return [float(o.item()) for o in out]

    def make_context(self, spec, variant=0):
        lib = self.lib
        prec = nbtest.FP32 if spec["prec"] == "fp32" else nbtest.FP64
        force = nbtest.PAIRWISE if spec["force"] == "PAIRWISE" else nbtest.ADAPTIVE
        ctx = lib.context(spec["n"], prec, device=self.local)
        if variant:
            ctx.set_option("force_variant", variant)
        for k, val in self.opts:
            ctx.set_option(k, val)
        if self.world > 1:
            lib.check(lib.l.nb_comm_init(ctx.h, self.comm_id(), self.rank, self.world))
        ctx.set_methods(nbtest.VV, force, 1)
        return ctx, prec, force

    def parity(self, ctx, spec, pm0, v0, force, prec, total_steps, dt_sim, single_limit=2 ** 21):
        """Untimed. Collective (every rank calls): accelerations + state of the stepped system; rank 0 checks sampled
        particles (both sides of every slice boundary, the first and last real particle, random ones) against the long-double
        CPU arbiter, and — multi-GPU, systems up to single_limit particles — the positions against a single-GPU context that
        repeats the same steps on rank 0's device."""
        n, world, lib = spec["n"], self.world, self.lib
        n_sample = 256 if n <= 2 ** 21 else 96          # the arbiter costs n_sample x n long-double interactions on the host
        a = ctx.accel()
        p, v = ctx.download()
        order = ctx.order()
        out = None
        if self.rank == 0:
            if self.oracle is None:
                self.oracle = nbtest.Oracle()
                self.oracle.set_threads(host_cores(self.oracle))
            slots = {0, n - 1}
            for r in range(world):
                b, cnt = lib.partition_bounds(n, r, world)
                for s in (b - 1, b, b + 1, b + cnt - 1):
                    if 0 <= s < n:
                        slots.add(s)
            idx = set(int(order[s]) for s in slots)
            rng = np.random.default_rng(1)
            while len(idx) < n_sample:
                idx.add(int(rng.integers(0, n)))
            idx = np.array(sorted(idx), dtype=np.int64)
            t0 = time.perf_counter()
            ref = self.oracle.forces_subset(p, force, idx)
            err = nbtest.rel_err(a[idx], ref)
            tol = TOL_FP32 if prec == nbtest.FP32 else TOL_FP64
            out = {"max_rel_err": float(err.max()), "median_rel_err": float(np.median(err)), "n_checked": int(len(idx)),
                   "tolerance": tol, "vs": "oracle/ long-double arbiter (nbo_forces_subset) on the state after the timed steps",
                   "includes": "both sides of every slice boundary in device order, first/last slot, random particles",
                   "arbiter_seconds": time.perf_counter() - t0, "multi_vs_single_pos": None}
            ok = bool(np.isfinite(a).all() and err.max() <= tol)
            if world > 1 and n <= single_limit:
                with lib.context(n, prec, device=self.local) as one:
                    one.set_methods(nbtest.VV, force, 1)
                    one.upload(pm0, v0); one.compute_accel(); one.step(dt_sim, total_steps)
                    p1, v1 = one.download()
                scale = float(np.abs(p1[:, :3]).max())
                dpos = float(np.abs(p[:, :3] - p1[:, :3]).max())
                dvel = float(np.abs(v[:, :3] - v1[:, :3]).max())
                out.update({"multi_vs_single_pos": dpos / scale, "multi_vs_single_pos_abs_AU": dpos, "multi_vs_single_vel_abs": dvel,
                            "multi_vs_single_steps": total_steps})
                ok = ok and dpos / scale <= 1e-9
            elif world > 1:
                out["multi_vs_single_note"] = f"skipped above {single_limit} particles (a single-GPU repeat would take minutes)"
            out["ok"] = ok
        if world > 1:
            self.dist.barrier()
        return out

    def run(self, spec, K, W, dt_sim=0.04, variant=0, want_parity=True, keep_ctx=False, progress=False):
        torch, lib, rank, world = self.torch, self.lib, self.rank, self.world
        say = (lambda m: print(f"[bench {spec.get('id', '')}] {m}", file=sys.stderr, flush=True)) if (progress and rank == 0) else (lambda m: None)
        t_wall = time.perf_counter()
        pm, v = generate(spec)
        say(f"generated {spec['n']} particles in {time.perf_counter() - t_wall:.1f} s")
        ctx, prec, force = self.make_context(spec, variant)
        ctx.upload(pm, v)                       # inputs resident in HBM before the timed region
        ctx.compute_accel()
        ctx.sync()
        say(f"upload + initial force evaluation done at {time.perf_counter() - t_wall:.1f} s")
        slice_work = None
        if world > 1 and spec["n"] <= 2 ** 21:
            # diagnostic, untimed: one more evaluation of the own slice with no exchange in front of it (the tiles are all there
            # after the upload), so the ranks do not wait for each other: the time each rank's share of the work takes
            ctx.force_timing(reset=True); ctx.compute_accel(); ctx.sync()
            slice_work = self.gather(ctx.force_timing(reset=True)[0])
        if W > 0:
            ctx.step(dt_sim, W); ctx.sync()
            say(f"{W} warm-up step(s) done at {time.perf_counter() - t_wall:.1f} s")
        sampler = ClockSampler(self.local)      # every rank watches its own GPU; rank 0's record is the line's "clocks"
        self.barrier()
        sampler.start(); sampler.wait_first()
        launches0 = ctx.launches()
        # Device timing: CUDA events recorded on the context's own stream (nb_timer_start/stop); the force-kernel time is
        # the mean over the K evaluations of the event pairs the context records around every force pass.
        ctx.force_timing(reset=True)
        self.barrier()
        ctx.timer_start()                       # CUDA event on the context's stream
        ctx.step(dt_sim, K)                     # K steps enqueued back to back, no host synchronisation in between
        t_local = ctx.timer_stop() * 1e-3       # CUDA event, waits for the stream
        near_ms, _ = ctx.near_timing()
        fk_ms, fk_evals = ctx.force_timing(reset=True)
        torch.cuda.synchronize()
        launches = ctx.launches() - launches0
        clocks = sampler.stop()
        per_rank_mhz = self.gather(clocks["sm_mhz"] or 0.0)
        per_rank_slow = self.gather(1.0 if set(clocks["reasons"]) & {"hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown"} else 0.0)
        say(f"{K} timed step(s) done at {time.perf_counter() - t_wall:.1f} s")
        t_step = self.max_over_ranks(t_local) / K
        per_rank_force = self.gather(fk_ms)
        per_rank_near = self.gather(near_ms)
        per_rank_step = self.gather(t_local / K * 1e3)
        n = spec["n"]
        interactions = float(n) * float(n)
        res = {"id": spec.get("id"), "workload": describe(spec), "n": n, "value": interactions / t_step / 1e9, "unit": "Gint/s",
               "ms_per_step": t_step * 1e3, "steps_per_s": 1.0 / t_step, "steps": K, "warmup": W,
               "dtype": "f32" if prec == nbtest.FP32 else "f64", "n_gpus": world, "gpu_launches": launches, "clocks": clocks,
               "per_rank": {"force_ms": {"min": min(per_rank_force), "mean": sum(per_rank_force) / world, "max": max(per_rank_force)},
                            "near_kernel_span_ms": {"min": min(per_rank_near), "mean": sum(per_rank_near) / world, "max": max(per_rank_near)},
                            "step_ms": {"min": min(per_rank_step), "mean": sum(per_rank_step) / world, "max": max(per_rank_step)},
                            "force_ms_by_rank": per_rank_force, "slice_work_ms_by_rank": slice_work, "sm_mhz_by_rank": per_rank_mhz, "ranks_with_slowdown": int(sum(per_rank_slow))}}
        res["_fk_ms"] = max(per_rank_force)
        res["_interactions"] = interactions
        if want_parity:
            par = self.parity(ctx, spec, pm, v, force, prec, W + K, dt_sim)
            say(f"parity check done at {time.perf_counter() - t_wall:.1f} s")
            if par is not None:
                res["parity"] = par
        if keep_ctx:
            return res, ctx, (pm, v, prec, force)
        ctx.close()
        return res, None, None


def finish_entry(res, spec, runner, peaks, peak_src, rates, n_sm):
    f_max = float(peaks.get("sm_max_mhz", 1965.0)) * 1e6
    fk_ms = res.pop("_fk_ms"); inter = res.pop("_interactions")
    res["roofline"] = roofline_for(spec, n_sm, f_max, peak_src, rates, inter / runner.world, fk_ms, res["ms_per_step"], runner.world)
    return res


EXTRA_CONFIGS = [
    {"id": "C2", "init": "random", "n": 65536, "force": "PAIRWISE", "prec": "fp32", "steps": 200, "warmup": 20},
    {"id": "C4", "init": "galaxy", "n": 262144, "force": "ADAPTIVE_MUTUAL", "prec": "fp32", "steps": 20, "warmup": 3},
    {"id": "C4-fp64mode", "init": "galaxy", "n": 262144, "force": "ADAPTIVE_MUTUAL", "prec": "fp64", "steps": 8, "warmup": 3},
    {"id": "fp64-pairwise", "init": "spiral", "n": 262144, "force": "PAIRWISE", "prec": "fp64", "steps": 8, "warmup": 3},
    {"id": "C5", "init": "spiral", "n": 16777216, "force": "PAIRWISE", "prec": "fp32", "steps": 1, "warmup": 0},
]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--particles", "--n", dest="n", type=int, default=1048576, help="particles; use --particles under torchrun (its parser treats --n as ambiguous)")
    ap.add_argument("--init", default="spiral", choices=["spiral", "random", "galaxy"])
    ap.add_argument("--force", default="PAIRWISE", choices=["PAIRWISE", "ADAPTIVE_MUTUAL"])
    ap.add_argument("--precision", default="fp32", choices=["fp32", "fp64"])
    ap.add_argument("--variant", type=int, default=0, help="force kernel tile staging: 0 auto, 1 loads, 2 bulk copy")
    ap.add_argument("--ref-n", type=int, default=32768, help="sample size of the --impl reference arm")
    ap.add_argument("--configs", default="all", help="extra BASELINE.json configurations in the same line: all | none | comma list of "
                                                     + ",".join(c["id"] for c in EXTRA_CONFIGS))
    ap.add_argument("--opt", action="append", default=[], help="nb_set_option key=value applied to every context (tuning experiments)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-parity", action="store_true")
    args = ap.parse_args()

    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference_arm(args, rank, world)
        return

    import torch
    import torch.distributed as dist
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (the product has no CPU path; use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    assert world == args.gpus or world == 1 and args.gpus == 1, f"--gpus {args.gpus} but WORLD_SIZE={world}"

    lib = nbtest.Lib()
    runner = Runner(lib, rank, world, local_rank, args.opt)
    peaks, peak_src = measured_peaks()
    rates = pipe_rates()
    n_sm = torch.cuda.get_device_properties(local_rank).multi_processor_count
    K, W, dt_sim = args.steps, max(args.warmup, 0), 0.04
    head = {"id": "headline", "init": args.init, "n": args.n, "force": args.force, "prec": args.precision}
    n = args.n

    res, ctx, (pm_np, v_np, prec, force) = runner.run(head, K, W, dt_sim, variant=args.variant, want_parity=not args.no_parity,
                                                      keep_ctx=True)
    interactions = float(n) * float(n)

    # ---- e2e: the reference-facing call with HOST buffers (runBenchmarkOnGPU semantics), copies inside the timing
    e2e = None
    if not args.no_e2e:
        pm_t, pm = pinned(n); v_t, v = pinned(n)
        pm[:] = pm_np; v[:] = v_np
        out_pm_t, out_pm = pinned(n); out_v_t, out_v = pinned(n)
    if not args.no_e2e and world == 1:
        Ke = max(1, min(K, 3))
        lib.check(lib.l.nb_run_benchmark(n, nbtest._p(pm), nbtest._p(v), nbtest.VV, force, dt_sim, 0, 256, prec, 1, local_rank))
        torch.cuda.synchronize()
        t_sum = 0.0
        for _ in range(Ke):
            out_pm[:] = pm; out_v[:] = v      # refresh of the in/out host buffers: not part of the call, not timed
            t0 = time.perf_counter()
            lib.check(lib.l.nb_run_benchmark(n, nbtest._p(out_pm), nbtest._p(out_v), nbtest.VV, force, dt_sim, 1, 256,
                                             prec, 1, local_rank))
            t_sum += time.perf_counter() - t0  # the call returns after its D2H copy and nb_destroy: fully synchronous
        t_e = t_sum / Ke
        e2e = {"value": 2.0 * interactions / t_e / 1e9, "unit": "Gint/s", "h2d_bytes_per_step": 64 * n,
               "d2h_bytes_per_step": 64 * n, "ms_per_call": t_e * 1e3,
               "call": "nb_run_benchmark(steps=1) on pinned host Particles arrays: create + H2D + initial force + 1 VV "
                       "step + D2H + destroy = 2 force evaluations per call"}
    elif not args.no_e2e:
        # multi-GPU: upload (H2D) + step + download (D2H) per timed call through the context API
        Ke = max(1, min(K, 3))
        out_idx_t = torch.empty(n, dtype=torch.int32).pin_memory(); out_idx = out_idx_t.numpy().view(np.uint32)
        ctx.upload(pm, v); ctx.step(dt_sim, 1); ctx.download_slice(out_pm, out_v, out_idx)      # untimed first call
        runner.barrier(); t0 = time.perf_counter()
        for _ in range(Ke):
            ctx.upload(pm, v); ctx.step(dt_sim, 1); rows = ctx.download_slice(out_pm, out_v, out_idx)[0].shape[0]
        torch.cuda.synchronize()
        t_e = runner.max_over_ranks(time.perf_counter() - t0) / Ke
        e2e = {"value": 2.0 * interactions / t_e / 1e9, "unit": "Gint/s", "h2d_bytes_per_step": 64 * n,
               "d2h_bytes_per_step": 68 * n, "ms_per_call": t_e * 1e3,
               "call": "per call, all ranks together: nb_upload (each rank copies its share of the rows H2D, all-gather over NVLink) + "
                       "nb_step(1) (initial force + 1 VV step) + nb_download_slice (each rank copies its own rows and their caller "
                       "indices D2H): 2 force evaluations per call; bytes are the sum over ranks"}
    ctx.close()

    # ---- the other BASELINE.json configurations at this GPU count
    want = [c["id"] for c in EXTRA_CONFIGS] if args.configs == "all" else ([] if args.configs == "none" else args.configs.split(","))
    extra = []
    for spec in EXTRA_CONFIGS:
        if spec["id"] not in want:
            continue
        r, _, _ = runner.run(spec, spec["steps"], spec["warmup"], dt_sim, want_parity=not args.no_parity, progress=(spec["n"] > 4000000))
        if rank == 0:
            r = finish_entry(r, spec, runner, peaks, peak_src, rates, n_sm)
            if spec["steps"] < 3 or spec["warmup"] < 3:
                r["note"] = ("one timed step after the initial force evaluation (which is the warm-up: same kernels, same size); a step "
                             "takes tens of seconds on one GPU, so start-up effects are < 1e-4 of it")
            if not args.no_cpu_baseline and spec["id"] in ("C2", "C4"):
                r["cpu_baseline"] = cpu_baseline(spec)
            extra.append(r)

    bad = []
    if rank == 0:
        res = finish_entry(res, head, runner, peaks, peak_src, rates, n_sm)
        line = {"metric": "pair interactions/sec", "value": res["value"], "unit": "Gint/s", "n_gpus": world, "steps": K,
                "warmup": W, "ms_per_step": res["ms_per_step"], "steps_per_s": res["steps_per_s"], "higher_is_better": True,
                "scaling": "strong", "vs_baseline": None, "dtype": res["dtype"], "data": "synthetic",
                "config": {"workload": describe(head), "n": n, "dt": dt_sim,
                           "parallelism": f"i-slices x{world}" if world > 1 else "single GPU",
                           "inputs": "seeded numpy restatement of the reference generator (same distribution; the reference "
                                     "seeds from std::random_device)",
                           "l2_policy": "inputs larger than L2 are not needed: each step rewrites positions; packed tiles "
                                        f"({40 * n / 1e6:.0f} MB) stream from L2/HBM every step"},
                "roofline": res["roofline"], "clocks": res["clocks"], "gpu_launches": res["gpu_launches"], "per_rank": res["per_rank"]}
        if "parity" in res:
            line["parity"] = res["parity"]
        if e2e:
            line["e2e"] = e2e
        if not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline(head)
            if world == 1 and args.force == "PAIRWISE":
                rc = reference_cuda_bar(n)
                if rc:
                    line["reference_cuda_fp64"] = rc
        if extra:
            line["configs"] = extra
        print(json.dumps(line), flush=True)
        for r in [line] + extra:
            p = r.get("parity")
            if p and not p["ok"]:
                bad.append((r.get("id", "headline"), p["max_rel_err"], p.get("multi_vs_single_pos")))
    if world > 1:
        flag = torch.tensor([len(bad)], device="cuda")
        dist.broadcast(flag, 0)
        dist.barrier()
        dist.destroy_process_group()
        if int(flag.item()):
            if rank == 0:
                print(f"bench.py: PARITY FAILURE {bad}", file=sys.stderr)
            sys.exit(3)
    elif bad:
        print(f"bench.py: PARITY FAILURE {bad}", file=sys.stderr)
        sys.exit(3)


if __name__ == "__main__":
    main()
