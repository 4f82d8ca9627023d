#!/usr/bin/env python
"""bench.py — headline benchmark of the B200 all-pairs gravity + Velocity-Verlet step.

    python bench.py --gpus N --steps K --warmup W            # our arm (CUDA, through the C ABI)
    python bench.py --impl reference --gpus N --steps K ...   # the reference's own CPU implementation (host cores)
    torchrun ... bench.py --gpus N ...                        # N > 1: one rank per GPU, NCCL

Workload (BASELINE.json configs[2], the N=1M point of the metric): SPIRAL_GALAXY N = 1,048,576 (config.json
parameters), PAIRWISE + VELOCITY_VERLET, FP32 pair math with FP64 accumulation. One "step" = one integrator step =
one all-pairs force evaluation (N^2 pair interactions, self pairs counted). Prints ONE JSON line on rank 0.
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
import nbtest  # noqa: E402  (ctypes bindings + seeded generators; the oracle classes are used only for cpu_baseline)

N_SM_NOMINAL = 148
SLOTS_PER_INTERACTION = 13          # 13 FP32-pipe issue slots + 1 MUFU per PAIRWISE interaction


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return json.load(open(p)), "MEASURED_PEAKS.json"
    return {"sm_max_mhz": 1965.0, "hbm_gbs": 6650.0}, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """Samples nvidia-smi clocks / throttle reasons while the timed region runs."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

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


def ncu_traffic_bytes(n, force, prec, world):
    """dram__bytes_read.sum + dram__bytes_write.sum per force-kernel launch from the committed `ncu --set full` capture
    (profiles/), valid for the default workload only; None otherwise."""
    if not (n == 1048576 and force == nbtest.PAIRWISE and prec == nbtest.FP32 and world == 1):
        return None
    path = os.path.join(ROOT, "profiles", "r01_ncu_force_far_near_splits64.txt")
    try:
        tot = 0.0
        for line in open(path):
            if line.startswith("dram__bytes_read.sum [Mbyte]") or line.startswith("dram__bytes_write.sum [Mbyte]"):
                tot += float(line.split("=")[1]) * 1e6
        return tot or None
    except OSError:
        return None


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


def cpu_baseline(workload_n, steps=1, kind_pref="reference"):
    """The reference's own OpenMP CPU path (oracle/_ref = System::runBenchmark in CPU mode, compiled from the
    reference sources) on a bounded sample of the workload; falls back to the C restatement (kind "port")."""
    ref = nbtest.Ref.try_load() if kind_pref == "reference" else None
    eng = ref if ref is not None else nbtest.Oracle()
    kind = "reference" if ref is not None else "port"
    cores = host_cores(eng)
    eng.set_threads(cores)
    # sample size: ~10-30 s of CPU work. One VV step = 2 force evaluations = 2 n^2 interactions at ~0.27 Gint/s/core.
    n_s = 16384
    while n_s < 131072 and 2.0 * (2 * n_s) ** 2 / (0.27e9 * cores) < 12.0 * 1.0:
        n_s *= 2
    n_s = min(n_s, workload_n)
    pm, v = nbtest.gen_spiral(n_s, seed=42)
    t0 = time.perf_counter()
    if ref is not None:
        eng.run(pm, v, nbtest.VV, nbtest.PAIRWISE, 0.04, steps)
    else:
        eng.step(pm, v, nbtest.VV, nbtest.PAIRWISE, 0.04, steps)
    dt = time.perf_counter() - t0
    gints = 2.0 * steps * float(n_s) ** 2 / dt / 1e9
    return {"value": gints, "unit": "Gint/s", "cores": cores, "kind": kind, "seconds": dt,
            "sample": f"SPIRAL_GALAXY N={n_s} (same generator/parameters as the workload), {steps} Velocity-Verlet step(s) "
                      f"= {2 * steps} all-pairs force evaluations via System::runBenchmark CPU/OpenMP; all-pairs Gint/s is "
                      f"N-independent"}, n_s


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


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--particles", "--n", dest="n", type=int, default=1048576, help="particles (SPIRAL_GALAXY); use --particles under torchrun (its parser treats --n as ambiguous)")
    ap.add_argument("--force", default="PAIRWISE", choices=["PAIRWISE", "ADAPTIVE_MUTUAL"])
    ap.add_argument("--precision", default="fp32", choices=["fp32", "fp64"])
    ap.add_argument("--variant", type=int, default=0, help="force kernel tile staging: 0 auto, 1 loads, 2 bulk copy")
    ap.add_argument("--ref-n", type=int, default=32768, help="sample size of the --impl reference arm")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
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
    n = args.n
    force = nbtest.PAIRWISE if args.force == "PAIRWISE" else nbtest.ADAPTIVE
    prec = nbtest.FP32 if args.precision == "fp32" else nbtest.FP64
    K, W, dt_sim = args.steps, max(args.warmup, 0), 0.04

    # synthetic workload, generated once on the host (same seed on every rank) into pinned memory
    pm_np, v_np = nbtest.gen_spiral(n, seed=42)
    pm_t, pm = pinned(n); v_t, v = pinned(n)
    pm[:] = pm_np; v[:] = v_np
    out_pm_t, out_pm = pinned(n); out_v_t, out_v = pinned(n)

    ctx = lib.context(n, prec, device=local_rank)
    if args.variant:
        ctx.set_option("force_variant", args.variant)
    if world > 1:
        idbuf = torch.zeros(128, dtype=torch.uint8)
        if rank == 0:
            raw = (C.c_char * 128)()
            lib.check(lib.l.nb_comm_unique_id(raw))
            idbuf = torch.frombuffer(bytearray(raw.raw), dtype=torch.uint8).clone()
        idbuf = idbuf.cuda()
        dist.broadcast(idbuf, 0)
        raw = (C.c_char * 128).from_buffer_copy(bytes(idbuf.cpu().numpy().tobytes()))
        lib.check(lib.l.nb_comm_init(ctx.h, raw, rank, world))
    ctx.set_methods(nbtest.VV, force, 1)
    ctx.upload(pm, v)                       # inputs resident in HBM before the timed region
    ctx.compute_accel()
    ctx.sync()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    ctx.step(dt_sim, W); ctx.sync()
    sampler = ClockSampler(local_rank)
    barrier()
    if rank == 0:
        sampler.start()
    launches0 = ctx.launches()
    # Device timing: CUDA events recorded on the context's own stream (nb_timer_start/stop); the force-kernel time is
    # the mean over the K evaluations of the event pairs the context records around every force pass.
    ctx.force_timing(reset=True)
    barrier()
    ctx.timer_start()                       # CUDA event on the context's stream
    ctx.step(dt_sim, K)                     # K steps enqueued back to back, no host synchronisation in between
    t_local = ctx.timer_stop() * 1e-3       # CUDA event, waits for the stream
    fk_ms, fk_evals = ctx.force_timing(reset=True)
    torch.cuda.synchronize()
    launches = ctx.launches() - launches0
    clocks = sampler.stop() if rank == 0 else None
    t_all = torch.tensor([t_local], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t_all, op=dist.ReduceOp.MAX)
    t_step = float(t_all.item()) / K
    interactions = float(n) * float(n)
    value = interactions / t_step / 1e9

    # ---- e2e: the reference-facing call with HOST buffers (runBenchmarkOnGPU semantics), copies inside the timing
    e2e = None
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
        # multi-GPU: upload (H2D) + step + download (D2H of the rank's slice) per timed step through the context API
        Ke = max(1, min(K, 3))
        barrier(); t0 = time.perf_counter()
        for _ in range(Ke):
            ctx.upload(pm, v); ctx.step(dt_sim, 1); lib.check(lib.l.nb_download(ctx.h, nbtest._p(out_pm), nbtest._p(out_v)))
        torch.cuda.synchronize()
        te = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device="cuda")
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
        t_e = float(te.item()) / Ke
        e2e = {"value": 2.0 * interactions / t_e / 1e9, "unit": "Gint/s", "h2d_bytes_per_step": 64 * n * world,
               "d2h_bytes_per_step": 64 * n, "ms_per_call": t_e * 1e3,
               "call": "per step and rank: nb_upload (full system H2D) + nb_step(1) (initial force + 1 VV step) + "
                       "nb_download (slice D2H)"}

    if rank == 0:
        peaks, peak_src = measured_peaks()
        f_max = float(peaks.get("sm_max_mhz", 1965.0)) * 1e6
        n_sm = torch.cuda.get_device_properties(local_rank).multi_processor_count
        per_gpu_interactions = interactions / world
        ach_int = per_gpu_interactions / (fk_ms * 1e-3)
        peak_int = n_sm * (128.0 / SLOTS_PER_INTERACTION) * f_max
        roofline = {"bound": "fp32",
                    "bound_note": "compute-bound on FP32 issue (not hbm, not tensor): 13 FP32-pipe slots + 1 MUFU per PAIRWISE "
                                  "interaction (SURVEY.md 8d); achieved/peak are interactions/s x 13 slots x 2 flop",
                    "achieved": ach_int * SLOTS_PER_INTERACTION * 2 / 1e12,
                    "peak": peak_int * SLOTS_PER_INTERACTION * 2 / 1e12, "unit": "TFLOP/s",
                    "frac": ach_int / peak_int, "traffic": ncu_traffic_bytes(n, force, prec, world),
                    "traffic_note": "DRAM bytes per force evaluation (far + near kernel) from "
                                    "profiles/r01_ncu_force_far_near_splits64.txt: 42 MB of packed tiles re-read from L2 by every "
                                    "CTA; the rest is the 74 partial-sum slots (25 MB each, fixed-order reduction) written once",
                    "achieved_Tint_per_s": ach_int / 1e12, "peak_Tint_per_s": peak_int / 1e12,
                    "kernel": "force evaluation = force_fp32_r4_kernel (far, FP32) + force_near64_kernel (near, FP64, own stream); "
                              "per-GPU time from CUDA events on the context stream around both",
                    "kernel_ms": fk_ms, "share_of_step": fk_ms / (t_step * 1e3),
                    "peak_source": f"{n_sm} SMs x 128 FP32 lanes x sm_max_mhz from {peak_src}; no FP32 entry in "
                                   f"MEASURED_PEAKS.json (tools/pipe_bench measures the FFMA/FFMA2/MUFU rates)",
                    "hbm_check_GBps": per_gpu_interactions * 32.0 / 512.0 / (fk_ms * 1e-3) / 1e9}
        line = {"metric": "pair interactions/sec", "value": value, "unit": "Gint/s", "n_gpus": world, "steps": K,
                "warmup": W, "ms_per_step": t_step * 1e3, "steps_per_s": 1.0 / t_step, "higher_is_better": True,
                "scaling": "strong", "vs_baseline": None,
                "dtype": "f32" if prec == nbtest.FP32 else "f64", "data": "synthetic",
                "config": {"workload": f"SPIRAL_GALAXY N={n} (config.json parameters, seed 42), {args.force} + "
                                       f"VELOCITY_VERLET, {'FP32 pair math (hi+lo positions) + FP64 accumulation' if prec == nbtest.FP32 else 'FP64'}",
                           "n": n, "dt": dt_sim, "parallelism": f"i-slices x{world}" if world > 1 else "single GPU",
                           "l2_policy": "inputs larger than L2 are not needed: each step rewrites positions; packed tiles "
                                        f"({32 * n / 1e6:.0f} MB) stream from L2/HBM every step"},
                "roofline": roofline, "clocks": clocks, "gpu_launches": launches}
        if e2e:
            line["e2e"] = e2e
        if not args.no_cpu_baseline and world >= 1:
            cb, _ = cpu_baseline(n)
            line["cpu_baseline"] = cb
            if world == 1:
                ctx.close()                      # free the device before the reference's kernels allocate
                rc = reference_cuda_bar(n)
                if rc:
                    line["reference_cuda_fp64"] = rc
        print(json.dumps(line), flush=True)
    ctx.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
