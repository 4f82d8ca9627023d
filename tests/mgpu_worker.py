"""Multi-GPU worker (one process per GPU, launched by torchrun from tests/test_multigpu.py or by hand):
   torchrun --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tests/mgpu_worker.py [n] [fp32|fp64] [pairwise|adaptive]
Checks the sliced context (NCCL all-gather per step) against a single-GPU context and the CPU oracle."""
import ctypes as C
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import nbtest  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 20000
prec = nbtest.FP64 if (len(sys.argv) > 2 and sys.argv[2] == "fp64") else nbtest.FP32
force = nbtest.ADAPTIVE if (len(sys.argv) > 3 and sys.argv[3] == "adaptive") else nbtest.PAIRWISE
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
lib = nbtest.Lib()
pm, v = nbtest.gen_spiral(n, seed=77)

idbuf = torch.zeros(128, dtype=torch.uint8, device="cuda")
if rank == 0:
    raw = (C.c_char * 128)()
    lib.check(lib.l.nb_comm_unique_id(raw))
    idbuf = torch.frombuffer(bytearray(raw.raw), dtype=torch.uint8).cuda()
dist.broadcast(idbuf, 0)
raw = (C.c_char * 128).from_buffer_copy(idbuf.cpu().numpy().tobytes())

tol = 1e-5 if prec == nbtest.FP32 else 1e-12
with lib.context(n, prec, device=local) as ctx:
    lib.check(lib.l.nb_comm_init(ctx.h, raw, rank, world))
    ctx.set_methods(nbtest.VV, force, 1)
    ctx.set_option("resort_interval", 2)                  # exercise the collective re-sort inside the 3 steps
    ctx.upload(pm, v)
    ctx.compute_accel()
    a_multi = ctx.accel()
    ctx.step(0.04, 3)
    e_mid = ctx.energy()                                  # collective: slice potentials + all-reduce; must not disturb the run
    ctx.step(0.04, 1)
    p_multi, v_multi = ctx.download()
    # egress without the gather: every rank's own rows + their caller indices; together they must be the downloaded state
    ps, vs, idx = ctx.download_slice()
    full = torch.zeros((n, 8), dtype=torch.float64, device="cuda")
    full[torch.from_numpy(idx.astype(np.int64)).cuda()] = torch.from_numpy(np.hstack([ps, vs])).cuda()
    rows = torch.tensor([len(idx)], device="cuda")
    dist.all_reduce(full); dist.all_reduce(rows)
    slice_ok = int(rows.item()) == n and np.array_equal(full.cpu().numpy(), np.hstack([p_multi, v_multi]))
    mom = ctx.momentum()
    e_multi = ctx.energy()
    snap6 = np.zeros((n, 6)); sys2 = np.zeros(2); stat2 = np.zeros((n, 2))
    lib.check(lib.l.nb_snapshot(ctx.h, C.c_double(0.16), nbtest._p(snap6), nbtest._p(sys2), nbtest._p(stat2)))
with lib.context(n, prec, device=local) as ctx:          # the same thing on one GPU
    ctx.set_methods(nbtest.VV, force, 1)
    ctx.upload(pm, v)
    ctx.compute_accel()
    a_single = ctx.accel()
    ctx.step(0.04, 4)
    p_single, v_single = ctx.download()
    e_single = ctx.energy()
err_a = nbtest.rel_err(a_multi, a_single).max()
err_p = np.abs(p_multi[:, :3] - p_single[:, :3]).max()
err_v = np.abs(v_multi[:, :3] - v_single[:, :3]).max()
ok = slice_ok and err_a <= tol and err_p <= (1e-5 if prec == nbtest.FP32 else 1e-9) and err_v <= (1e-5 if prec == nbtest.FP32 else 1e-9)
ok = ok and np.linalg.norm(mom[:3]) < 1e-6 * np.abs(pm[:, 3:4] * v[:, :3]).sum()
# energy: FP64 in both precisions -> the sliced sum equals the single-GPU one to summation-order round-off of the state
etol = 1e-6 if prec == nbtest.FP32 else 1e-11
err_e = abs(e_multi[0] - e_single[0]) / abs(e_single[0])
ok = ok and err_e <= etol and abs(e_multi[1] + e_multi[2] - e_multi[0]) <= 1e-12 * abs(e_multi[0]) and np.isfinite(e_mid).all()
# snapshot rows = (x,y,z,vx,vy,vz) in the caller's order, (time, energy), (mass, 0): every rank gets the whole system
ok = ok and np.array_equal(snap6[:, :3], p_multi[:, :3]) and np.array_equal(snap6[:, 3:], v_multi[:, :3])
ok = ok and sys2[0] == 0.16 and sys2[1] == e_multi[0] and np.array_equal(stat2[:, 1], pm[:, 3]) and np.array_equal(stat2[:, 0], np.arange(n))
if n <= 40000 and rank == 0:
    oracle = nbtest.Oracle()
    err_o = nbtest.rel_err(a_multi, oracle.forces(pm, force)).max()
    ok = ok and err_o <= (1.5e-5 if prec == nbtest.FP32 else 1e-12)
    print(f"rank0: vs oracle {err_o:.2e}")
print(f"rank {rank}/{world}: accel multi-vs-single {err_a:.2e}  pos {err_p:.2e}  vel {err_v:.2e}  |P|={np.linalg.norm(mom[:3]):.2e}  dE {err_e:.2e}  ok={ok}", flush=True)
flag = torch.tensor([1 if ok else 0], device="cuda")
dist.all_reduce(flag, op=dist.ReduceOp.MIN)
dist.destroy_process_group()
sys.exit(0 if int(flag.item()) == 1 else 1)
