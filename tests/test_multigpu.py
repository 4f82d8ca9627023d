"""Multi-rank tests. CPU (gloo, world_size 2): the host-side partition + exchange logic the multi-GPU path implements,
replayed with numpy/oracle slices. GPU (needs >= 2 devices, skipped otherwise): tests/mgpu_worker.py under torchrun."""
import os
import subprocess
import sys

import numpy as np
import pytest

import nbtest

HERE = os.path.dirname(os.path.abspath(__file__))

GLOO_WORKER = r'''
import os, sys
import numpy as np
import torch, torch.distributed as dist
sys.path.insert(0, sys.argv[1])
import nbtest
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
dist.init_process_group("gloo")
lib, oracle = nbtest.Lib(), nbtest.Oracle()
n = 3000
pm, v = nbtest.gen_random(n, seed=5)
b, c = lib.partition_bounds(n, rank, world)           # the product's own partition arithmetic (host only, no GPU)
# every rank evaluates its own i-slice against ALL j (what each GPU does), then the slices are all-gathered
for method in (nbtest.PAIRWISE, nbtest.ADAPTIVE):
    mine = oracle.forces_subset(pm, method, np.arange(b, b + c))
    sizes = [lib.partition_bounds(n, r, world)[1] for r in range(world)]
    parts = [torch.zeros((s, 3), dtype=torch.float64) for s in sizes]
    dist.all_gather(parts, torch.from_numpy(mine)) if len(set(sizes)) == 1 else None
    if len(set(sizes)) != 1:                          # ragged last slice: gather by padding
        pad = max(sizes)
        buf = torch.zeros((pad, 3), dtype=torch.float64); buf[:c] = torch.from_numpy(mine)
        outs = [torch.zeros((pad, 3), dtype=torch.float64) for _ in range(world)]
        dist.all_gather(outs, buf)
        parts = [o[:s] for o, s in zip(outs, sizes)]
    full = torch.cat(parts).numpy()
    ref = oracle.forces(pm, method)
    assert nbtest.rel_err(full, ref).max() < 1e-13, nbtest.rel_err(full, ref).max()
# one distributed Velocity-Verlet step: slice-local drift/kick + all-gather of positions + all-reduce of the COM sums
dt = 0.01
msum = torch.tensor([(pm[b:b+c, 3:4] * v[b:b+c, :3]).sum(0).tolist() + [pm[b:b+c, 3].sum()]], dtype=torch.float64)
dist.all_reduce(msum)
vcm = (msum[0, :3] / msum[0, 3]).numpy()
a0 = oracle.forces(pm, nbtest.PAIRWISE)
pos = pm.copy(); vel = v.copy()
vel[b:b+c, :3] -= vcm
pos[b:b+c, :3] += vel[b:b+c, :3] * dt + 0.5 * a0[b:b+c] * dt * dt
sl = torch.zeros((max(lib.partition_bounds(n, r, world)[1] for r in range(world)), 3), dtype=torch.float64)
sl[:c] = torch.from_numpy(pos[b:b+c, :3])
outs = [torch.zeros_like(sl) for _ in range(world)]
dist.all_gather(outs, sl)
for r in range(world):
    rb, rc = lib.partition_bounds(n, r, world)
    pos[rb:rb+rc, :3] = outs[r][:rc].numpy()
a1 = oracle.forces_subset(pos, nbtest.PAIRWISE, np.arange(b, b + c))
vel[b:b+c, :3] += 0.5 * (a0[b:b+c] + a1) * dt
p_ref, v_ref = oracle.step(pm, v, nbtest.VV, nbtest.PAIRWISE, dt, 1)
assert np.abs(pos[:, :3] - p_ref[:, :3]).max() < 1e-12
assert np.abs(vel[b:b+c, :3] - v_ref[b:b+c, :3]).max() < 1e-12
# report points: every rank sums the kinetic energy and the pair potential of its own i-slice against ALL j, then one
# all-reduce of (K, sum m_i phi_i) — the collective nb_energy (csrc/nb_context.cu) replayed with numpy
G = 39.47841760435743
d = pm[b:b+c, None, :3] - pm[None, :, :3]
r = np.sqrt((d * d).sum(-1))
with np.errstate(divide="ignore"):
    inv = np.where(r > 0, 1.0 / r, 0.0)
phi = (inv * pm[None, :, 3]).sum(1)
part = torch.tensor([0.5 * (pm[b:b+c, 3] * (v[b:b+c, :3] ** 2).sum(1)).sum(), (pm[b:b+c, 3] * phi).sum()], dtype=torch.float64)
dist.all_reduce(part)
e_total = float(part[0]) - 0.5 * G * float(part[1])
e_ref = oracle.energy(pm, v)
assert abs(e_total - e_ref) <= 1e-12 * abs(e_ref), (e_total, e_ref)
dist.destroy_process_group()
print("gloo rank", rank, "ok")
'''


def test_partitioned_step_with_gloo_world_size_2(tmp_path):
    script = tmp_path / "gloo_worker.py"
    script.write_text(GLOO_WORKER)
    env = dict(os.environ, MASTER_ADDR="127.0.0.1", OMP_NUM_THREADS="2")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                        "--master-addr", "127.0.0.1", "--master-port", "29533", str(script), HERE],
                       capture_output=True, text=True, env=env, timeout=300)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-3000:]
    assert r.stdout.count("ok") == 2


@pytest.mark.gpu
@pytest.mark.parametrize("args", [["20000", "fp32", "pairwise"], ["9000", "fp64", "adaptive"], ["5000", "fp32", "adaptive"]],
                         ids=["fp32-pairwise", "fp64-adaptive", "fp32-adaptive-ragged"])
def test_sliced_context_matches_single_gpu(args):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs >= 2 GPUs")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                        "--master-addr", "127.0.0.1", "--master-port", "29534", os.path.join(HERE, "mgpu_worker.py")] + args,
                       capture_output=True, text=True, timeout=600)
    print("\n".join(l for l in r.stdout.splitlines() if l.startswith("rank")))      # shown with -s: kept as the NCCL-path log
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
