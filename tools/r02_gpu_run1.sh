#!/bin/bash
# round-2 GPU session 1 (one GPU): tests, bench with the extra configs, ncu captures of the kernels that had no profile
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q -s > gpurun_out/r02_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -n 3 gpurun_out/r02_pytest_gpu.log
grep -E "all-particle|nb_force_noalloc" gpurun_out/r02_pytest_gpu.log
python bench.py --configs C2,C4,C4-fp64mode,fp64-pairwise --steps 5 --warmup 3 > gpurun_out/r02_bench1.json 2> gpurun_out/r02_bench1.err; echo "bench rc=$?"; tail -n 5 gpurun_out/r02_bench1.err
NCU="ncu --set full --clock-control none --import-source on -k regex:force_ --launch-skip 2 -c 2 -f"
B="python bench.py --steps 1 --warmup 0 --configs none --no-e2e --no-cpu-baseline --no-parity"
$NCU -o gpurun_out/r02_ncu_c2 $B --particles 65536 --init random > gpurun_out/ncu_c2.log 2>&1; echo "ncu c2 rc=$?"
$NCU -o gpurun_out/r02_ncu_c4_fp32 $B --particles 262144 --init galaxy --force ADAPTIVE_MUTUAL > gpurun_out/ncu_c4.log 2>&1; echo "ncu c4 rc=$?"
ncu --set full --clock-control none --import-source on -k regex:force_ --launch-skip 1 -c 1 -f -o gpurun_out/r02_ncu_fp64_pairwise $B --particles 262144 --init spiral --precision fp64 > gpurun_out/ncu_fp64p.log 2>&1; echo "ncu fp64p rc=$?"
ncu --set full --clock-control none --import-source on -k regex:force_ --launch-skip 1 -c 1 -f -o gpurun_out/r02_ncu_fp64_adaptive $B --particles 262144 --init galaxy --force ADAPTIVE_MUTUAL --precision fp64 > gpurun_out/ncu_fp64a.log 2>&1; echo "ncu fp64a rc=$?"
ls -la gpurun_out | head -30
