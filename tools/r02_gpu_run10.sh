#!/bin/bash
# round-2 GPU session 10 (one GPU): ncu --set full of the far + near kernels at C2, 1M and C4 with the owner-per-CTA near kernel
mkdir -p gpurun_out
B="python bench.py --steps 1 --warmup 0 --configs none --no-e2e --no-cpu-baseline --no-parity"
NCU="ncu --set full --clock-control none --import-source on -k regex:force_ --launch-skip 2 -c 2 -f"
$NCU -o gpurun_out/r02_ncu_c2 $B --particles 65536 --init random > gpurun_out/ncu_c2.log 2>&1; echo "ncu c2 rc=$?"
$NCU -o gpurun_out/r02_ncu_1M $B > gpurun_out/ncu_1M.log 2>&1; echo "ncu 1M rc=$?"
$NCU -o gpurun_out/r02_ncu_c4_fp32 $B --particles 262144 --init galaxy --force ADAPTIVE_MUTUAL > gpurun_out/ncu_c4.log 2>&1; echo "ncu c4 rc=$?"
