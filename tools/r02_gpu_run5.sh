#!/bin/bash
# round-2 GPU session 5 (one GPU): the full default bench line (all configs incl. C5 = 16M on one GPU), launch list and
# ncu captures of the re-shaped near kernel (C2 and 1M) and of the FP64 kernel with 4 i-particles per thread
mkdir -p gpurun_out
( time python bench.py > gpurun_out/r02_bench_default.json 2> gpurun_out/r02_bench_default.err ) 2> gpurun_out/r02_bench_default.time; echo "bench rc=$?"; cat gpurun_out/r02_bench_default.time; tail -n 12 gpurun_out/r02_bench_default.err
B="python bench.py --steps 1 --warmup 0 --configs none --no-e2e --no-cpu-baseline --no-parity"
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_ncu_launches_1M.csv python bench.py --steps 2 --warmup 1 --configs none --no-e2e --no-cpu-baseline --no-parity > gpurun_out/ncu_l.log 2>&1; echo "ncu launches rc=$?"
NCU="ncu --set full --clock-control none --import-source on -k regex:force_ --launch-skip 2 -c 2 -f"
$NCU -o gpurun_out/r02_ncu_1M $B > gpurun_out/ncu_1M.log 2>&1; echo "ncu 1M rc=$?"
$NCU -o gpurun_out/r02_ncu_c2 $B --particles 65536 --init random > gpurun_out/ncu_c2.log 2>&1; echo "ncu c2 rc=$?"
ncu --set full --clock-control none --import-source on -k regex:force_ --launch-skip 1 -c 1 -f -o gpurun_out/r02_ncu_fp64_pairwise $B --particles 262144 --init spiral --precision fp64 > gpurun_out/ncu_fp64p.log 2>&1; echo "ncu fp64p rc=$?"
