#!/bin/bash
# round-2 GPU session 13 (one GPU): final state — full GPU test suite, smoke, reference arm, the default bench line (all configs),
# launch lists and ncu --set full captures of the hot kernels, FILE_CSV loop timing
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q -s > gpurun_out/r02_pytest_gpu_final.log 2>&1; echo "pytest rc=$?"; tail -n 3 gpurun_out/r02_pytest_gpu_final.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r02_bench_reference_arm.json 2>/dev/null; echo "ref arm rc=$?"
( time python bench.py > gpurun_out/r02_bench_final_1gpu.json 2> gpurun_out/r02_bench_final_1gpu.err ) 2> gpurun_out/r02_bench_final_1gpu.time; echo "bench rc=$?"; tail -n 3 gpurun_out/r02_bench_final_1gpu.time
python tools/sim_timing.py 2>&1 | tail -4
B="python bench.py --steps 1 --warmup 0 --configs none --no-e2e --no-cpu-baseline --no-parity"
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_ncu_launches_1M.csv python bench.py --steps 2 --warmup 1 --configs none --no-e2e --no-cpu-baseline --no-parity > gpurun_out/ncu_l.log 2>&1; echo "ncu launches rc=$?"
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_ncu_launches_c2.csv python bench.py --steps 2 --warmup 1 --configs none --no-e2e --no-cpu-baseline --no-parity --particles 65536 --init random > gpurun_out/ncu_l.log 2>&1
NCU="ncu --set full --clock-control none --import-source on -k regex:force_ --launch-skip 2 -c 2 -f"
$NCU -o gpurun_out/r02_ncu_1M $B > gpurun_out/ncu_1M.log 2>&1; echo "ncu 1M rc=$?"
$NCU -o gpurun_out/r02_ncu_c2 $B --particles 65536 --init random > gpurun_out/ncu_c2.log 2>&1; echo "ncu c2 rc=$?"
$NCU -o gpurun_out/r02_ncu_c4_fp32 $B --particles 262144 --init galaxy --force ADAPTIVE_MUTUAL > gpurun_out/ncu_c4.log 2>&1; echo "ncu c4 rc=$?"
ncu --set full --clock-control none --import-source on -k regex:force_ --launch-skip 1 -c 1 -f -o gpurun_out/r02_ncu_fp64_pairwise $B --particles 262144 --init spiral --precision fp64 > gpurun_out/ncu_fp64p.log 2>&1; echo "ncu fp64p rc=$?"
ncu --set full --clock-control none --import-source on -k regex:potential -c 1 -f -o gpurun_out/r02_ncu_energy python tools/sim_timing.py 65536 > gpurun_out/ncu_energy.log 2>&1; echo "ncu energy rc=$?"
