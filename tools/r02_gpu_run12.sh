#!/bin/bash
# round-2 GPU session 12 (one GPU): FP64 kernel shape variants; near-kernel split sweep at C2 / C4 / 1M
mkdir -p gpurun_out
B="python bench.py --configs none --no-cpu-baseline --no-e2e --no-parity"
python -m pytest tests -m gpu -x -q > gpurun_out/r02_pytest_gpu12.log 2>&1; echo "pytest rc=$?"; tail -n 4 gpurun_out/r02_pytest_gpu12.log
for V in "" _va _vb _vc; do
  NB_LIB_PATH=$PWD/n-body-simulation_b200/libnbody_b200$V.so $B --particles 262144 --precision fp64 --steps 8 --warmup 3 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('fp64 pairwise variant \"$V\"', round(d['value'],1), round(d['ms_per_step'],3))"
done
for S in 1 2 3 4 8 16; do
  $B --particles 65536 --init random --steps 100 --warmup 10 --opt near_splits=$S 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('C2 near_splits $S', round(d['value'],1), round(d['ms_per_step'],4))"
done
for S in 2 4 8 16; do
  $B --particles 262144 --init galaxy --force ADAPTIVE_MUTUAL --steps 10 --warmup 3 --opt near_splits=$S 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('C4 near_splits $S', round(d['value'],1), round(d['ms_per_step'],4))"
done
for S in 16 32 64; do
  $B --steps 3 --warmup 1 --opt near_splits=$S 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('1M near_splits $S', round(d['value'],1), round(d['ms_per_step'],4))"
done
