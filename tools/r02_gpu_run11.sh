#!/bin/bash
# round-2 GPU session 11 (one GPU): near kernel with parallel sub-tile classification and prefetched staging
mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "acceleration or full_size or step_parity or all_particles or plummer" > gpurun_out/r02_pytest_gpu11.log 2>&1; echo "pytest rc=$?"; tail -n 3 gpurun_out/r02_pytest_gpu11.log
python bench.py --configs C2,C4 --steps 5 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/r02_bench11.json 2> gpurun_out/r02_bench11.err; echo "bench rc=$?"; tail -n 3 gpurun_out/r02_bench11.err
python - <<'PY'
import json
d = json.loads(open("gpurun_out/r02_bench11.json").read().strip().splitlines()[-1])
print("headline", round(d["value"], 1), round(d["ms_per_step"], 3), "frac", round(d["roofline"]["frac"], 4), "parity", d["parity"]["max_rel_err"])
for c in d.get("configs", []):
    print("   ", c["id"], round(c["value"], 1), round(c["ms_per_step"], 3), "frac", round(c["roofline"]["frac"], 3), "parity", c["parity"]["max_rel_err"])
PY
for NP in "65536 --init random" "1048576 --init spiral" "262144 --init galaxy --force ADAPTIVE_MUTUAL"; do
ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/l.csv python bench.py --steps 2 --warmup 1 --configs none --no-e2e --no-cpu-baseline --no-parity --particles $NP > gpurun_out/ncu_l.log 2>&1; echo "ncu launches $NP rc=$?"
grep -E "near64|r4_kernel" gpurun_out/l.csv | tail -2 | awk -F'","' '{print $5, $9, $NF}'
done
