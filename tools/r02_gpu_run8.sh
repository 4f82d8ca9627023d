#!/bin/bash
# round-2 GPU session 8 (one GPU): near-split heuristic; FP64 kernel unroll variants
mkdir -p gpurun_out
python bench.py --configs C2,C4,fp64-pairwise --steps 5 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/r02_bench8.json 2> gpurun_out/r02_bench8.err; echo "bench rc=$?"; tail -n 3 gpurun_out/r02_bench8.err
for U in 1 4; do NB_LIB_PATH=$PWD/n-body-simulation_b200/libnbody_b200_u$U.so python bench.py --particles 262144 --precision fp64 --configs none --steps 8 --warmup 3 --no-cpu-baseline --no-e2e --no-parity > gpurun_out/r02_bench8_u$U.json 2>/dev/null; done
python - <<'PY'
import json
d = json.loads(open("gpurun_out/r02_bench8.json").read().strip().splitlines()[-1])
print("headline", round(d["value"], 1), round(d["ms_per_step"], 3), "frac", round(d["roofline"]["frac"], 4), "parity", d["parity"]["max_rel_err"], "near span", d["per_rank"]["near_kernel_span_ms"]["max"])
for c in d.get("configs", []):
    print("   ", c["id"], round(c["value"], 1), round(c["ms_per_step"], 3), "frac", round(c["roofline"]["frac"], 3), "parity", c["parity"]["max_rel_err"])
for U in (1, 4):
    d = json.loads(open(f"gpurun_out/r02_bench8_u{U}.json").read().strip().splitlines()[-1])
    print("fp64 pairwise 262144 unroll", U, round(d["value"], 1), round(d["ms_per_step"], 3))
PY
for NP in "65536 --init random" "1048576 --init spiral" "262144 --init galaxy --force ADAPTIVE_MUTUAL"; do
ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/l.csv python bench.py --steps 2 --warmup 1 --configs none --no-e2e --no-cpu-baseline --no-parity --particles $NP > gpurun_out/ncu_l.log 2>&1; echo "ncu launches $NP rc=$?"
grep -E "near64|r4_kernel" gpurun_out/l.csv | tail -2 | awk -F'","' '{print $5, $9, $NF}'
done
