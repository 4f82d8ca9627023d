#!/bin/bash
# round-2 GPU session 7 (one GPU): near kernel with one owner per 256-thread CTA; FILE_CSV loop against the round-1 library
mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "acceleration or full_size or step_parity" > gpurun_out/r02_pytest_gpu7.log 2>&1; echo "pytest rc=$?"; tail -n 3 gpurun_out/r02_pytest_gpu7.log
python bench.py --configs C2,C4 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r02_bench7.json 2> gpurun_out/r02_bench7.err; echo "bench rc=$?"; tail -n 3 gpurun_out/r02_bench7.err
python - <<'PY'
import json
for f in ("r02_bench7",):
    d = json.loads(open(f"gpurun_out/{f}.json").read().strip().splitlines()[-1])
    print(f, "headline", round(d["value"], 1), round(d["ms_per_step"], 3), "frac", round(d["roofline"]["frac"], 4), "parity", d["parity"]["max_rel_err"], "e2e", d.get("e2e", {}).get("value"))
    for c in d.get("configs", []):
        print("   ", c["id"], round(c["value"], 1), round(c["ms_per_step"], 3), "frac", round(c["roofline"]["frac"], 3), "parity", c["parity"]["max_rel_err"])
PY
echo "--- FILE_CSV loop, this library"; python tools/sim_timing.py 2>&1 | tail -5





for NP in "65536 --init random" "1048576 --init spiral" "262144 --init galaxy --force ADAPTIVE_MUTUAL"; do
ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/l.csv python bench.py --steps 2 --warmup 1 --configs none --no-e2e --no-cpu-baseline --no-parity --particles $NP > gpurun_out/ncu_l.log 2>&1; echo "ncu launches $NP rc=$?"
grep -E "near64|r4_kernel" gpurun_out/l.csv | tail -2 | awk -F'","' '{print $5, $9, $NF}'
done
echo "--- FILE_CSV loop, round-1 library"; NB_LIB_PATH=$PWD/n-body-simulation_b200/libnbody_b200_r01.so python tools/sim_timing.py 2>&1 | tail -5
