#!/bin/bash
# round-2 GPU session 6 (one GPU): near field fused into the far kernel's warps (A/B against the two-kernel form), triangle
# energy kernel, snapshot ring of nb_run_simulation, persistent render scratch
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r02_pytest_gpu6.log 2>&1; echo "pytest rc=$?"; tail -n 15 gpurun_out/r02_pytest_gpu6.log
python bench.py --configs C2,C4 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r02_bench6_fused.json 2> gpurun_out/r02_bench6_fused.err; echo "bench fused rc=$?"; tail -n 3 gpurun_out/r02_bench6_fused.err
python bench.py --configs C2,C4 --steps 5 --warmup 3 --no-cpu-baseline --no-e2e --opt fuse_near=0 > gpurun_out/r02_bench6_split.json 2> gpurun_out/r02_bench6_split.err; echo "bench split rc=$?"
python - <<'PY'
import json
for f in ("r02_bench6_fused", "r02_bench6_split"):
    d = json.loads(open(f"gpurun_out/{f}.json").read().strip().splitlines()[-1])
    print(f, "headline", round(d["value"], 1), round(d["ms_per_step"], 3), "frac", round(d["roofline"]["frac"], 4), "parity", d["parity"]["max_rel_err"], "e2e", d.get("e2e", {}).get("value"))
    for c in d.get("configs", []):
        print("   ", c["id"], round(c["value"], 1), round(c["ms_per_step"], 3), "frac", round(c["roofline"]["frac"], 3), "parity", c["parity"]["max_rel_err"])
PY
python tools/sim_timing.py 2>&1 | tail -5
