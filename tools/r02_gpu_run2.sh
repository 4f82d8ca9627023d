#!/bin/bash
# round-2 GPU session 2 (one GPU): new near kernel (128-thread / 64-register CTAs beside the far kernel), far kernel at 112
# registers, FP64 kernel with 4 i-particles per thread; A/B against a library whose far kernel keeps 128 registers
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q -s > gpurun_out/r02_pytest_gpu2.log 2>&1; echo "pytest rc=$?"; tail -n 3 gpurun_out/r02_pytest_gpu2.log
grep -E "all-particle|nb_force_noalloc" gpurun_out/r02_pytest_gpu2.log
python bench.py --configs C2,C4,C4-fp64mode,fp64-pairwise --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r02_bench2.json 2> gpurun_out/r02_bench2.err; echo "bench rc=$?"; tail -n 5 gpurun_out/r02_bench2.err
NB_LIB_PATH=$PWD/n-body-simulation_b200/libnbody_b200_far128.so python bench.py --configs C2,C4 --steps 5 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/r02_bench2_far128.json 2> gpurun_out/r02_bench2_far128.err; echo "bench far128 rc=$?"
python - <<'PY'
import json
for f in ("r02_bench2", "r02_bench2_far128"):
    d = json.loads(open(f"gpurun_out/{f}.json").read().strip().splitlines()[-1])
    print(f, "headline", round(d["value"], 1), round(d["ms_per_step"], 3), "near span", round(d["per_rank"]["near_kernel_span_ms"]["max"], 2), "parity", d["parity"]["max_rel_err"])
    for c in d.get("configs", []):
        print("   ", c["id"], round(c["value"], 1), round(c["ms_per_step"], 3), "frac", round(c["roofline"]["frac"], 3), "near", round(c["per_rank"]["near_kernel_span_ms"]["max"], 3), "parity", c["parity"]["max_rel_err"])
PY
