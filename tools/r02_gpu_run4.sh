#!/bin/bash
# round-2 GPU session 4 (four GPUs): sliced upload / slice download / remote pass on its own stream; per-rank times and clocks
mkdir -p gpurun_out
python -m pytest tests/test_multigpu.py -m gpu -v -s > gpurun_out/r02_mgpu_pytest_b.log 2>&1; echo "pytest rc=$?"; grep -E "rank [01]/|passed|failed" gpurun_out/r02_mgpu_pytest_b.log
for G in 4 2; do
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $G --master-addr 127.0.0.1 --master-port 2954$G \
  bench.py --gpus $G --steps 5 --warmup 3 --configs C2,C4 --no-cpu-baseline > gpurun_out/r02_bench_${G}gpu_b.json 2> gpurun_out/r02_bench_${G}gpu_b.err; echo "bench $G rc=$?"; tail -n 2 gpurun_out/r02_bench_${G}gpu_b.err
python - <<PY
import json
d = json.loads(open("gpurun_out/r02_bench_${G}gpu_b.json").read().strip().splitlines()[-1])
print("headline", round(d["value"], 1), round(d["ms_per_step"], 3), "e2e", round(d["e2e"]["value"], 1), round(d["e2e"]["ms_per_call"], 1), "parity", d["parity"]["max_rel_err"], d["parity"]["multi_vs_single_pos"], d["parity"]["ok"])
print("per_rank", d["per_rank"])
for c in d.get("configs", []):
    print("   ", c["id"], round(c["value"], 1), round(c["ms_per_step"], 3), "frac", round(c["roofline"]["frac"], 3), "parity", c["parity"]["max_rel_err"], c["parity"]["multi_vs_single_pos"], c["per_rank"]["force_ms_by_rank"])
PY
done
