#!/bin/bash
# round-2 GPU session 15 (N GPUs, N = first argument): final state — the default bench line at N GPUs (all configs; "nocfg5" as the
# second argument leaves C5 out), and with "mgpu" as the third argument the NCCL pytest cases with their worker log
N=$1; mkdir -p gpurun_out
CFG="all"; [ "$2" = "nocfg5" ] && CFG="C2,C4,C4-fp64mode,fp64-pairwise"
if [ "$3" = "mgpu" ]; then
  python -m pytest tests/test_multigpu.py -m gpu -v -s > gpurun_out/r02_mgpu_pytest.log 2>&1; echo "mgpu pytest rc=$?"; grep -E "rank [01]/|passed|failed" gpurun_out/r02_mgpu_pytest.log
fi
( time timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2954$N \
  bench.py --gpus $N --configs $CFG > gpurun_out/r02_bench_${N}gpu.json 2> gpurun_out/r02_bench_${N}gpu.err ) 2> gpurun_out/r02_bench_${N}gpu.time; echo "bench $N rc=$?"; tail -n 3 gpurun_out/r02_bench_${N}gpu.time
python - <<PY
import json
d = json.loads(open("gpurun_out/r02_bench_${N}gpu.json").read().strip().splitlines()[-1])
print("headline", round(d["value"], 1), round(d["ms_per_step"], 3), "e2e", round(d["e2e"]["value"], 1), round(d["e2e"]["ms_per_call"], 1), "parity", d["parity"]["max_rel_err"], d["parity"]["multi_vs_single_pos"], d["parity"]["ok"])
print("   slice work", [round(x, 2) for x in d["per_rank"]["slice_work_ms_by_rank"]], "span", [round(x, 2) for x in d["per_rank"]["force_ms_by_rank"]])
for c in d.get("configs", []):
    print("   ", c["id"], round(c["value"], 1), round(c["ms_per_step"], 3), "frac", round(c["roofline"]["frac"], 3), "parity", c["parity"]["max_rel_err"], c["parity"]["multi_vs_single_pos"], [round(x, 1) for x in c["per_rank"]["force_ms_by_rank"]])
PY
