#!/bin/bash
# round-2 GPU session 9 (eight GPUs): the default line at 8 GPUs (all configs incl. C5 = 16M), then 4 GPUs without C5
mkdir -p gpurun_out
nvidia-smi -L | wc -l
( time timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29548 \
  bench.py --gpus 8 > gpurun_out/r02_bench_8gpu.json 2> gpurun_out/r02_bench_8gpu.err ) 2> gpurun_out/r02_bench_8gpu.time; echo "bench 8 rc=$?"; tail -n 4 gpurun_out/r02_bench_8gpu.time; grep -v "^\*\|OMP_NUM" gpurun_out/r02_bench_8gpu.err | tail -n 8
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29544 \
  bench.py --gpus 4 --configs C2,C4 --no-cpu-baseline > gpurun_out/r02_bench_4gpu_c.json 2> gpurun_out/r02_bench_4gpu_c.err; echo "bench 4 rc=$?"
python - <<'PY'
import json
for f in ("r02_bench_8gpu", "r02_bench_4gpu_c"):
    d = json.loads(open(f"gpurun_out/{f}.json").read().strip().splitlines()[-1])
    print(f, "headline", round(d["value"], 1), round(d["ms_per_step"], 3), "e2e", round(d["e2e"]["value"], 1), round(d["e2e"]["ms_per_call"], 1), "parity", d["parity"]["max_rel_err"], d["parity"]["multi_vs_single_pos"], d["parity"]["ok"])
    print("   slice work", [round(x, 2) for x in d["per_rank"]["slice_work_ms_by_rank"]], "span", [round(x, 2) for x in d["per_rank"]["force_ms_by_rank"]], "mhz", d["per_rank"]["sm_mhz_by_rank"])
    for c in d.get("configs", []):
        print("   ", c["id"], round(c["value"], 1), round(c["ms_per_step"], 3), "frac", round(c["roofline"]["frac"], 3), "parity", c["parity"]["max_rel_err"], c["parity"]["multi_vs_single_pos"], "slice work", c["per_rank"]["slice_work_ms_by_rank"] and [round(x, 3) for x in c["per_rank"]["slice_work_ms_by_rank"]])
PY
