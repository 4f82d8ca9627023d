#!/bin/bash
# round-2 GPU session 16 (N GPUs): C5 alone after the NCCL warm-up in nb_comm_init (headline shortened to N = 262144)
N=$1; mkdir -p gpurun_out
timeout 800 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2956$N \
  bench.py --gpus $N --particles 262144 --steps 3 --warmup 3 --configs C5 --no-cpu-baseline > gpurun_out/r02_bench_${N}gpu_c5.json 2> gpurun_out/r02_bench_${N}gpu_c5.err; echo "bench rc=$?"
python - <<PY
import json
d = json.loads(open("gpurun_out/r02_bench_${N}gpu_c5.json").read().strip().splitlines()[-1])
print("headline 262144", round(d["value"], 1), round(d["ms_per_step"], 3), d["parity"]["ok"])
for c in d.get("configs", []):
    print("   ", c["id"], round(c["value"], 1), round(c["ms_per_step"], 3), "frac", round(c["roofline"]["frac"], 3), "parity", c["parity"]["max_rel_err"], [round(x, 1) for x in c["per_rank"]["force_ms_by_rank"]], c["clocks"])
PY
