#!/bin/bash
# round-2 GPU session 3 (two GPUs): the NCCL path — pytest multi-GPU cases with their worker log, bench.py --gpus 2 with parity
mkdir -p gpurun_out
nvidia-smi -L
python -m pytest tests/test_multigpu.py -m gpu -v -s > gpurun_out/r02_mgpu_pytest.log 2>&1; echo "pytest rc=$?"; grep -E "^rank|passed|failed|PASSED|FAILED" gpurun_out/r02_mgpu_pytest.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29541 \
  bench.py --gpus 2 --steps 5 --warmup 3 --configs C2,C4,C4-fp64mode > gpurun_out/r02_bench_2gpu.json 2> gpurun_out/r02_bench_2gpu.err; echo "bench rc=$?"; tail -n 3 gpurun_out/r02_bench_2gpu.err
python - <<'PY'
import json
d = json.loads(open("gpurun_out/r02_bench_2gpu.json").read().strip().splitlines()[-1])
print("headline", round(d["value"], 1), round(d["ms_per_step"], 3), "e2e", round(d["e2e"]["value"], 1), round(d["e2e"]["ms_per_call"], 1), "parity", d["parity"])
print("per_rank", d["per_rank"])
for c in d.get("configs", []):
    print("   ", c["id"], round(c["value"], 1), round(c["ms_per_step"], 3), "frac", round(c["roofline"]["frac"], 3), "parity", c["parity"]["max_rel_err"], c["parity"]["multi_vs_single_pos"], c["per_rank"]["force_ms"])
PY
