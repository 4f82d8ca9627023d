#!/bin/bash
# tools/scale_run.sh — strong-scaling runs on one box: bench.py at N=1M for 1/2/4/8 GPUs and N=16M for 8 GPUs.
# Usage (GPU box with 8 GPUs): bash tools/scale_run.sh [ngpu_max]
MAXG=${1:-8}
mkdir -p gpurun_out
for n in 1 2 4 8; do
  [ $n -gt $MAXG ] && break
  if [ $n = 1 ]; then
    python bench.py --gpus 1 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/scale_1M_$n.json 2> gpurun_out/scale_1M_$n.err
  else
    timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2954$n \
      bench.py --gpus $n --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/scale_1M_$n.json 2> gpurun_out/scale_1M_$n.err
  fi
  tail -n 2 gpurun_out/scale_1M_$n.err
  python -c "import json; d=json.loads(open('gpurun_out/scale_1M_$n.json').read().strip().splitlines()[-1]); print('N=1M gpus', d['n_gpus'], 'Gint/s', round(d['value'],1), 'ms/step', round(d['ms_per_step'],2), 'frac', round(d['roofline']['frac'],3), 'e2e', round(d['e2e']['value'],1))"
done
if [ $MAXG -ge 8 ]; then
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29559 \
    bench.py --gpus 8 --particles 16777216 --steps 2 --warmup 1 --no-cpu-baseline --no-e2e > gpurun_out/scale_16M_8.json 2> gpurun_out/scale_16M_8.err
  tail -n 2 gpurun_out/scale_16M_8.err
  python -c "import json; d=json.loads(open('gpurun_out/scale_16M_8.json').read().strip().splitlines()[-1]); print('N=16M gpus', d['n_gpus'], 'Gint/s', round(d['value'],1), 'ms/step', round(d['ms_per_step'],2), 'frac', round(d['roofline']['frac'],3))"
fi
