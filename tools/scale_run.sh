#!/bin/bash
# 1/2/4/8-GPU bench lines of the default config on one box (run under `gpurun --gpus 8`); output in gpurun_out/<tag>/
TAG=${1:-scale}; O=gpurun_out/$TAG; mkdir -p $O
python bench.py --gpus 1 --no-cpu-baseline --no-ref-cuda > $O/scale_n1.json 2> $O/scale_n1.err
for N in 2 4 8; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500 + N)) bench.py --gpus $N --no-cpu-baseline --no-ref-cuda > $O/scale_n$N.json 2> $O/scale_n$N.err
done
for N in 1 2 4 8; do python -c "import json,sys; d=json.loads(open('$O/scale_n$N.json').read().strip().splitlines()[-1]); print($N, round(d['ms_per_step'],3), round(d['value'],1), round(d['e2e']['value'],1))"; done
