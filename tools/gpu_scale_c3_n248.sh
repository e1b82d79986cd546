O=gpurun_out/scale_r03c; mkdir -p $O
for N in 2 4 8; do
  timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500 + N)) bench.py --gpus $N --no-cpu-baseline --no-ref-cuda > $O/c3_n$N.json 2> $O/c3_n$N.err
  python -c "
import json
d=json.loads(open('$O/c3_n$N.json').read().strip().splitlines()[-1]); r=d['roofline']
print($N, 'ms/step', round(d['ms_per_step'],3), 'Mrays/s', round(d['value'],1), 'e2e', round(d['e2e']['value'],1), 'kernel', r['kernel_ms_ranks'], 'exchange', round(r['exchange_ms'],3), 'variant', r['kernel_variant'])"
done
