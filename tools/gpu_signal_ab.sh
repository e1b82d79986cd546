# A/B of the completion signal of the peer-memory frame exchange (flag words vs the 4-byte all-reduce), C3, N = 2, 4, 8.
# usage (8-GPU box): tools/gpu_retry.sh --gpus 8 --timeout 1200 -- 'bash tools/gpu_signal_ab.sh'
O=gpurun_out/signal_r04; mkdir -p $O
echo "(2-GPU flag test passed in the previous call)"
for N in 2 4 8; do
  REPS="nccl peer"; [ $N = 8 ] && REPS="nccl peer peer2 nccl2"
  for S0 in $REPS; do S=${S0%2}
    timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500 + N)) bench.py --gpus $N --steps 20 --warmup 5 --completion $S --no-cpu-baseline --no-ref-cuda > $O/c3_n${N}_$S0.json 2> $O/c3_n${N}_$S0.err
    python -c "
import json
d=json.loads(open('$O/c3_n${N}_$S0.json').read().strip().splitlines()[-1]); r=d['roofline']
print($N, '$S', 'ms/step', round(d['ms_per_step'],3), 'Mrays/s', round(d['value'],1), 'e2e', round(d['e2e']['value'],1), 'kernel', r['kernel_ms_ranks'], 'exchange', round(r['exchange_ms'],3))" | tee -a $O/summary.txt
  done
done
