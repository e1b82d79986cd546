#!/bin/bash
# Round-2 multi-GPU evidence on one 8xB200 box (run under `gpurun --gpus 8`): C3 at 1/2/4/8 GPUs, C4 and C5 at 8 GPUs (the configs
# BASELINE.json defines there), the C++ drop-in on 8 devices in ONE process (RTB_DEVICES=8) against its single-device frame.
TAG=${1:-scale_r03}; O=gpurun_out/$TAG; mkdir -p $O
run() { # N config extra...
  N=$1; C=$2; shift 2
  if [ $N -eq 1 ]; then timeout 400 python bench.py --gpus 1 --config $C --no-cpu-baseline --no-ref-cuda "$@" > $O/${C}_n$N.json 2> $O/${C}_n$N.err
  else timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500 + N)) bench.py --gpus $N --config $C --no-cpu-baseline --no-ref-cuda "$@" > $O/${C}_n$N.json 2> $O/${C}_n$N.err; fi
  python - <<PY
import json
try:
    d=json.loads(open('$O/${C}_n$N.json').read().strip().splitlines()[-1]); r=d.get('roofline') or {}
    print('$C', $N, 'ms/step', round(d['ms_per_step'],3), 'Mrays/s', round(d['value'],1), 'e2e', round(d['e2e']['value'],1), 'kernel ms ranks', r.get('kernel_ms_ranks'), 'variant', r.get('kernel_variant'))
except Exception as e: print('$C', $N, 'FAILED', e)
PY
}
for N in 1 2 4 8; do run $N c3; done
run 8 c4 --steps 3
run 8 c5
run 8 c2
# C++ drop-in, one process, 8 devices
python - <<'PY'
import sys; sys.path.insert(0,'.')
from cuda_raytracing_b200 import synth
v,f=synth.bumpy_sphere_mesh(200,400); synth.write_obj('/tmp/m.obj',v,f)
PY
for n in 1 8; do RTB_DEVICES=$n cuda_raytracing_b200/cpp/demo /tmp/m.obj /tmp/o$n.ppm 20 2 3840 2160 > $O/demo_dev$n.log 2>&1; tail -2 $O/demo_dev$n.log; done
cmp /tmp/o1.ppm /tmp/o8.ppm && echo "C++ demo: 8-device frame == 1-device frame (byte-identical)" | tee $O/demo_cmp.txt
