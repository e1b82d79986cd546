#!/bin/bash
# 32-bit (11/11/10) vs 39-bit Morton keys at 10 M triangles: build time, trace quality (nodes per ray), C5 frame
TAG=${1:-r04i}; O=gpurun_out/$TAG; mkdir -p $O
timeout 400 python -m pytest tests/test_parity_gpu.py -m gpu -q -x -k "bvh_invariants or c5_full_size or c1_obj or edge" > $O/pytest_build.log 2>&1; echo "rc=$?" >> $O/pytest_build.log; tail -3 $O/pytest_build.log
for KB in 64 32; do
  RTB_OPTIONS=key_bits=$KB timeout 200 python tools/build_prof.py 1582 3161 4 > $O/build_prof_kb$KB.log 2>&1; tail -2 $O/build_prof_kb$KB.log
  RTB_OPTIONS=key_bits=$KB timeout 120 python bench.py --config c5 --no-cpu-baseline --no-ref-cuda > $O/bench_c5_kb$KB.json 2> $O/bench_c5_kb$KB.err
  python -c "
import json
d=json.loads(open('$O/bench_c5_kb$KB.json').read().strip().splitlines()[-1]); r=d['roofline']
print('key_bits', $KB, 'ms/step', round(d['ms_per_step'],3), 'kernel_ms', round(r['kernel_ms'],3), 'nodes/ray', round(r['nodes_per_ray'],3), 'tris/ray', round(r['tris_per_ray'],3), 'blas', d['details']['blas'])"
done
