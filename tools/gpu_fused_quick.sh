#!/bin/bash
# quick check of a builder change: builder tests, A/B timing fused vs one-pass, launch list
TAG=${1:-r04c}; O=gpurun_out/$TAG; mkdir -p $O
timeout 600 python -m pytest tests/test_parity_gpu.py -m gpu -q -x -k "bvh_invariants or c5_full_size or edge or c1_obj or async_rebuild or outgrows or instance_update" > $O/pytest_build.log 2>&1; echo "rc=$?" >> $O/pytest_build.log; tail -6 $O/pytest_build.log
for F in 0 1; do
  RTB_BUILD_FUSED=$F timeout 200 python tools/build_prof.py 1582 3161 4 > $O/build_prof_fused$F.log 2>&1; tail -2 $O/build_prof_fused$F.log
  RTB_BUILD_FUSED=$F timeout 200 python tools/build_prof.py 500 1000 4 > $O/build_prof_1m_fused$F.log 2>&1; tail -1 $O/build_prof_1m_fused$F.log
done
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file $O/build_launches.csv python tools/build_prof.py 1582 3161 2 > $O/ncu_build.log 2>&1
python - <<PY
import csv
rows=[r for r in csv.reader(open('$O/build_launches.csv')) if len(r)>5]
for i,r in enumerate(rows):
    if 'Kernel Name' in r: hdr=r; rows=rows[i+1:]; break
ki=hdr.index('Kernel Name'); vi=hdr.index('Metric Value'); ii=hdr.index('ID')
seq=[(r[ki].split('(')[0][:40], float(r[vi].replace(',',''))/1000) for r in rows if r[ii].isdigit()]
a=[i for i,s in enumerate(seq) if 'k_scene_bounds' in s[0]][-1]
print([ (n,round(t,1)) for n,t in seq[a:] if 'rs_' not in n], round(sum(t for n,t in seq[a:]),1))
PY
