#!/bin/bash
# fused fit + chunk-local collapse: sanitizer on the invariants test, builder tests, A/B timing, launch list, C5 / C3 lines
TAG=${1:-r04b}; O=gpurun_out/$TAG; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -q -x > $O/pytest_all.log 2>&1; echo "pytest_all rc=$?" >> $O/pytest_all.log; tail -6 $O/pytest_all.log
for F in 0 1; do
  RTB_BUILD_FUSED=$F timeout 200 python tools/build_prof.py 1582 3161 4 > $O/build_prof_fused$F.log 2>&1; cat $O/build_prof_fused$F.log
  RTB_BUILD_FUSED=$F timeout 200 python tools/build_prof.py 500 1000 4 > $O/build_prof_1m_fused$F.log 2>&1; tail -2 $O/build_prof_1m_fused$F.log
done
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file $O/build_launches.csv python tools/build_prof.py 1582 3161 2 > $O/ncu_build.log 2>&1
timeout 120 python bench.py --config c5 --no-cpu-baseline --no-ref-cuda > $O/bench_c5.json 2> $O/bench_c5.err; cat $O/bench_c5.json
timeout 120 python bench.py --no-cpu-baseline --no-ref-cuda > $O/bench_c3.json 2> $O/bench_c3.err; cat $O/bench_c3.json
timeout 120 python bench.py --config c2 --no-cpu-baseline --no-ref-cuda > $O/bench_c2.json 2> $O/bench_c2.err; cut -c1-200 $O/bench_c2.json
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
