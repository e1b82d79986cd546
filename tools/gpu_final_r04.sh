#!/bin/bash
# final evidence of the round: GPU suite, smoke, default bench line (C3 with CPU baseline and reference kernel), C5 / C2 lines, builder launch list
TAG=${1:-r04j}; O=gpurun_out/$TAG; mkdir -p $O
timeout 900 python -m pytest tests -m gpu -q -x > $O/pytest_all.log 2>&1; echo "pytest_all rc=$?" >> $O/pytest_all.log; tail -4 $O/pytest_all.log
timeout 200 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > $O/smoke.log 2>&1; tail -1 $O/smoke.log
timeout 400 python bench.py > $O/bench_default.json 2> $O/bench_default.err; cut -c1-300 $O/bench_default.json
timeout 120 python bench.py --config c5 --no-cpu-baseline --no-ref-cuda > $O/bench_c5.json 2> $O/bench_c5.err; cut -c1-200 $O/bench_c5.json
timeout 120 python bench.py --config c2 --no-cpu-baseline --no-ref-cuda > $O/bench_c2.json 2> $O/bench_c2.err; cut -c1-200 $O/bench_c2.json
timeout 200 python tools/build_prof.py 1582 3161 4 > $O/build_prof.log 2>&1; tail -2 $O/build_prof.log
timeout 200 python tools/build_prof.py 500 1000 4 > $O/build_prof_1m.log 2>&1; tail -1 $O/build_prof_1m.log
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file $O/build_launches.csv python tools/build_prof.py 1582 3161 2 > $O/ncu_build.log 2>&1
python - <<PY
import csv
rows=[r for r in csv.reader(open('$O/build_launches.csv')) if len(r)>5]
for i,r in enumerate(rows):
    if 'Kernel Name' in r: hdr=r; rows=rows[i+1:]; break
ki=hdr.index('Kernel Name'); vi=hdr.index('Metric Value'); ii=hdr.index('ID')
seq=[(r[ki].split('(')[0][:40], float(r[vi].replace(',',''))/1000) for r in rows if r[ii].isdigit()]
a=[i for i,s in enumerate(seq) if 'k_scene_bounds' in s[0]][-1]
agg={}
for n,t in seq[a:]: agg[n]=agg.get(n,0)+t
print({k:round(v,1) for k,v in agg.items()}, round(sum(agg.values()),1))
PY
