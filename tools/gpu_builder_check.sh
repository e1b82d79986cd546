#!/bin/bash
# builder change check: tests touching the builder / textures, build timing, ncu launch list of the builder, C5 bench
TAG=${1:-run}; O=gpurun_out/$TAG; mkdir -p $O
timeout 600 python -m pytest tests -m gpu -q -x > $O/pytest_all.log 2>&1; echo "pytest_all rc=$?" >> $O/pytest_all.log; tail -4 $O/pytest_all.log
timeout 200 python tools/build_prof.py 1582 3161 5 > $O/build_prof.log 2>&1; cat $O/build_prof.log
timeout 200 python tools/build_prof.py 500 1000 5 > $O/build_prof_1m.log 2>&1; cat $O/build_prof_1m.log
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file $O/build_launches.csv python tools/build_prof.py 1582 3161 2 > $O/ncu_build.log 2>&1
timeout 120 python bench.py --config c5 --no-cpu-baseline --no-ref-cuda > $O/bench_c5.json 2> $O/bench_c5.err; cat $O/bench_c5.json
