#!/bin/bash
# One GPU visit: parity tests, bench lines, launch list and a full ncu capture of the queue tracer.
# usage (under gpurun): bash tools/gpu_round.sh <tag> [skip-launches] [count]
TAG=${1:-run}; SKIP=${2:-30}; CNT=${3:-3}
O=gpurun_out/$TAG; mkdir -p $O
python -m pytest tests -m gpu -x -q > $O/pytest.log 2>&1; echo "pytest rc=$?" >> $O/pytest.log
tail -3 $O/pytest.log
timeout 300 python bench.py --no-cpu-baseline --no-ref-cuda > $O/bench_c3.json 2> $O/bench_c3.err; cat $O/bench_c3.json
timeout 300 python bench.py --config c2 --no-cpu-baseline --no-ref-cuda > $O/bench_c2.json 2> $O/bench_c2.err; cat $O/bench_c2.json
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-ref-cuda"
$CMD > $O/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_c3.csv $CMD > $O/ncu_launch.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:"k_mega|k_trace_queue" -s $SKIP -c $CNT -o $O/prof_trace $CMD > $O/ncu_full.log 2>&1
