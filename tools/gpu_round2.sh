#!/bin/bash
# One GPU visit for the instance-box change: the tests next to it, an A/B of the box budget on C3/C4, a bench line, the whole
# GPU suite, then the other configs while time remains.   usage (under gpurun): bash tools/gpu_round2.sh <tag>
TAG=${1:-run}; O=gpurun_out/$TAG; mkdir -p $O
timeout 300 python -m pytest tests -m gpu -q -k "rebuild or instance_boxes or instances_textures or empty_and_degenerate" > $O/pytest_new.log 2>&1
echo "pytest_new rc=$?" >> $O/pytest_new.log; tail -5 $O/pytest_new.log
timeout 240 python tools/tune.py c3 inst_cut=1,256,1,256,64,16 > $O/tune_c3_inst_cut.log 2> $O/tune_c3.err; cat $O/tune_c3_inst_cut.log
timeout 300 python bench.py --no-ref-cuda > $O/bench_c3.json 2> $O/bench_c3.err; cat $O/bench_c3.json
timeout 600 python -m pytest tests -m gpu -q -x > $O/pytest_all.log 2>&1
echo "pytest_all rc=$?" >> $O/pytest_all.log; tail -5 $O/pytest_all.log
timeout 120 python bench.py --config c5 --no-cpu-baseline --no-ref-cuda > $O/bench_c5.json 2> $O/bench_c5.err; cat $O/bench_c5.json
timeout 120 python bench.py --config c2 --no-cpu-baseline --no-ref-cuda > $O/bench_c2.json 2> $O/bench_c2.err; cat $O/bench_c2.json
timeout 200 python tools/tune.py c4 inst_cut=1,256 > $O/tune_c4_inst_cut.log 2> $O/tune_c4.err; cat $O/tune_c4_inst_cut.log
