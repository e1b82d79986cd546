#!/bin/bash
# Round-end visit: GPU suite, smoke, the bench lines of every config + the reference arm, full ncu capture of the builder kernels.
TAG=${1:-run}; O=gpurun_out/$TAG; mkdir -p $O
timeout 600 python -m pytest tests -m gpu -q -x > $O/pytest_all.log 2>&1; echo "pytest_all rc=$?" >> $O/pytest_all.log; tail -3 $O/pytest_all.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "smoke rc=$?" >> $O/smoke.log; tail -4 $O/smoke.log
timeout 400 python bench.py > $O/bench_c3_default.json 2> $O/bench_c3_default.err; cut -c1-400 $O/bench_c3_default.json
timeout 200 python bench.py --config c5 --no-cpu-baseline --no-ref-cuda > $O/bench_c5.json 2> $O/bench_c5.err; cut -c1-200 $O/bench_c5.json
timeout 200 python bench.py --config c2 --no-cpu-baseline --no-ref-cuda > $O/bench_c2.json 2> $O/bench_c2.err; cut -c1-200 $O/bench_c2.json
timeout 200 python bench.py --config c1 --no-cpu-baseline --no-ref-cuda > $O/bench_c1.json 2> $O/bench_c1.err; cut -c1-200 $O/bench_c1.json
timeout 300 python bench.py --config c4 --no-cpu-baseline --no-ref-cuda --steps 3 > $O/bench_c4.json 2> $O/bench_c4.err; cut -c1-200 $O/bench_c4.json
timeout 200 python tools/build_prof.py 1582 3161 3 > $O/build_prof.log 2>&1; cat $O/build_prof.log
timeout 300 ncu --set full --clock-control none --import-source on -k regex:"k_fit|k_collapse|k_scene_bounds|k_emit_tris|k_hierarchy" -c 6 -o $O/prof_build python tools/build_prof.py 1582 3161 1 > $O/ncu_full.log 2>&1; tail -2 $O/ncu_full.log
timeout 300 python bench.py --impl reference > $O/bench_c3_reference_arm.json 2> $O/bench_ref.err; cut -c1-300 $O/bench_c3_reference_arm.json
