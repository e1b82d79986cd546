#!/bin/bash
# Round-2 evidence in one GPU visit: bench lines of every config, launch list + ncu --set full captures of the kernels that matter
# (k_megaq on C3, k_mega on a 1/8 share, k_megaq<PATH> on C4 at 2 spp), builder launch list.  usage (under gpurun): bash tools/gpu_r03.sh <tag>
TAG=${1:-r03}; O=gpurun_out/$TAG; mkdir -p $O
timeout 600 python bench.py > $O/bench_c3_default.json 2> $O/bench_c3_default.err; cut -c1-300 $O/bench_c3_default.json
timeout 300 python bench.py --impl reference --steps 3 > $O/bench_c3_reference_arm.json 2> $O/bench_ref.err; cut -c1-300 $O/bench_c3_reference_arm.json
for c in c1 c2 c5; do timeout 300 python bench.py --config $c --no-cpu-baseline --no-ref-cuda > $O/bench_$c.json 2> $O/bench_$c.err; cut -c1-200 $O/bench_$c.json; done
timeout 400 python bench.py --config c4 --no-cpu-baseline --no-ref-cuda --steps 3 > $O/bench_c4.json 2> $O/bench_c4.err; cut -c1-200 $O/bench_c4.json
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-ref-cuda"
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_c3.csv $CMD > $O/ncu_launch.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_megaq -s 3 -c 1 -o $O/prof_megaq_c3 $CMD > $O/ncu_megaq.log 2>&1
ncu --set full --clock-control none --import-source on -k k_mega -s 3 -c 1 -o $O/prof_mega_share8 python tools/tune.py c3 tile_count=8 fused=1 > $O/ncu_mega8.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_megaq -s 2 -c 1 -o $O/prof_megaq_c4 python tools/tune.py c4 fused=2 > $O/ncu_megaq_c4.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file $O/build_launches.csv python tools/build_prof.py 1582 3161 2 > $O/ncu_build.log 2>&1
python tools/tune.py c3 tile_count=8 fused=1,2 cost_block=3,2,1,0 > $O/tune_share8.log 2>&1; cat $O/tune_share8.log
python tools/tune.py c3 tile_count=4 fused=1,2 cost_block=3,1,0 > $O/tune_share4.log 2>&1; cat $O/tune_share4.log
