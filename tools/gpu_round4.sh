O=gpurun_out/r02a; mkdir -p $O
for v in "" col10.so col12.so; do
  echo "== variant '$v'" >> $O/collapse_occ.log
  RTB_VARIANT=$v timeout 120 python tools/build_prof.py 1582 3161 4 >> $O/collapse_occ.log 2>&1
  RTB_VARIANT=$v timeout 120 python tools/build_prof.py 500 1000 4 >> $O/collapse_occ.log 2>&1
done
grep "==\|rebuild 2" $O/collapse_occ.log
timeout 600 python -m pytest tests -m gpu -q -x > $O/pytest_all.log 2>&1; echo "pytest_all rc=$?" >> $O/pytest_all.log; tail -3 $O/pytest_all.log
timeout 120 python bench.py --config c5 --no-cpu-baseline --no-ref-cuda > $O/bench_c5.json 2> $O/bench_c5.err; cut -c1-220 $O/bench_c5.json
timeout 200 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file $O/build_launches.csv python tools/build_prof.py 1582 3161 1 > $O/ncu_build.log 2>&1
