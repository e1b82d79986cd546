O=gpurun_out/r01t; mkdir -p $O
for v in 1000 500 700 1400 2000 3000; do PRE="sah_node=$v" timeout 120 python tools/tune.py c3 spec_thr=4 2>>$O/err.log | sed "s/^/sah_node=$v /" ; done > $O/tune_c3_sah_node.log
cat $O/tune_c3_sah_node.log
for v in 1000 500 2000; do PRE="sah_node=$v" timeout 120 python tools/tune.py c2 spec_thr=4 2>>$O/err.log | sed "s/^/sah_node=$v /" ; done > $O/tune_c2_sah_node.log
cat $O/tune_c2_sah_node.log
