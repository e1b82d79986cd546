#!/bin/bash
# ncu --set full of the fused builder kernels (10 M triangles)
TAG=${1:-r04e}; O=gpurun_out/$TAG; mkdir -p $O
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"k_fit_collapse|k_fit_top_fused|k_collapse" -c 3 -o $O/fit python tools/build_prof.py 1582 3161 1 > $O/ncu.log 2>&1
tail -3 $O/ncu.log
ncu -i $O/fit.ncu-rep --page raw --csv > $O/fit_raw.csv 2>/dev/null
ncu -i $O/fit.ncu-rep --page details > $O/fit_details.txt 2>/dev/null
ls -la $O
