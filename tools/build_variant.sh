#!/bin/bash
# Build an A/B variant of librtb200.so with extra nvcc flags into tools/variants/ (git-ignored; travels to the GPU box).
# usage: tools/build_variant.sh <name.so> <flags...>;  then RTB_VARIANT=<name.so> python tools/tune.py ...
set -e
NAME=$1; shift
ROOT=$(cd "$(dirname "$0")/.." && pwd)
TMP=$(mktemp -d); mkdir -p $ROOT/tools/variants
cd $ROOT/cuda_raytracing_b200/csrc
FL="-gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -Xcompiler -fPIC,-ffp-contract=off,-fvisibility=hidden"
for f in rtb_api rtb_build rtb_render rtb_wavefront rtb_mega; do /usr/local/cuda/bin/nvcc $FL "$@" -c $f.cu -o $TMP/$f.o & done; wait
/usr/local/cuda/bin/nvcc $FL -x cu -c rtb_obj.cpp -o $TMP/rtb_obj.o
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -shared -o $ROOT/tools/variants/$NAME $TMP/*.o
rm -rf $TMP; ls -la $ROOT/tools/variants/$NAME
