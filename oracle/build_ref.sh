#!/usr/bin/env bash
# oracle/build_ref.sh -- compile the reference's own sources, from where they lie under
# /root/reference, into oracle/_ref/ (git-ignored, travels to the GPU box).  TEST INFRASTRUCTURE.
#   libref_host.so       host C++ form  (g++ -O2 -ffp-contract=off, fake cudart)      patches P1+P2+P3
#   libref_cuda.so       reference CUDA kernel, sm_100a, P2 (32-bit Q_rsqrt) + P3 AOVs, -fmad as nvcc default
#   libref_cuda_asis.so  reference CUDA kernel exactly as shipped (UB Q_rsqrt on LP64), for the record
# No reference source is copied into the repo: the patched raycast.cu lives in a mktemp dir that is
# deleted on exit; everything else is #included from /root/reference directly.
set -euo pipefail
HERE="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
REF="${REF_DIR:-/root/reference/CudaRaytracer}"
OUT="$HERE/_ref"
if [ ! -d "$REF" ]; then echo "build_ref: $REF not present (GPU box?) -- keeping prebuilt $OUT"; exit 0; fi
mkdir -p "$OUT"
TMP="$(mktemp -d)"; trap 'rm -rf "$TMP"' EXIT
CXX=/usr/bin/g++
NVCC="${NVCC:-/usr/local/cuda/bin/nvcc}"

# P3: one statement after the accept block's last assignment (raycast.cu:125); P1: return by value (raycast.cu:21)
sed -e 's/hit_info.material = material;/hit_info.material = material; REF_AOV_RECORD(ray, index, mesh_idx, distance);/' \
    "$REF/raycast.cu" > "$TMP/raycast_p3.cu"
grep -q REF_AOV_RECORD "$TMP/raycast_p3.cu"
mkdir -p "$TMP/host" "$TMP/cuda" "$TMP/asis"
sed -e 's/__device__ HitInfo& cast_ray/__device__ HitInfo cast_ray/' "$TMP/raycast_p3.cu" > "$TMP/host/raycast_patched.cu"
grep -q '__device__ HitInfo cast_ray' "$TMP/host/raycast_patched.cu"
cp "$TMP/raycast_p3.cu" "$TMP/cuda/raycast_patched.cu"
cp "$REF/raycast.cu" "$TMP/asis/raycast_patched.cu"

echo "[build_ref] host form"
$CXX -std=c++17 -O2 -ffp-contract=off -w -fPIC -shared -pthread -DREF_HOST -DREF_P2 -DREF_P3 \
    -I "$TMP/host" -I "$HERE/shim" -I /usr/local/cuda/include -I "$REF" \
    -x c++ "$HERE/ref_harness.cpp" -o "$OUT/libref_host.so"

if [ "${REF_SKIP_CUDA:-0}" != "1" ]; then
  echo "[build_ref] CUDA form (P2+P3)"
  $NVCC -std=c++17 -O3 -gencode arch=compute_100a,code=sm_100a -lineinfo -w -Xcompiler -fPIC -shared -DREF_P2 -DREF_P3 \
      -I "$TMP/cuda" -I "$HERE/shim" -I "$REF" -x cu "$HERE/ref_harness.cpp" -o "$OUT/libref_cuda.so"
  echo "[build_ref] CUDA form (as shipped)"
  $NVCC -std=c++17 -O3 -gencode arch=compute_100a,code=sm_100a -lineinfo -w -Xcompiler -fPIC -shared \
      -I "$TMP/asis" -I "$HERE/shim" -I "$REF" -x cu "$HERE/ref_harness.cpp" -o "$OUT/libref_cuda_asis.so"
fi
ls -la "$OUT"
