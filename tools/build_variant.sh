#!/bin/bash
# Build an alternative libgsr_b200 with extra -D flags for A/B kernel experiments:
#   tools/build_variant.sh alt "-DRB_MIN_CTAS=4"   ->  gaussian-splatting-npu_b200/gsr_b200/libgsr_b200_alt.so
# Use with GSR_B200_LIB=<path> python bench.py ... (ctypes legs only).
set -e
name=$1; defs=$2
root=$(cd "$(dirname "$0")/.." && pwd)
src=$root/gaussian-splatting-npu_b200/csrc
out=$root/gaussian-splatting-npu_b200/build/var_$name
mkdir -p $out
for f in gsr_api preprocess_fwd binning radix_sort render_fwd render_bwd preprocess_bwd; do
  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC --expt-relaxed-constexpr $defs -c $src/$f.cu -o $out/$f.o &
done
wait
nvcc -shared -gencode arch=compute_100a,code=sm_100a -o $root/gaussian-splatting-npu_b200/gsr_b200/libgsr_b200_$name.so $out/*.o -cudart shared
echo built libgsr_b200_$name.so
