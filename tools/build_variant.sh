#!/bin/bash
# Build an alternative libgsr_b200 with extra -D flags for A/B kernel experiments:
#   tools/build_variant.sh alt "-DRB_MIN_CTAS=4" [file.cu ...]  ->  gaussian-splatting-npu_b200/gsr_b200/libgsr_b200_alt.so
# Only the listed files (default: the two blend kernels) are recompiled with the flags; the other objects come from
# the regular build (gaussian-splatting-npu_b200/build/*.o).  Use with tools/ab_stage.py or GSR_B200_LIB=<path>.
set -e
name=$1; defs=$2; shift 2 || true
files=${@:-render_fwd.cu render_bwd.cu}
root=$(cd "$(dirname "$0")/.." && pwd)
src=$root/gaussian-splatting-npu_b200/csrc
obj=$root/gaussian-splatting-npu_b200/build
out=$obj/var_$name
mkdir -p $out
objs=""
for o in $obj/*.o; do
  b=$(basename $o .o)
  if echo " $files " | grep -q " $b.cu "; then
    nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC --expt-relaxed-constexpr $defs -c $src/$b.cu -o $out/$b.o &
    objs="$objs $out/$b.o"
  else
    objs="$objs $o"
  fi
done
wait
nvcc -shared -gencode arch=compute_100a,code=sm_100a -o $root/gaussian-splatting-npu_b200/gsr_b200/libgsr_b200_$name.so $objs -cudart shared
echo built libgsr_b200_$name.so
