#!/bin/bash
# Builds libguava_rasterizer_b200.so variants that differ only in the backward compositor's launch bound
# (GUAVA_BWD_MIN_CTAS) into guava_renderer_b200/lib/variants/ for an A/B run on the GPU box (tools/experiments/gpu_bwd_variants.sh).
set -eu
cd "$(dirname "$0")/../.."   # needs tools/experiments/r01_v40_v41_compositor_variants.patch applied (git apply)
python guava_renderer_b200/build.py > /dev/null
L=guava_renderer_b200/lib; V=$L/variants; mkdir -p $V
cp $L/libguava_rasterizer_b200.so $V/lib_bwd12.so
for n in "$@"; do   # 13, 14: launch bound only; pN: packed alpha chain (GUAVA_BWD_PACKED=1) at bound N; qN: + packed moments / sums (=2)
  D="-DGUAVA_BWD_MIN_CTAS=${n#p}"; [ "${n#p}" != "$n" ] && D="$D -DGUAVA_BWD_PACKED=1"
  [ "${n#q}" != "$n" ] && D="-DGUAVA_BWD_MIN_CTAS=${n#q} -DGUAVA_BWD_PACKED=2"
  nvcc -c guava_renderer_b200/csrc/render_backward.cu -o $V/render_backward_$n.o $D \
    -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -Xcompiler -O2 \
    -Iinclude -Iguava_renderer_b200/csrc -Xptxas -v 2>&1 | grep -A1 "render_backward" | grep "Used\|spill" || true
  objs=$(ls $L/obj/*.o | grep -v render_backward)
  nvcc -shared -o $V/lib_bwd$n.so $objs $V/render_backward_$n.o -gencode arch=compute_100a,code=sm_100a
  rm $V/render_backward_$n.o
done
ls -la $V
