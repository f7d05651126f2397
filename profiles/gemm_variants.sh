#!/bin/bash
# kernel-experiment driver: GEMM parity tests + shape microbench on every variant library under build_variants/
cd "$(dirname "$0")/.."
for f in build_variants/libmab200_*.so; do
  echo "=== $(basename $f .so | sed s/libmab200_//)"
  MAB_LIB=$PWD/$f timeout 300 python -m pytest tests/test_gpu_gemm.py -x -q -m gpu 2>&1 | tail -1
  MAB_LIB=$PWD/$f timeout 120 python profiles/gemm_shapes.py 2>&1 | awk '{print $1, $2, $3, $(NF-3), $(NF-2), $(NF-1)}'
done
