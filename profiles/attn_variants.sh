#!/bin/bash
# kernel-experiment driver: attention parity test + microbench on every variant library under build_variants/
cd "$(dirname "$0")/.."
for f in build_variants/libmab200_*.so; do
  echo "=== $(basename $f .so | sed s/libmab200_//)"
  MAB_LIB=$PWD/$f timeout 300 python -m pytest tests/test_gpu_attention.py -x -q -m gpu 2>&1 | tail -1
  MAB_LIB=$PWD/$f timeout 120 python profiles/attn_only.py 2>&1 | tail -4 | awk '{print $2, $3, $4, $(NF-1)}'
done
