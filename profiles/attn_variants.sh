#!/bin/bash
# kernel-experiment driver: runs the attention microbench on every variant library under build_variants/
for f in /root/repo/build_variants/libmab200_*.so; do
  echo "=== $(basename $f .so | sed s/libmab200_//)"
  MAB_LIB=$f python profiles/attn_only.py 2>&1 | tail -4 | awk '{print $2, $3, $4, $(NF-1)}'
done
