#!/bin/bash
# A/B of prebuilt library variants: bash scripts/gpu_variants.sh min2 min3
PK=dist-gpu-accelerated-tree-search_b200
cp $PK/libbbgpu.so /tmp/libbbgpu_orig.so
for v in "$@"; do
  cp $PK/variants/libbbgpu_$v.so $PK/libbbgpu.so
  echo "=== variant $v"
  timeout 300 python scripts/gpu_quick.py 2>&1 | sed -n '2,4p;7p'
done
cp /tmp/libbbgpu_orig.so $PK/libbbgpu.so
