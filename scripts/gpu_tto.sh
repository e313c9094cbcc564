#!/bin/bash
mkdir -p gpurun_out/r2
for cfg in "14 1" "14 0" "14 2" "29 2" "22 2" "21 2" "24 2"; do
  timeout 150 python scripts/gpu_tto.py $cfg 262144 2>&1 | tail -1
done | tee gpurun_out/r2/time_to_optimum.txt
for cfg in "14 2" "29 2" "21 2"; do
  echo "reference GPU driver, -u 0: $cfg"; ( cd gpurun_out/r2 && timeout 300 ../../oracle/_ref/ref_pfsp_multigpu_cuda.out -i ${cfg% *} -l ${cfg#* } -u 0 -D 1 -C 0 2>&1 | tail -5 | head -4 | tr '\n' ' ' ); echo
done | tee -a gpurun_out/r2/time_to_optimum.txt
