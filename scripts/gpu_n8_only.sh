#!/bin/bash
set -u
mkdir -p gpurun_out
bash scripts/gpu_scaling.sh "8" "pfsp:21:2 nq:17"
cd gpurun_out && ../dist-gpu-accelerated-tree-search_b200/host/pfsp_multigpu_cuda.out -i 21 -l 2 -D 8 -C 0 -M 1000000 | tail -6 | head -5; ../dist-gpu-accelerated-tree-search_b200/host/nqueens_multigpu_cuda.out -N 17 -D 8 -M 2000000 | tail -5 | head -4
