#!/bin/bash
set -u
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q -k "multi_gpu" 2>&1 | tail -2
bash scripts/gpu_scaling.sh "1 2 4 8" "pfsp:21:2 nq:17"
cd gpurun_out && ../dist-gpu-accelerated-tree-search_b200/host/pfsp_multigpu_cuda.out -i 21 -l 2 -D 8 -C 0 -M 1000000 | tail -12; ../dist-gpu-accelerated-tree-search_b200/host/nqueens_multigpu_cuda.out -N 17 -D 8 -M 2000000 | tail -6; cat multigpu.csv | cut -c1-300
