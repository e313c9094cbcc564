#!/bin/bash
# head-to-head on the box: the reference's own GPU drivers rebuilt for sm_100a (oracle/_ref, built by `make -C oracle
# refgpu` where /root/reference exists) against this repo's drop-in executables, same flags
set -u
REF=$PWD/oracle/_ref
OURS=$PWD/dist-gpu-accelerated-tree-search_b200/host
mkdir -p gpurun_out/refgpu && cd gpurun_out/refgpu
for args in "-i 29 -l 2" "-i 30 -l 2" "-i 22 -l 2" "-i 14 -l 1" "-i 30 -l 0"; do
  echo "### pfsp $args -D 1 -C 0"
  timeout 300 $REF/ref_pfsp_multigpu_cuda.out $args -D 1 -C 0 2>&1 | grep -A4 "^Size of the explored tree" | tail -4 | tr '\n' ' '; echo " <- reference (sm_100a rebuild)"
  timeout 300 $OURS/pfsp_multigpu_cuda.out $args -D 1 -C 0 2>&1 | grep -A4 "^Size of the explored tree" | tail -4 | tr '\n' ' '; echo " <- this repo"
  [ -x $REF/ref_pfsp_multigpu_with_bbgpu.out ] && { timeout 300 $REF/ref_pfsp_multigpu_with_bbgpu.out $args -D 1 -C 0 2>&1 | grep -A4 "^Size of the explored tree" | tail -4 | tr '\n' ' '; echo " <- unmodified reference driver on this repo's evaluate_gpu (drop-in boundary)"; }
done
for args in "-N 15" "-N 16"; do
  echo "### nqueens $args -D 1"
  timeout 300 $REF/ref_nqueens_multigpu_cuda.out $args -D 1 2>&1 | tail -4 | head -3 | tr '\n' ' '; echo " <- reference (sm_100a rebuild)"
  timeout 300 $OURS/nqueens_multigpu_cuda.out $args -D 1 2>&1 | tail -4 | head -3 | tr '\n' ' '; echo " <- this repo"
done
