#!/bin/bash
# Pins and head-to-head on the box: the reference's OWN GPU drivers (unmodified sources rebuilt for sm_100a by
# `make -C oracle refgpu`, oracle/_ref/) run on the headline configurations next to this repo's drop-in executables.
# Full stdout of every run is kept under gpurun_out/pins/ ; scripts/pins_to_golden.py turns the reference's outputs
# into tests/golden/ref_gpu_runs.json.
#   usage: scripts/gpu_pin_reference.sh [D]        (D = number of GPUs, default 1)
set -u
D=${1:-1}
REF=$PWD/oracle/_ref
OURS=$PWD/dist-gpu-accelerated-tree-search_b200/host
OUT=$PWD/gpurun_out/pins
mkdir -p $OUT && cd $OUT
nvidia-smi -L > nvidia_smi_L.txt 2>&1
nvidia-smi --query-gpu=name,driver_version,clocks.max.sm --format=csv >> nvidia_smi_L.txt 2>&1
# a throw-away process first so that every timed binary sees a warm driver
$OURS/nqueens_multigpu_cuda.out -N 8 -D 1 > warm.txt 2>&1

run() { # tag, binary, args...
  local tag=$1; shift
  local t0=$(date +%s.%N)
  echo "\$ $(basename $1) ${@:2}" > $tag.txt
  timeout 1500 "$@" >> $tag.txt 2>&1
  echo "rc=$?" >> $tag.txt
  local t1=$(date +%s.%N)
  awk -v a=$t0 -v b=$t1 'BEGIN{printf "process_wall_s=%.4f\n", b-a}' >> $tag.txt
}
CFGS=("ta014_lb1:-i 14 -l 1" "ta021_lb2:-i 21 -l 2" "ta024_lb2:-i 24 -l 2")
# every reference run first: they are the pins
for cfg in "${CFGS[@]}"; do
  tag=${cfg%%:*}; args=${cfg#*:}
  run ref_${tag}_D$D $REF/ref_pfsp_multigpu_cuda.out $args -u 1 -D $D -C 0
done
run ref_nq17_D$D $REF/ref_nqueens_multigpu_cuda.out -N 17 -g 1 -D $D
for cfg in "${CFGS[@]}"; do
  tag=${cfg%%:*}; args=${cfg#*:}
  run ours_${tag}_D$D $OURS/pfsp_multigpu_cuda.out $args -u 1 -D $D -C 0
done
run ours_nq17_D$D $OURS/nqueens_multigpu_cuda.out -N 17 -g 1 -D $D
rm -f multigpu.csv singlegpu.csv
tail -n 12 ref_*.txt ours_*.txt
