#!/bin/bash
# round 2, four GPUs: push kernel shape (loads in flight, CTAs) where the exchange dominates
mkdir -p gpurun_out
for T in 8192 1024; do
  echo "## N=4 M=65536x64 T=$T" >> gpurun_out/sweep4.txt
  US="1 4" CS="128 296" bash profiles/sweep_push.sh 4 $T 65536 >> gpurun_out/sweep4.txt 2>&1
done
cat gpurun_out/sweep4.txt
