#!/bin/bash
# round 2, two GPUs: multimem.st (NVSwitch multicast) against unicast peer stores for the push of score columns
mkdir -p gpurun_out
O=gpurun_out/mm_summary.txt
: > $O
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/ubench_multimem profiles/ubench_multimem.cu -lcuda >> $O 2>&1
nvidia-smi topo -m >> $O 2>&1
timeout 120 /tmp/ubench_multimem 8192 8192 >> $O 2>&1; echo "rc=$?" >> $O
timeout 120 /tmp/ubench_multimem 1024 8192 >> $O 2>&1; echo "rc=$?" >> $O
cat $O
