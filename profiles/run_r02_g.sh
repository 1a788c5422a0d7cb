#!/bin/bash
# round 2, seventh GPU pass: plugin check (8 + timing cases), first run of the changed bench.py, row-piece write ubench
mkdir -p gpurun_out
O=gpurun_out/g_summary.txt
: > $O
timeout 900 python -m pytest tests/test_gpu_rasr_plugin.py -q -m gpu -x -s > gpurun_out/g_plugin.log 2>&1; echo "plugin rc=$?" >> $O
grep -E "PASS|FAIL|TIMING|passed|failed" gpurun_out/g_plugin.log | tail -60 >> $O
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/g_bench.json 2> gpurun_out/g_bench.err; echo "bench rc=$?" >> $O
tail -5 gpurun_out/g_bench.err >> $O
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/ubench_rowpiece profiles/ubench_rowpiece.cu >> $O 2>&1
for T in 8000 64000; do /tmp/ubench_rowpiece $T 4501 >> gpurun_out/g_rowpiece.txt 2>&1; done
cat gpurun_out/g_rowpiece.txt >> $O
cat $O
head -c 3000 gpurun_out/g_bench.json
