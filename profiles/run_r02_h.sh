#!/bin/bash
# round 2, eighth GPU pass: whole GPU suite after the plugin rewrite, the gather time-out handling and batch-int
mkdir -p gpurun_out
O=gpurun_out/h_summary.txt
: > $O
timeout 1500 python -m pytest tests -q -m gpu -x > gpurun_out/h_suite.log 2>&1; echo "suite rc=$?" >> $O
tail -25 gpurun_out/h_suite.log >> $O
timeout 600 python -m pytest tests/test_gpu_rasr_plugin.py -q -m gpu -x -s 2>&1 | grep -E "PASS|FAIL|SKIP|TIMING|passed|failed" | tail -60 >> $O
cat $O
