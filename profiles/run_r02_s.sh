#!/bin/bash
# round 2: full-sum epilogue, grouped in-place exponentials against the serial chain
mkdir -p gpurun_out
O=gpurun_out/s_summary.txt
: > $O
run() { echo "## $*" >> $O; env "$@" >> $O 2>&1; }
run python profiles/time_case.py cfg3 sum 10000 5 0
run GMM_TC_SUM=serial python profiles/time_case.py cfg3 sum 10000 5 0
run python profiles/time_case.py cfg1 sum 32000 5 0
run GMM_TC_SUM=serial python profiles/time_case.py cfg1 sum 32000 5 0
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_rasr_plugin.py -q -m gpu -x > gpurun_out/s_parity.log 2>&1; echo "parity rc=$?" >> $O; tail -3 gpurun_out/s_parity.log >> $O
cat $O
