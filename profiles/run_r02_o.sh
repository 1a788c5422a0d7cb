#!/bin/bash
# round 2: full-sum epilogue with the unpredicated full-segment path
mkdir -p gpurun_out
O=gpurun_out/o_summary.txt
: > $O
run() { echo "## $*" >> $O; env "$@" >> $O 2>&1; }
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_rasr_plugin.py tests/test_gpu_accumulate.py -q -m gpu -x > gpurun_out/o_parity.log 2>&1; echo "parity rc=$?" >> $O; tail -4 gpurun_out/o_parity.log >> $O
run python profiles/time_case.py cfg3 sum 10000 5 0
run python profiles/time_case.py cfg1 sum 32000 5 0
run python profiles/time_case.py cfg3 sum 10000 5
cat $O
