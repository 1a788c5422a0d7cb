#!/bin/bash
# round 2: full-sum scorer with eight epilogue warps, two-pass form with an add tree
mkdir -p gpurun_out
O=gpurun_out/q_summary.txt
: > $O
run() { echo "## $*" >> $O; env "$@" >> $O 2>&1; }
run GMM_TC_EPI_WARPS=8 python profiles/time_case.py cfg3 sum 10000 5 0
run python profiles/time_case.py cfg3 sum 10000 5 0
run GMM_TC_EPI_WARPS=8 python profiles/time_case.py cfg1 sum 32000 5 0
run python profiles/time_case.py cfg1 sum 32000 5 0
GMM_TC_EPI_WARPS=8 timeout 600 python -m pytest tests/test_gpu_parity.py -q -m gpu -x -k "sum or posterior or golden_float or random" > gpurun_out/q_parity.log 2>&1; echo "parity(8 warps) rc=$?" >> $O; tail -3 gpurun_out/q_parity.log >> $O
run GMM_TC_DEBUG=1 GMM_TC_EPI_WARPS=8 python profiles/time_case.py cfg3 sum 10000 2 0
cat $O
