#!/bin/bash
# round 2, fifth GPU pass: phase timings of the two tensor kernels (GMM_TC_DEBUG / GMM_QTC_DEBUG), 64-byte swizzle default
mkdir -p gpurun_out
O=gpurun_out/e_timing.txt
: > $O
run() { echo "## $*" >> $O; env "$@" >> $O 2>&1; }
timeout 300 python -m pytest tests/test_gpu_parity.py tests/test_gpu_subset.py tests/test_gpu_gather.py -q -m gpu -x > gpurun_out/e_parity.log 2>&1; echo "parity rc=$?" >> $O; tail -2 gpurun_out/e_parity.log >> $O
run python profiles/time_case.py cfg1 max 32000 20 0
run GMM_TC_DEBUG=1 python profiles/time_case.py cfg1 max 32000 1 0
run GMM_TC_DEBUG=1 GMM_TC_STORES=staged python profiles/time_case.py cfg1 max 32000 1 0
run GMM_TC_DEBUG=1 python profiles/time_case.py cfg3 max 10000 1 0
run GMM_QTC_DEBUG=1 python profiles/time_case.py cfg2 quant 8000 1 0
run GMM_QTC_DEBUG=1 python profiles/time_case.py cfg2 quantscore 8000 1 0
run GMM_QTC_DEBUG=1 GMM_QTC_EPI_WARPS=8 python profiles/time_case.py cfg2 quant 8000 1 0
run GMM_QTC_DEBUG=1 python profiles/time_case.py cfg2 quant 64000 1 0
cat $O
