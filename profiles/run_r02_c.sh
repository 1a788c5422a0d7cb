#!/bin/bash
# round 2, third GPU pass: quantized kernel with the constants inside the contraction, float kernel with direct row-piece
# stores -- whole suite, then timings.
mkdir -p gpurun_out
O=gpurun_out/c_timing.txt
: > $O
timeout 900 python -m pytest tests -q -m gpu -x > gpurun_out/c_suite.log 2>&1; echo "suite rc=$?" >> $O
tail -4 gpurun_out/c_suite.log >> $O
run() { echo "## $*" >> $O; env "$@" >> $O 2>&1; }
run python profiles/time_case.py cfg1 max 32000 20
run python profiles/time_case.py cfg1 max 32000 20 0
run GMM_TC_STORES=staged python profiles/time_case.py cfg1 max 32000 20 0
run GMM_TC_DEBUG=1 python profiles/time_case.py cfg1 max 32000 2 0
run python profiles/time_case.py cfg3 max 10000 10
run python profiles/time_case.py cfg3 max 10000 10 0
run python profiles/time_case.py cfg2 quant 8000 20
run python profiles/time_case.py cfg2 quant 8000 20 0
run python profiles/time_case.py cfg2 quantscore 8000 20
run python profiles/time_case.py cfg2 quant 64000 10
run GMM_QTC_EPI_WARPS=8 python profiles/time_case.py cfg2 quant 8000 20
run GMM_QTC_EPI_WARPS=8 python profiles/time_case.py cfg2 quantscore 8000 20
run GMM_QTC_GENERIC=1 python profiles/time_case.py cfg2 quant 8000 20
cat $O
