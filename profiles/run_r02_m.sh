#!/bin/bash
# round 2: whole suite with the final kernels, then timings
mkdir -p gpurun_out
O=gpurun_out/m_summary.txt
: > $O
run() { echo "## $*" >> $O; env "$@" >> $O 2>&1; }
timeout 1500 python -m pytest tests -q -m gpu -x > gpurun_out/m_suite.log 2>&1; echo "suite rc=$?" >> $O; tail -4 gpurun_out/m_suite.log >> $O
run python profiles/time_case.py cfg2 quant 8000 20 0
run python profiles/time_case.py cfg2 quantscore 8000 20 0
run python profiles/time_case.py cfg2 quant 64000 10 0
run python profiles/time_case.py cfg1 max 8000 20 0
run python profiles/time_case.py cfg1 max 32000 20 0
run python profiles/time_case.py cfg3 max 10000 10 0
run python profiles/time_case.py cfg3 sum 10000 5 0
run GMM_QTC_DEBUG=1 python profiles/time_case.py cfg2 quant 8000 2 0
grep -v "epilogue warp" $O
