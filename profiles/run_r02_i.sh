#!/bin/bash
# round 2, ninth GPU pass: suite after batch-int rule / gather changes; quantized kernel with two epilogue groups
mkdir -p gpurun_out
O=gpurun_out/i_summary.txt
: > $O
run() { echo "## $*" >> $O; env "$@" >> $O 2>&1; }
timeout 1500 python -m pytest tests -q -m gpu -x > gpurun_out/i_suite.log 2>&1; echo "suite rc=$?" >> $O
tail -8 gpurun_out/i_suite.log >> $O
timeout 600 python -m pytest tests/test_gpu_rasr_plugin.py -q -m gpu -x -s 2>&1 | grep -E "FAIL|SKIP|TIMING|passed|failed|batch-dia" | tail -30 >> $O
run python profiles/time_case.py cfg2 quant 8000 20 0
run python profiles/time_case.py cfg2 quantscore 8000 20 0
run python profiles/time_case.py cfg2 quant 64000 10 0
run GMM_QTC_GROUPS=1 python profiles/time_case.py cfg2 quant 8000 20 0
run GMM_QTC_EPI_WARPS=16 python profiles/time_case.py cfg2 quant 8000 20 0
run GMM_QTC_DEBUG=1 python profiles/time_case.py cfg2 quant 8000 1 0
run GMM_QTC_DEBUG=1 python profiles/time_case.py cfg2 quantscore 8000 1 0
cat $O
