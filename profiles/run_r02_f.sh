#!/bin/bash
# round 2, sixth GPU pass: vector store path of the quantized kernel
mkdir -p gpurun_out
O=gpurun_out/f_timing.txt
: > $O
run() { echo "## $*" >> $O; env "$@" >> $O 2>&1; }
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_rasr_plugin.py tests/test_zz_no_alloc.py -q -m gpu -x > gpurun_out/f_parity.log 2>&1; echo "parity rc=$?" >> $O; tail -15 gpurun_out/f_parity.log >> $O
run python profiles/time_case.py cfg2 quant 8000 20 0
run python profiles/time_case.py cfg2 quantscore 8000 20 0
run python profiles/time_case.py cfg2 quant 64000 10 0
run GMM_QTC_EPI_WARPS=8 python profiles/time_case.py cfg2 quant 8000 20 0
run GMM_QTC_EPI_WARPS=8 python profiles/time_case.py cfg2 quantscore 8000 20 0
run GMM_QTC_STORES=scalar python profiles/time_case.py cfg2 quant 8000 20 0
run GMM_QTC_DEBUG=1 python profiles/time_case.py cfg2 quant 8000 1 0
run GMM_QTC_DEBUG=1 python profiles/time_case.py cfg2 quantscore 8000 1 0
run GMM_QTC_DEBUG=1 GMM_QTC_EPI_WARPS=8 python profiles/time_case.py cfg2 quant 8000 1 0
cat $O
