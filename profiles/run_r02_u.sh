#!/bin/bash
# round 2: float kernel, row pieces written by the TMA engine (2-D tensor-map stores)
mkdir -p gpurun_out
O=gpurun_out/u_summary.txt
: > $O
run() { echo "## $*" >> $O; env "$@" >> $O 2>&1; }
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_fp8_split.py tests/test_gpu_gather.py tests/test_gpu_align.py -q -m gpu -x > gpurun_out/u_parity.log 2>&1; echo "parity rc=$?" >> $O; tail -5 gpurun_out/u_parity.log >> $O
run python profiles/time_case.py cfg1 max 32000 20 0
run GMM_TC_STORES=vector python profiles/time_case.py cfg1 max 32000 20 0
run python profiles/time_case.py cfg1 max 8000 20 0
run GMM_TC_STORES=vector python profiles/time_case.py cfg1 max 8000 20 0
run python profiles/time_case.py cfg3 max 10000 10 0
run GMM_TC_DEBUG=1 python profiles/time_case.py cfg1 max 32000 2 0
cat $O
