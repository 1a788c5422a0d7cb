#!/bin/bash
mkdir -p gpurun_out
O=gpurun_out/j_summary.txt
: > $O
run() { echo "## $*" >> $O; env "$@" >> $O 2>&1; }
run GMM_QTC_DEBUG=1 python profiles/time_case.py cfg2 quant 8000 3 0
grep -v "epilogue warp" $O
