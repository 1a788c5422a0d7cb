#!/bin/bash
mkdir -p gpurun_out
O=gpurun_out/j_summary.txt
: > $O
run() { echo "## $*" >> $O; env "$@" >> $O 2>&1; }
for s in 1 3 7 9 21 32 63; do
run GMM_QTC_SPLIT=$s python profiles/time_case.py cfg2 quant 8000 20 0
done
run GMM_QTC_DEBUG=1 GMM_QTC_SPLIT=63 python profiles/time_case.py cfg2 quant 8000 2 0
run GMM_QTC_DEBUG=1 GMM_QTC_SPLIT=21 python profiles/time_case.py cfg2 quant 8000 2 0
for s in 11 50 100 250 500; do
run GMM_QTC_SPLIT=$s python profiles/time_case.py cfg2 quant 64000 10 0
done
grep -v "epilogue warp" $O
