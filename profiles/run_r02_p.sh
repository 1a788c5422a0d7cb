#!/bin/bash
mkdir -p gpurun_out
O=gpurun_out/p_summary.txt
: > $O
run() { echo "## $*" >> $O; env "$@" >> $O 2>&1; }
run GMM_TC_DEBUG=1 python profiles/time_case.py cfg1 max 32000 3 0
run GMM_TC_DEBUG=1 python profiles/time_case.py cfg1 max 8000 3 0
run GMM_TC_DEBUG=1 python profiles/time_case.py cfg3 max 10000 3 0
cat $O
