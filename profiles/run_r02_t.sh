#!/bin/bash
mkdir -p gpurun_out
O=gpurun_out/t_summary.txt
: > $O
run() { echo "## $*" >> $O; env "$@" >> $O 2>&1; }
run python profiles/time_case.py cfg1 max 32000 20 0
run GMM_TC_DRAIN=bulk python profiles/time_case.py cfg1 max 32000 20 0
run python profiles/time_case.py cfg3 max 10000 10 0
run GMM_TC_DRAIN=bulk python profiles/time_case.py cfg3 max 10000 10 0
run GMM_TC_DEBUG=1 GMM_TC_DRAIN=bulk python profiles/time_case.py cfg1 max 32000 2 0
cat $O
