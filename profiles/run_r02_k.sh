#!/bin/bash
# round 2: phase timers of the float kernel (max and sum) for DESIGN.md
mkdir -p gpurun_out
O=gpurun_out/k_summary.txt
: > $O
run() { echo "## $*" >> $O; env "$@" >> $O 2>&1; }
run GMM_TC_DEBUG=1 python profiles/time_case.py cfg3 sum 10000 2
run GMM_TC_DEBUG=1 python profiles/time_case.py cfg3 max 10000 2
run GMM_TC_DEBUG=1 python profiles/time_case.py cfg1 max 32000 2
run python profiles/time_case.py cfg3 sum 10000 5
run python profiles/time_case.py cfg1 sum 32000 5
cat $O
