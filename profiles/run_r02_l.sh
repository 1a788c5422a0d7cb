#!/bin/bash
# round 2: float kernel row-piece stores (padded rows, ld = 4504)
mkdir -p gpurun_out
O=gpurun_out/l_summary.txt
: > $O
run() { echo "## $*" >> $O; env "$@" >> $O 2>&1; }
run python profiles/time_case.py cfg1 max 32000 20 0
run GMM_TC_STORES=lane python profiles/time_case.py cfg1 max 32000 20 0
run GMM_TC_STORES=staged python profiles/time_case.py cfg1 max 32000 20 0
run python profiles/time_case.py cfg1 max 32000 20
run python profiles/time_case.py cfg1 max 8000 20 0
run GMM_TC_STORES=lane python profiles/time_case.py cfg1 max 8000 20 0
run GMM_TC_GROUPS=2 python profiles/time_case.py cfg1 max 32000 20 0
run GMM_TC_DEBUG=1 python profiles/time_case.py cfg1 max 32000 2 0
run GMM_TC_DEBUG=1 GMM_TC_STORES=lane python profiles/time_case.py cfg1 max 32000 2 0
cat $O
