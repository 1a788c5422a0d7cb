#!/bin/bash
# round 2, fourth GPU pass: e4m3 correction operands (whole suite), A-ring depth experiments (64-byte swizzle)
mkdir -p gpurun_out
O=gpurun_out/d_timing.txt
: > $O
timeout 900 python -m pytest tests -q -m gpu -x -s > gpurun_out/d_suite.log 2>&1; echo "suite rc=$?" >> $O
grep -a "last mean_scale" gpurun_out/d_suite.log >> $O
tail -4 gpurun_out/d_suite.log >> $O
run() { echo "## $*" >> $O; env "$@" >> $O 2>&1; }
run python profiles/time_case.py cfg1 max 32000 20 0
run GMM_TC_SWIZZLE=64 python profiles/time_case.py cfg1 max 32000 20 0
run GMM_TC_DEBUG=1 GMM_TC_SWIZZLE=64 python profiles/time_case.py cfg1 max 32000 2 0
run GMM_TC_STORES=staged GMM_TC_SWIZZLE=64 python profiles/time_case.py cfg1 max 32000 20 0
run GMM_TC_SPLIT=fp16 python profiles/time_case.py cfg1 max 32000 20 0
run GMM_TC_SPLIT=fp16 GMM_TC_SWIZZLE=64 python profiles/time_case.py cfg1 max 32000 20 0
run python profiles/time_case.py cfg3 max 10000 10 0
run GMM_TC_SWIZZLE=64 python profiles/time_case.py cfg3 max 10000 10 0
run GMM_TC_DEBUG=1 GMM_TC_SWIZZLE=64 python profiles/time_case.py cfg3 max 10000 2 0
run python profiles/time_case.py cfg3 sum 10000 10 0
run GMM_TC_SWIZZLE=64 python profiles/time_case.py cfg3 sum 10000 10 0
cat $O
