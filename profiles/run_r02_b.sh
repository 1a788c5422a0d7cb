#!/bin/bash
# round 2, second GPU pass: float-kernel epilogue experiments (early accumulator release, 16 epilogue warps), MMA-thread
# wait shares, fresh ncu captures of the default kernels.
mkdir -p gpurun_out
O=gpurun_out/b_timing.txt
: > $O
run() { echo "## $*" >> $O; env "$@" >> $O 2>&1; }
for T in 32000 8000; do
run GMM_TC_DRAIN=chunked python profiles/time_case.py cfg1 max $T 20
run python profiles/time_case.py cfg1 max $T 20
run GMM_TC_EPI_WARPS=16 python profiles/time_case.py cfg1 max $T 20
done
run GMM_TC_DRAIN=chunked python profiles/time_case.py cfg3 max 10000 10
run python profiles/time_case.py cfg3 max 10000 10
run GMM_TC_EPI_WARPS=16 python profiles/time_case.py cfg3 max 10000 10
run GMM_TC_DEBUG=1 GMM_TC_DRAIN=chunked python profiles/time_case.py cfg1 max 32000 3
run GMM_TC_DEBUG=1 python profiles/time_case.py cfg1 max 32000 3
run GMM_TC_DEBUG=1 python profiles/time_case.py cfg3 max 10000 3
run python profiles/time_case.py cfg1 maxbest 32000 10
run python profiles/time_case.py cfg2 quant 8000 20
run python profiles/time_case.py cfg2 quant 64000 10
run GMM_QTC_EPI_WARPS=8 python profiles/time_case.py cfg2 quant 8000 20
timeout 300 python -m pytest tests/test_gpu_parity.py -q -m gpu -x > gpurun_out/b_parity.log 2>&1; echo "parity rc=$?" >> $O
# ncu: full captures of the default kernels (plain run first)
for c in "cfg1 max 32768" "cfg3 max 10000" "cfg3 sum 10000" "cfg2 quant 8000"; do
  n=$(echo $c | tr ' ' '_')
  python profiles/prof_case.py $c 3 > gpurun_out/b_plain_$n.log 2>&1 && \
  ncu --set full --clock-control none --import-source on -k regex:'tc_score|qtc_score' -s 1 -c 1 -f -o gpurun_out/prof_${n}_r02 \
      python profiles/prof_case.py $c 3 > gpurun_out/b_ncu_$n.log 2>&1
  python profiles/summarize.py gpurun_out/prof_${n}_r02.ncu-rep > gpurun_out/prof_${n}_r02.txt 2>&1
done
cat $O
