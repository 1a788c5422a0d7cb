#!/bin/bash
# round 2, final captures (one GPU): each program first runs plain and must exit 0, then under ncu.
#   * ncu --set full of one launch of each scoring kernel (the second launch of the process) -> gpurun_out/prof_*_r02.{ncu-rep,txt}
#   * launch list of `python bench.py --steps 2 --warmup 1` (gpu__time_duration, no clock control) -> launches_bench_r02.csv
#   * the bench itself, plain, default flags -> bench_n1_r02.json, and the reference arm
mkdir -p gpurun_out
O=gpurun_out/capture_summary.txt
: > $O
cap() {  # name, kernel regex, prof_case args...
  local name=$1 rex=$2; shift 2
  python profiles/prof_case.py "$@" 3 > gpurun_out/plain_$name.log 2>&1 || { echo "$name: plain run failed" >> $O; return; }
  ncu --set full --clock-control none --import-source on -k regex:"$rex" -s 1 -c 1 -f -o gpurun_out/$name \
      python profiles/prof_case.py "$@" 3 > gpurun_out/ncu_$name.log 2>&1
  python profiles/summarize.py gpurun_out/$name.ncu-rep > gpurun_out/$name.txt 2>&1
  echo "$name: $(grep -c . gpurun_out/$name.txt) summary lines" >> $O
}
cap prof_tc_cfg1_max_32768_r02 'tc_score' cfg1 max 32768
cap prof_tc_cfg3_max_10000_r02 'tc_score' cfg3 max 10000
cap prof_tc_cfg3_sum_10000_r02 'tc_score' cfg3 sum 10000
cap prof_quant_cfg2_r02 'qtc_score' cfg2 quant 8000
python bench.py --steps 2 --warmup 1 > gpurun_out/bench_short.json 2> gpurun_out/bench_short.err && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_bench_r02.csv \
    python bench.py --steps 2 --warmup 1 > gpurun_out/bench_under_ncu.log 2>&1
echo "launch list: $(wc -l < gpurun_out/launches_bench_r02.csv) lines" >> $O
python bench.py > gpurun_out/bench_n1_r02.json 2> gpurun_out/bench_n1_r02.err; echo "bench rc=$?" >> $O
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref_n1_r02.json 2> gpurun_out/bench_ref_n1_r02.err; echo "reference arm rc=$?" >> $O
rm -f gpurun_out/*.ncu-rep.tmp
ls -la gpurun_out/*.ncu-rep >> $O
cat $O
head -c 1500 gpurun_out/bench_n1_r02.json
