# Round-1 capture recipe (run under gpurun): plain run first, then the same command under ncu.
set -x
python bench.py --steps 10 --warmup 3 > gpurun_out/bench_full.json 2> gpurun_out/bench_full.err || exit 1
python bench.py --steps 2 --warmup 1 --no-cpu --no-extra > gpurun_out/plain_bench.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_r01.csv \
    python bench.py --steps 2 --warmup 1 --no-cpu --no-extra > gpurun_out/ncu_bench.log 2>&1
for c in "cfg1 max 512000" "cfg3 max 10000" "cfg3 sum 10000"; do
  n=$(echo $c | tr ' ' '_')
  python profiles/prof_case.py $c 3 > gpurun_out/plain_$n.log 2>&1 && \
  ncu --set full --clock-control none --import-source on -k regex:tc_score -s 1 -c 1 -f -o gpurun_out/prof_tc_${n}_r01 \
      python profiles/prof_case.py $c 3 > gpurun_out/ncu_$n.log 2>&1
done
