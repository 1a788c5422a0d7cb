# Round-1 capture recipe, second half of the round (L2-sized launches, two-pass full-sum epilogue, 16-warp quantized
# epilogue, compact-block alignment pipeline).  Run under gpurun: every command first plain, then the same under ncu.
set -x
python bench.py --steps 20 --warmup 3 > gpurun_out/bench_n1_r01.json 2> gpurun_out/bench_full.err || exit 1
python bench.py --steps 2 --warmup 1 --no-cpu --no-extra > gpurun_out/plain_bench.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file gpurun_out/launches_bench_r01.csv \
    python bench.py --steps 2 --warmup 1 --no-cpu --no-extra > gpurun_out/ncu_bench.log 2>&1
for c in "cfg1 max 32768" "cfg3 max 10000" "cfg3 sum 10000"; do
  n=$(echo $c | tr ' ' '_')
  python profiles/prof_case.py $c 3 > gpurun_out/plain_$n.log 2>&1 && \
  ncu --set full --clock-control none --import-source on -k regex:tc_score -s 1 -c 1 -f -o gpurun_out/prof_tc_${n}_r01 \
      python profiles/prof_case.py $c 3 > gpurun_out/ncu_$n.log 2>&1
done
python profiles/prof_case.py cfg2 quant 8000 3 > gpurun_out/plain_quant.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:qtc_score -s 1 -c 1 -f -o gpurun_out/prof_quant_cfg2_r01 \
    python profiles/prof_case.py cfg2 quant 8000 3 > gpurun_out/ncu_quant.log 2>&1
# the alignment pipeline: compact-block scoring (subset tensor pass) and the alignment kernel reading it
export PIPELINE_CASES=1:296
python profiles/time_pipeline.py 1024 > gpurun_out/plain_pipeline.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:align_ranked -s 2 -c 1 -f -o gpurun_out/prof_align_compact_r01 \
    python profiles/time_pipeline.py 1024 > gpurun_out/ncu_pipe_align.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:tc_score -s 2 -c 1 -f -o gpurun_out/prof_tc_subset_r01 \
    python profiles/time_pipeline.py 1024 > gpurun_out/ncu_pipe_tc.log 2>&1
cat gpurun_out/plain_pipeline.log
