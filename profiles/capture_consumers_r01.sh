# Round-1 capture of the consumer kernels (run under gpurun): plain run first, then the same command under ncu.
set -x
export ALIGN_RANKED_ONLY=1
python profiles/time_align.py 512 1000 300 > gpurun_out/plain_align.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:align_ranked -s 1 -c 1 -f -o gpurun_out/prof_align_512x1000_r01 \
    python profiles/time_align.py 512 1000 300 > gpurun_out/ncu_align.log 2>&1
python profiles/time_accumulate.py 512000 > gpurun_out/plain_accum.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:accumulate_kernel -s 1 -c 1 -f -o gpurun_out/prof_accum_512000_r01 \
    python profiles/time_accumulate.py 512000 > gpurun_out/ncu_accum.log 2>&1
python profiles/time_lda.py 512 1000 > gpurun_out/plain_lda.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:lda_kernel -s 1 -c 1 -f -o gpurun_out/prof_lda_512000_r01 \
    python profiles/time_lda.py 512 1000 > gpurun_out/ncu_lda.log 2>&1
cat gpurun_out/plain_align.log gpurun_out/plain_accum.log gpurun_out/plain_lda.log
