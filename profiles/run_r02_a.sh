#!/bin/bash
# round 2, first GPU pass.  Phase 1 (library as built before the quantized-kernel change): the no-alloc entry point
# alone, the whole suite under the new default split, benches.  Phase 2: rebuild with the output-templated quantized
# kernel, its tests, bench again.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,power.limit --format=csv > gpurun_out/a_gpu.txt
timeout 300 python -m pytest tests/test_zz_no_alloc.py -q -m gpu > gpurun_out/a_noalloc_alone.log 2>&1; echo "noalloc_alone rc=$?" | tee -a gpurun_out/a_rc.txt
timeout 900 python -m pytest tests -q -m gpu -rs -s --deselect tests/test_gpu_parity.py::test_quantized_output_sets_bit_exact > gpurun_out/a_suite_default.log 2>&1; echo "suite_default rc=$?" | tee -a gpurun_out/a_rc.txt
GMM_TC_SPLIT=fp16 timeout 900 python -m pytest tests -q -m gpu --deselect tests/test_gpu_fp8_split.py --deselect tests/test_gpu_parity.py::test_quantized_output_sets_bit_exact > gpurun_out/a_suite_fp16.log 2>&1; echo "suite_fp16 rc=$?" | tee -a gpurun_out/a_rc.txt
timeout 600 python bench.py > gpurun_out/a_bench_default.json 2> gpurun_out/a_bench_default.err; echo "bench_default rc=$?" | tee -a gpurun_out/a_rc.txt
GMM_TC_SPLIT=fp16 timeout 600 python bench.py > gpurun_out/a_bench_fp16.json 2> gpurun_out/a_bench_fp16.err; echo "bench_fp16 rc=$?" | tee -a gpurun_out/a_rc.txt
# phase 2
( make -s -C rasr_b200/csrc > gpurun_out/a_make.log 2>&1 ); echo "make rc=$?" | tee -a gpurun_out/a_rc.txt
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_rasr_plugin.py tests/test_zz_no_alloc.py -q -m gpu -k "quant or plugin or static" > gpurun_out/a_quant_new.log 2>&1; echo "quant_new rc=$?" | tee -a gpurun_out/a_rc.txt
timeout 600 python bench.py > gpurun_out/a_bench_newqtc.json 2> gpurun_out/a_bench_newqtc.err; echo "bench_newqtc rc=$?" | tee -a gpurun_out/a_rc.txt
tail -5 gpurun_out/a_noalloc_alone.log gpurun_out/a_suite_default.log gpurun_out/a_suite_fp16.log gpurun_out/a_quant_new.log
