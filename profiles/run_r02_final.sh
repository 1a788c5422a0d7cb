#!/bin/bash
# round 2, final state: whole GPU suite, smoke(), bench.py (default flags) and the reference arm
mkdir -p gpurun_out
O=gpurun_out/final_summary.txt
: > $O
timeout 1500 python -m pytest tests -q -m gpu > gpurun_out/final_suite.log 2>&1; echo "suite rc=$?" >> $O; tail -4 gpurun_out/final_suite.log >> $O
python -c "import __graft_entry__ as g; g.smoke()" >> $O 2>&1; echo "smoke rc=$?" >> $O
python bench.py > gpurun_out/bench_n1_r02.json 2> gpurun_out/bench_n1_r02.err; echo "bench rc=$?" >> $O
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref_n1_r02.json 2> gpurun_out/bench_ref_n1_r02.err; echo "reference arm rc=$?" >> $O
cat $O
head -c 600 gpurun_out/bench_n1_r02.json
