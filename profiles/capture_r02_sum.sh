#!/bin/bash
# round 2, end: the full-sum kernel after the register rearrangement -- plain run first, then one launch under ncu
mkdir -p gpurun_out
name=prof_tc_cfg3_sum_10000_r02
python profiles/prof_case.py cfg3 sum 10000 3 > gpurun_out/plain_$name.log 2>&1 || { echo "plain run failed"; exit 1; }
ncu --set full --clock-control none --import-source on -k regex:tc_score -s 1 -c 1 -f -o gpurun_out/$name \
    python profiles/prof_case.py cfg3 sum 10000 3 > gpurun_out/ncu_$name.log 2>&1
python profiles/summarize.py gpurun_out/$name.ncu-rep > gpurun_out/$name.txt 2>&1
rm -f gpurun_out/$name.ncu-rep
cat gpurun_out/$name.txt
