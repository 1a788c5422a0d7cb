#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_rasr_plugin.py tests/test_gpu_parity.py -q -m gpu -x -s 2>&1 | grep -E "PASS|FAIL|SKIP|TIMING|passed|failed" | grep -E "batch|TIMING|passed|failed|FAIL" | tail -40
