# A/B runs under environment switches of the tensor kernel (development log; results -> gpurun_out/ab.txt)
for v in "GMM_TC_EPI_WARPS=8" "GMM_TC_EPI_WARPS=16"; do
echo "env: $v"
env $v GMM_TC_DEBUG=1 timeout 120 python profiles/prof_case.py cfg1 max 32000 2 2>&1 | grep gmm_tc | tail -1
env $v GMM_TC_DEBUG=1 timeout 120 python profiles/prof_case.py cfg3 max 10000 2 2>&1 | grep gmm_tc | tail -1
env $v timeout 200 python bench.py --steps 10 --warmup 3 --no-cpu 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(' cfg1 ms', round(d['ms_per_step'],4), ' cfg3max ms', round(d['extra']['cfg3_max']['ms_per_step'],4), ' cfg3sum', round(d['extra']['cfg3_sum']['ms_per_step'],4), 'e2e', round(d['e2e']['value']))"
done
