#!/bin/bash
# round 2, four GPUs: the two-rank exchange test, bench.py at N=4 (extra.cfg5_mixture_sharded), reference arm at N=4
mkdir -p gpurun_out
O=gpurun_out/n4_summary.txt
: > $O
nvidia-smi -L >> $O
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29631 bench.py --gpus 4 --steps 10 --warmup 3 > gpurun_out/n4_bench.json 2> gpurun_out/n4_bench.err; echo "bench rc=$?" >> $O
tail -5 gpurun_out/n4_bench.err >> $O
python - >> $O <<'PY'
import json
try:
    l = json.loads(open('gpurun_out/n4_bench.json').read().strip().splitlines()[-1])
    print("value", l["value"], "e2e", l["e2e"]["value"])
    print(json.dumps(l["extra"].get("cfg5_mixture_sharded"), indent=1))
except Exception as e:
    print("no bench line", e)
PY
cat $O
