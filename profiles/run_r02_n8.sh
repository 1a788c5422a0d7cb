#!/bin/bash
# round 2, eight GPUs: bench.py at N=8 (weak scaling + extra.cfg5_mixture_sharded)
mkdir -p gpurun_out
O=gpurun_out/n8_summary.txt
: > $O
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29621 bench.py --gpus 8 --steps 10 --warmup 3 > gpurun_out/n8_bench.json 2> gpurun_out/n8_bench.err; echo "bench rc=$?" >> $O
tail -3 gpurun_out/n8_bench.err >> $O
python - >> $O <<'PY'
import json
try:
    l = json.loads(open('gpurun_out/n8_bench.json').read().strip().splitlines()[-1])
    print("value", l["value"], "e2e", l["e2e"]["value"], "align", l["e2e_alignment"]["value"], "train", l["e2e_training_pass"]["value"])
    print(json.dumps(l["extra"].get("cfg5_mixture_sharded"), indent=1))
except Exception as e:
    print("no bench line", e)
PY
cat $O
