#!/bin/bash
# Sweep of the push kernel's shape (loads in flight per thread, CTAs) under torchrun; N from $1, frames from $2.
N=${1:-2}; T=${2:-1024}; M=${3:-16384}
for U in ${US:-1 2 4}; do for C in ${CS:-32 64 128}; do
  echo -n "unroll $U ctas $C: "
  GMM_GATHER_UNROLL=$U GMM_GATHER_CTAS=$C timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N \
    --master-addr 127.0.0.1 --master-port 29533 profiles/bench_sharded.py $M 64 $T 20 2>/dev/null | grep "^{" | tail -1 | \
    python -c "import json,sys; d=json.loads(sys.stdin.read()); print(' '.join(f'{k[12:]}={v:.3f}' for k,v in d.items() if k.startswith('ms_per_step')), all(v for k,v in d.items() if k.startswith('bit_')))"
done; done
