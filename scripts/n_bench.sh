#!/bin/bash
# usage: scripts/n_bench.sh <n_gpus> [bench.py args...]  -> one compact line per run
n=$1; shift
timeout 280 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $n --steps 5 --warmup 3 --no-e2e "$@" > gpurun_out/nb.out 2>&1
grep '^{"metric"' gpurun_out/nb.out | python -c "
import json,sys
for l in sys.stdin:
    d=json.loads(l); print('$*', 'ms/step', round(d['ms_per_step'],2), 'G/s', round(d['value']/1e9,1), {k[3:]:round(v,2) for k,v in d['phases']['ms_per_step'].items()}, 'verified', d['verified'])"
grep -E "Error|error|Traceback" gpurun_out/nb.out | head -5
