#!/bin/bash
# 1 -> 8 GPU run of bench.py the way the driver launches it (needs gpurun --gpus 8)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
for n in 8 4 2; do
python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29600 + n)) bench.py --gpus $n --steps 20 --warmup 3 > gpurun_out/r02_scale_n$n.json 2> gpurun_out/r02_scale_n$n.err
python - <<PY
import json
d=json.loads(open('gpurun_out/r02_scale_n$n.json').read().strip().splitlines()[-1])
t=d['legs']['transcribe_256']
print('n=$n', 'encoder clips/s', round(d['value'],1), 'e2e', round(d['e2e']['value'],1), '| transcribe_256 clips/s', round(t['clips_per_s'],1), 'tokens/s', round(t['tokens_per_s']), 'micro', t['micro_batch'])
PY
done
