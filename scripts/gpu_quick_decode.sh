#!/bin/bash
# quick A/B of a decode-path change: parity (tiny model, every batch size; full-size teacher-forced logits), decode table
cd "$(dirname "$0")/.."
timeout 1200 python -m pytest tests/test_gpu_model.py tests/test_gpu_config4.py -m gpu -x -q -k "every_batch_size or teacher_forced or compacted or tokens" 2>&1 | tail -2
timeout 900 python scripts/bench_decode.py --batches ${BATCHES:-1,8,16,32,64,128} --kv 224 --reps 30 2>&1 | python -c "
import sys, json
for ln in sys.stdin:
    if ln.startswith('{'):
        d = json.loads(ln); print(d['batch'], round(d['us_per_step'], 1), d['launches_per_step'])
    else: print(ln.strip())"
