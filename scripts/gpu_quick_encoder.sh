#!/bin/bash
# quick A/B of an encoder change: parity of the tiny / HF cases, chained-launch timeline, Medium one-clip bench line
cd "$(dirname "$0")/.."
timeout 600 python -m pytest tests/test_gpu_model.py -m gpu -x -q -k "encoder or hf" 2>&1 | tail -2
timeout 300 python scripts/chain_timeline.py medium 1 2>&1 | grep -E "^it |end of|error"
timeout 600 python bench.py --steps 20 --warmup 3 --legs none --no-cpu ${BENCH_ARGS:-} 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print(round(d['value'],1), round(d['ms_per_step'],3), round(d['roofline']['achieved'],1), round(d['roofline']['frac'],3), d['breakdown_ms_per_step'])"
