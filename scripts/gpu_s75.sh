set -x
cd /root/repo
timeout 1200 python -m pytest tests/test_gpu_model.py tests/test_gpu_config4.py tests/test_gpu_mel.py -m gpu -x -q 2>&1 | tail -8
for v in 1 2; do
WQ4_DECODE_CHAIN=$v timeout 600 python scripts/bench_decode.py --batches 18,32,64,128 --kv 224 --reps 30 2>&1 | python -c "
import sys, json
for ln in sys.stdin:
    if ln.startswith('{'):
        d = json.loads(ln); print('chain=$v', d['batch'], round(d['us_per_step'], 1), d['launches_per_step'])
    else: print(ln.strip())
"
done
