set -x
cd /root/repo
timeout 900 python -m pytest tests/test_gpu_model.py -m gpu -x -q -s -k "encoder or hf or teacher" 2>&1 | tail -8
timeout 300 python scripts/chain_timeline.py medium 1 > gpurun_out/s67_tl.txt 2>&1; grep -E "^it |end of|error" gpurun_out/s67_tl.txt
for v in 1 0; do
WQ4_CHAIN_SPLITK=$v timeout 600 python bench.py --steps 20 --warmup 3 --legs none --no-cpu 2>> gpurun_out/s67.err | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1])
print('splitk=$v', round(d['value'],1), round(d['ms_per_step'],3), round(d['roofline']['achieved'],1), round(d['roofline']['frac'],3), d['breakdown_ms_per_step'])
"
done
