set -x
cd /root/repo
WQ4_CHAIN_TILE_M=192 timeout 600 python bench.py --steps 20 --warmup 3 --legs none --no-cpu > gpurun_out/s65_b1_tm192.json 2> gpurun_out/s65.err
timeout 600 python bench.py --model large_v3 --batch 8 --steps 10 --warmup 3 --legs none --no-cpu > gpurun_out/s65_lv3b8_chain.json 2>> gpurun_out/s65.err
WQ4_CHAIN_TILE_M=128 timeout 600 python bench.py --model large_v3 --batch 8 --steps 10 --warmup 3 --legs none --no-cpu > gpurun_out/s65_lv3b8_chain128.json 2>> gpurun_out/s65.err
WQ4_CHAIN=0 timeout 600 python bench.py --model large_v3 --batch 8 --steps 10 --warmup 3 --legs none --no-cpu > gpurun_out/s65_lv3b8_nochain.json 2>> gpurun_out/s65.err
tail -3 gpurun_out/s65.err
for f in gpurun_out/s65_*.json; do echo $f; python -c "
import json,sys
d=json.loads(open('$f').read().strip().splitlines()[-1])
print(d['value'], d['ms_per_step'], d['roofline']['achieved'], d['roofline']['frac'], d['breakdown_ms_per_step'])
"; done
