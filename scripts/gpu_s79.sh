cd /root/repo
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 --steps 20 --warmup 3 > gpurun_out/s79_bench_2gpu.json 2> gpurun_out/s79.err
tail -3 gpurun_out/s79.err
python -c "
import json
d=json.loads(open('gpurun_out/s79_bench_2gpu.json').read().strip().splitlines()[-1])
print(d['value'], d['n_gpus'], d['ms_per_step'], d['e2e']['value'])
print(d.get('legs',{}).get('transcribe_256'))
"
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29534 bench.py --impl reference --gpus 2 --steps 1 --warmup 0 2>/dev/null | tail -1 | cut -c1-400
