set -x
cd /root/repo
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -30 > gpurun_out/s60_pytest_all.txt
timeout 900 python bench.py --steps 20 --warmup 3 > gpurun_out/s60_bench.json 2> gpurun_out/s60_bench.err
tail -5 gpurun_out/s60_bench.err
timeout 300 python scripts/decode2_timeline.py 224 > gpurun_out/s60_timeline.txt 2>&1
timeout 300 python scripts/bench_decode.py --batches 1,2,4,8,16,32,64 --kv 224 --reps 50 >> gpurun_out/s60_timeline.txt 2>&1
