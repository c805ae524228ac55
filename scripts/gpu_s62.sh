set -x
cd /root/repo
timeout 900 python -m pytest tests/test_gpu_model.py -m gpu -x -q -s -k "encoder or hf or teacher" 2>&1 | tail -25 > gpurun_out/s62_tests.txt
cat gpurun_out/s62_tests.txt
timeout 600 python bench.py --steps 20 --warmup 3 --legs none --no-cpu > gpurun_out/s62_bench.json 2> gpurun_out/s62_bench.err
tail -3 gpurun_out/s62_bench.err; cat gpurun_out/s62_bench.json
WQ4_CHAIN=0 timeout 600 python bench.py --steps 20 --warmup 3 --legs none --no-cpu > gpurun_out/s62_bench_nochain.json 2>> gpurun_out/s62_bench.err
cat gpurun_out/s62_bench_nochain.json
