set -x
cd /root/repo
timeout 900 python -m pytest tests/test_gpu_model.py -m gpu -x -q -s -k "encoder or hf or teacher" 2>&1 | tail -12 > gpurun_out/s64_tests.txt
cat gpurun_out/s64_tests.txt
timeout 300 python scripts/chain_timeline.py medium 1 > gpurun_out/s64_tl.txt 2>&1; tail -32 gpurun_out/s64_tl.txt
timeout 600 python bench.py --steps 20 --warmup 3 --legs none --no-cpu > gpurun_out/s64_bench.json 2> gpurun_out/s64_bench.err
tail -3 gpurun_out/s64_bench.err; cat gpurun_out/s64_bench.json
