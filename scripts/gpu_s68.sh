set -x
cd /root/repo
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -12 > gpurun_out/s68_pytest_all.txt
cat gpurun_out/s68_pytest_all.txt
timeout 900 python bench.py --steps 20 --warmup 3 > gpurun_out/s68_bench.json 2> gpurun_out/s68_bench.err
tail -3 gpurun_out/s68_bench.err
