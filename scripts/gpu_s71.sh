set -x
cd /root/repo
timeout 900 python -m pytest tests/test_gpu_model.py tests/test_gpu_config4.py -m gpu -x -q 2>&1 | tail -6
timeout 300 python scripts/decode2_timeline.py 224 > gpurun_out/s71_tl.txt 2>&1; tail -34 gpurun_out/s71_tl.txt
timeout 300 python scripts/bench_decode.py --batches 1 --kv 224 --reps 100 2>&1 | tail -2
