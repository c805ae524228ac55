set -x
cd /root/repo
timeout 600 python -m pytest tests/test_gpu_model.py -m gpu -x -q -k "every_batch_size or compacted or budgets or tokens_match" 2>&1 | tail -15 > gpurun_out/s52_tests.txt
timeout 600 python -m pytest tests/test_gpu_config4.py -m gpu -x -q -s -k "teacher_forced or compacted or tokens" 2>&1 | tail -15 >> gpurun_out/s52_tests.txt
timeout 300 python scripts/decode2_timeline.py 224 > gpurun_out/s52_timeline.txt 2>&1
timeout 300 python scripts/bench_decode.py --batches 1 --kv 224 --reps 100 >> gpurun_out/s52_timeline.txt 2>&1
