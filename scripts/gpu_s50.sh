set -x
cd /root/repo
python -m pytest tests/test_gpu_config4.py tests/test_gpu_model.py tests/test_gpu_mel.py tests/test_gpu_skinny.py -m gpu -x -q -s 2>&1 | tail -80 > gpurun_out/s50_new_tests.txt
python -m pytest tests -m gpu -q 2>&1 | tail -30 > gpurun_out/s50_pytest_all.txt
