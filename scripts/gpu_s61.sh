set -x
cd /root/repo
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q -s 2>&1 | tail -15 > gpurun_out/s61_noalloc_alone.txt
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 > gpurun_out/s61_pytest_all.txt
cat gpurun_out/s61_noalloc_alone.txt gpurun_out/s61_pytest_all.txt
