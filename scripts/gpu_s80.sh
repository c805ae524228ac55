cd /root/repo
timeout 1800 python -m pytest tests -m gpu -x -q 2>&1 | tail -8 > gpurun_out/s80_pytest_all.txt
cat gpurun_out/s80_pytest_all.txt
timeout 900 python bench.py --steps 20 --warmup 3 > gpurun_out/s80_bench.json 2> gpurun_out/s80_bench.err
tail -2 gpurun_out/s80_bench.err
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3
