set -x
cd /root/repo
python bench.py --steps 20 --warmup 3 > gpurun_out/s51_bench.json 2> gpurun_out/s51_bench.err
tail -5 gpurun_out/s51_bench.err
