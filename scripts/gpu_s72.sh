set -x
cd /root/repo
timeout 600 python bench.py --steps 2 --warmup 3 --legs none --no-cpu > gpurun_out/s72_plain.json 2> gpurun_out/s72.err || exit 1
timeout 1200 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --csv --log-file gpurun_out/r02_ncu_launches_medium_b1.csv python bench.py --steps 2 --warmup 3 --legs none --no-cpu > gpurun_out/s72_ncu.log 2>&1
tail -2 gpurun_out/s72_ncu.log
wc -l gpurun_out/r02_ncu_launches_medium_b1.csv
