set -x
cd /root/repo
timeout 900 ncu --set full --clock-control none --import-source on -k regex:decode2_kernel -s 200 -c 1 -o gpurun_out/r02_decode2_b1 -f python scripts/decode2_timeline.py 224 > gpurun_out/s70_ncu.log 2>&1
tail -3 gpurun_out/s70_ncu.log
ls -la gpurun_out/r02_decode2_b1.ncu-rep
