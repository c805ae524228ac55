set -x
cd /root/repo
timeout 300 python scripts/prof_chain.py medium 1 || exit 1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:q4_chain -s 2 -c 1 -o gpurun_out/r02_chain_medium_b1 -f python scripts/prof_chain.py medium 1 > gpurun_out/s66_ncu.log 2>&1
tail -3 gpurun_out/s66_ncu.log
ls -la gpurun_out/*.ncu-rep
