cd /root/repo
timeout 300 python scripts/decode2_timeline.py 224 > gpurun_out/r02_decode2_timeline.txt 2>&1
timeout 900 python scripts/bench_decode.py --batches 1,2,4,8,16,32,64,128 --kv 224 --reps 50 > gpurun_out/r02_decode_step.txt 2>&1
timeout 300 python scripts/prof_chain.py large_v3 8 || exit 1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:q4_chain -s 2 -c 1 -o gpurun_out/r02_chain_lv3_b8 -f python scripts/prof_chain.py large_v3 8 > gpurun_out/ncu_lv3.log 2>&1
tail -2 gpurun_out/ncu_lv3.log
