#!/bin/bash
# Tuning sweep of the decode attention launch shape (splits, 8-warp/2-warp variant) through WQ4_DA_CROSS / WQ4_DA_SELF.
for B in 4 16; do
 echo "== B=$B default"; python scripts/bench_decode.py --batches $B 2>&1 | grep -o "us_per_step[^,]*"
 for c in "8,0" "4,0" "2,1" "1,1" "2,0" "4,1"; do echo "cross $c"; WQ4_DA_CROSS=$c python scripts/bench_decode.py --batches $B 2>&1 | grep -o "us_per_step[^,]*"; done
 for c in "1,1" "2,0" "1,0" "2,1"; do echo "self $c"; WQ4_DA_SELF=$c python scripts/bench_decode.py --batches $B 2>&1 | grep -o "us_per_step[^,]*"; done
done
