#!/bin/bash
# What a round-end check runs on the GPU box:  gpurun --timeout 3000 -- 'bash scripts/gpu_run_all.sh'
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -x -q 2>&1 | tail -8 | tee gpurun_out/pytest_gpu.txt
timeout 900 python bench.py --steps 20 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; tail -2 gpurun_out/bench.err
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -3
