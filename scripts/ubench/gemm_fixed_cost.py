import sys
sys.path.insert(0, '/root/repo/scripts'); sys.path.insert(0, '/root/repo')
import microbench as mb
for rows, n, k in ((128, 256, 128), (128, 256, 1024), (1500, 1024, 128), (1500, 1024, 256), (1500, 1024, 512), (1500, 1024, 1024), (1500, 1024, 2048), (1500, 1024, 4096), (750, 1024, 1024), (3000, 1024, 1024)):
    for mode in ("f16", "res"):
        dt, tile = mb.bench_gemm(rows, n, k, reps=20, mode=mode)
        print(f"rows {rows} N {n} K {k} mode {mode} tile {tile}: {dt*1e6:.2f} us", flush=True)
