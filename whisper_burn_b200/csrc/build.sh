#!/bin/bash
# Builds whisper_burn_b200/libwq4.so for sm_100a with nvcc (cross-compiles without a GPU).
set -euo pipefail
HERE="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
OUT="$HERE/../libwq4.so"
NVCC="${NVCC:-/usr/local/cuda/bin/nvcc}"
FLAGS=(-gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo
       -Xcompiler -fPIC,-fvisibility=hidden,-Wall -shared -cudart static
       --expt-relaxed-constexpr -Xptxas -v)
"$NVCC" "${FLAGS[@]}" -o "$OUT" "$HERE"/wq4_api.cu "$HERE"/wq4_tensor.cu "$HERE"/wq4_gemv.cu "$HERE"/wq4_gemm.cu "$HERE"/wq4_gemm2.cu "$HERE"/wq4_model_kernels.cu "$HERE"/wq4_decode.cu "$HERE"/wq4_attn.cu "$HERE"/wq4_mel.cu "$HERE"/wq4_whisper.cu "$@"
echo "built $OUT"
