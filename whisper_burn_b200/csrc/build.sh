#!/bin/bash
# Builds whisper_burn_b200/libwq4.so for sm_100a with nvcc (cross-compiles without a GPU).
# One object per source under csrc/build/ (compiled in parallel, only when the source or a header is newer), then one link.
set -euo pipefail
HERE="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
OUT="$HERE/../libwq4.so"
OBJ="$HERE/build"
NVCC="${NVCC:-/usr/local/cuda/bin/nvcc}"
FLAGS=(-gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo
       -Xcompiler -fPIC,-fvisibility=hidden,-Wall
       --expt-relaxed-constexpr -Xptxas -v)
SRCS=(wq4_api wq4_tensor wq4_gemv wq4_skinny wq4_gemm wq4_gemm2 wq4_chain wq4_model_kernels wq4_decode wq4_decode2 wq4_attn wq4_mel wq4_whisper wq4_noalloc)
# --clean: forget every object first (what __graft_entry__.build() does: the check must prove the sources compile)
if [ "${1:-}" = "--clean" ]; then shift; rm -rf "$OBJ"; fi
mkdir -p "$OBJ"
newest_hdr=$(ls -t "$HERE"/*.cuh "$HERE"/../../include/*.h "$HERE"/build.sh | head -1)
pids=()
for s in "${SRCS[@]}"; do
    o="$OBJ/$s.o"
    if [ ! -f "$o" ] || [ "$HERE/$s.cu" -nt "$o" ] || [ "$newest_hdr" -nt "$o" ]; then
        ( "$NVCC" "${FLAGS[@]}" "$@" -c -o "$o.tmp" "$HERE/$s.cu" > "$OBJ/$s.log" 2>&1 && mv "$o.tmp" "$o" ) &
        pids+=("$!:$s")
    fi
done
fail=0
for ps in "${pids[@]:-}"; do
    [ -z "$ps" ] && continue
    if ! wait "${ps%%:*}"; then fail=1; echo "compile failed: ${ps##*:}.cu"; cat "$OBJ/${ps##*:}.log"; fi
done
[ "$fail" = 0 ] || exit 1
objs=()
for s in "${SRCS[@]}"; do objs+=("$OBJ/$s.o"); done
"$NVCC" -gencode arch=compute_100a,code=sm_100a -shared -cudart static -Xcompiler -fPIC -o "$OUT" "${objs[@]}"
echo "built $OUT"
