#!/usr/bin/env bash
# Builds libdab_b200.so in-tree for sm_100a (B200).  nvcc cross-compiles without a GPU.
set -euo pipefail
cd "$(dirname "$0")"
OUT=../dab_b200/libdab_b200.so
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
FLAGS=(-gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -Xcompiler -fvisibility=hidden
       --expt-relaxed-constexpr -Xptxas -v)
mkdir -p obj
pids=()
objs=()
for f in pool.cu select.cu scan_simt.cu scan_umma.cu scan_umma_quad.cu store.cu filter.cu noalloc.cu exchange.cu sharded.cu docagg.cu; do
  [ -f "$f" ] || continue
  o=obj/${f%.cu}.o
  objs+=("$o")
  if [ ! -f "$o" ] || [ "$f" -nt "$o" ] || [ common.cuh -nt "$o" ] || [ search.cuh -nt "$o" ] || [ store.cuh -nt "$o" ] || [ umma.cuh -nt "$o" ] || [ ../../include/dab_b200.h -nt "$o" ]; then
    ( "$NVCC" "${FLAGS[@]}" -c "$f" -o "$o" > "obj/${f%.cu}.log" 2>&1 || { cat "obj/${f%.cu}.log"; exit 1; } ) &
    pids+=($!)
  fi
done
for p in "${pids[@]:-}"; do [ -n "$p" ] && wait "$p"; done
"$NVCC" -shared -gencode arch=compute_100a,code=sm_100a "${objs[@]}" -o "$OUT" -lcuda
echo "built $OUT"
