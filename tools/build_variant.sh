#!/usr/bin/env bash
# Builds an experimental variant of the library into exp/lib_<name>.so (git-ignored; travels to the GPU box):
#   tools/build_variant.sh <name> [-DMACRO=VALUE ...]
# Variants are always built with -DDAB_TUNING (environment switches enabled).  Timing: DAB_EXP_LIB=exp/lib_<name>.so
# python tools/exp_scan.py ...   Nothing built here is part of the product.
set -euo pipefail
name=$1; shift
root=$(cd "$(dirname "$0")/.." && pwd)
src=$root/django-assistant-bot_b200/csrc
out=$root/exp/obj_$name
mkdir -p "$out"
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
FLAGS=(-gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -Xcompiler -fvisibility=hidden
       --expt-relaxed-constexpr -DDAB_TUNING "$@")
pids=()
for f in "$src"/*.cu; do
  b=$(basename "${f%.cu}")
  ( "$NVCC" "${FLAGS[@]}" -c "$f" -o "$out/$b.o" > "$out/$b.log" 2>&1 || { cat "$out/$b.log"; exit 1; } ) &
  pids+=($!)
done
for p in "${pids[@]}"; do wait "$p"; done
"$NVCC" -shared -gencode arch=compute_100a,code=sm_100a "$out"/*.o -o "$root/exp/lib_$name.so" -lcuda
echo "built exp/lib_$name.so"
