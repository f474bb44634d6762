#!/usr/bin/env bash
# Builds libb200sd.so in-tree for sm_100a (cross-compiles without a GPU).
# fast-math stays off: GroupNorm / timestep-embedding parity needs IEEE sinf/cosf/expf/division.
set -euo pipefail
HERE="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
OUT="${B200SD_LIB_OUT:-$HERE/../_lib}"
mkdir -p "$OUT"
NVCC="${NVCC:-/usr/local/cuda/bin/nvcc}"
FLAGS=(${B200SD_EXTRA_NVCC_FLAGS:-} -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17
       -Xcompiler -fPIC -Xcompiler -fvisibility=hidden -Xptxas -v --expt-relaxed-constexpr)
OBJS=()
pids=()
for f in gemm attention bandwidth peer program; do
  ( "$NVCC" "${FLAGS[@]}" -c "$HERE/$f.cu" -o "$OUT/$f.o" > "$OUT/$f.ptxas.log" 2>&1 || { cat "$OUT/$f.ptxas.log"; exit 1; } ) &
  pids+=($!)
  OBJS+=("$OUT/$f.o")
done
for p in "${pids[@]}"; do wait "$p"; done
"$NVCC" -shared -o "$OUT/libb200sd.so" "${OBJS[@]}" -cudart static
echo "built $OUT/libb200sd.so"
