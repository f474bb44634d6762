#!/bin/bash
# Same-box A/B of pre-built library variants (stablediffusion_b200/_variants/<name>/libb200sd.so, built on the CPU box with
# B200SD_LIB_OUT=... B200SD_EXTRA_NVCC_FLAGS=... csrc/build.sh): op micro-bench per variant, optionally the sustained bench line.
# Usage: gpurun -- bash tools/ab_variants.sh "<bench_ops --only list>" [variants...]
ONLY=${1:-attn_self,attn_xl}; shift
VARS=${@:-$(ls stablediffusion_b200/_variants)}
mkdir -p gpurun_out
for v in $VARS; do
  echo "=== $v"
  B200SD_LIB_DIR=$PWD/stablediffusion_b200/_variants/$v timeout 300 python tools/bench_ops.py --only $ONLY --reps ${REPS:-7} --out gpurun_out/ab_$v.txt 2>&1 | grep -v Warn
  if [ "${STEP:-0}" = "1" ]; then
    B200SD_LIB_DIR=$PWD/stablediffusion_b200/_variants/$v timeout 300 python tools/profile_step.py --batch 8 --out gpurun_out/ab_step_$v.txt > /dev/null 2>&1; head -1 gpurun_out/ab_step_$v.txt
  fi
done
