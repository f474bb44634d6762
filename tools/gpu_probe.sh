#!/usr/bin/env bash
# Runs every GPU kernel test FUNCTION in its own process (a trapped kernel poisons the CUDA context of its
# process only), with a per-group timeout; summary in gpurun_out/probe_summary.txt.
set -u
mkdir -p gpurun_out
FILE="${1:-tests/test_kernels_gpu.py}"
OUT=gpurun_out/probe_summary.txt
: > "$OUT"; : > gpurun_out/probe_failures.log
nvidia-smi --query-gpu=name,driver_version,memory.total --format=csv >> "$OUT" 2>&1
python -m pytest "$FILE" -m gpu --collect-only -q 2>/dev/null | grep "::" | sed 's/\[.*//' | sort -u > gpurun_out/probe_tests.txt
while read -r t; do
  start=$(date +%s)
  timeout 300 python -m pytest "$t" -m gpu -q -s --no-header -p no:cacheprovider -rfE --tb=short > gpurun_out/probe_last.log 2>&1
  rc=$?
  end=$(date +%s)
  echo "$t rc=$rc $((end - start))s : $(tail -1 gpurun_out/probe_last.log)" >> "$OUT"
  grep -E "PARITY|E2E|UNFUSED|launches per" gpurun_out/probe_last.log | cut -c1-300 >> "$OUT"
  if [ $rc -ne 0 ]; then
    echo "----- $t" >> gpurun_out/probe_failures.log
    grep -E "^(FAILED|ERROR|E  )|Error|out of tolerance|rel-L2|trap|illegal|launch fail" gpurun_out/probe_last.log | cut -c1-300 | head -40 >> gpurun_out/probe_failures.log
  fi
done < gpurun_out/probe_tests.txt
echo "== summary"; cat "$OUT"; echo "== failures"; head -c 6000 gpurun_out/probe_failures.log
