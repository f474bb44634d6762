#!/bin/bash
# One-GPU round check: new kernel tests, per-op step tables, the bench line (and its A/B switches), ncu launch list + full
# captures of the two top kernels.  Usage: gpurun -- bash tools/round2_check.sh [tag]
set -u
T=${1:-r02}
OUT=gpurun_out
mkdir -p $OUT
if [ "${FULL:-0}" = "1" ]; then   # what the driver runs at round end: the whole GPU suite, then smoke()
  python -m pytest tests -m gpu -x -q -p no:cacheprovider > $OUT/${T}_pytest_gpu.log 2>&1; echo "pytest -m gpu rc=$?"; tail -3 $OUT/${T}_pytest_gpu.log
  grep -E "PARITY|PSNR|rel-L2" $OUT/${T}_pytest_gpu.log | head -40
  python -c "import __graft_entry__ as g; g.smoke()" > $OUT/${T}_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 $OUT/${T}_smoke.log
else
  python -m pytest tests/test_kernels_gpu.py tests/test_configs_gpu.py tests/test_zz_no_alloc.py -m gpu -x -q -p no:cacheprovider > $OUT/${T}_pytest_kernels.log 2>&1; echo "pytest rc=$?"; tail -3 $OUT/${T}_pytest_kernels.log
fi
python tools/profile_step.py --batch 8 --out $OUT/${T}_profile_step_b8.txt > /dev/null 2>&1; head -1 $OUT/${T}_profile_step_b8.txt
python tools/profile_step.py --batch 1 --out $OUT/${T}_profile_step_b1.txt > /dev/null 2>&1; head -1 $OUT/${T}_profile_step_b1.txt
python tools/profile_vae.py --out $OUT/${T}_profile_vae_b8.txt > /dev/null 2>&1; head -3 $OUT/${T}_profile_vae_b8.txt
BENCH="python bench.py --steps 3 --warmup 3 --no-cpu-baseline"
$BENCH > $OUT/${T}_bench.json 2> $OUT/${T}_bench.err; echo "bench rc=$?"
python -c "import json; d=json.load(open('$OUT/${T}_bench.json')); print('images/s', round(d['value'],3), 'e2e', round(d['e2e']['value'],3), 'step ms', round(d['unet_step_ms'],3), 'gemm frac', round(d['roofline']['frac'],3), 'attn TF', round(d['roofline']['attention_tflops'],1)); [print(' ', k, round(v['images_per_s'],3), 'step ms', round(v['unet_step_ms'],3)) for k,v in (d['extra_workloads'] or {}).items()]"
# cost of the batch-invariant mode on config 2 (same box, back to back)
B200SD_BATCH_INVARIANT=1 $BENCH --no-extras --no-yardstick > $OUT/${T}_bench_batch_invariant.json 2>/dev/null; echo "bench invariant rc=$?"
python -c "import json; a=json.load(open('$OUT/${T}_bench.json')); b=json.load(open('$OUT/${T}_bench_batch_invariant.json')); print('batch-invariant mode: images/s', round(b['value'],3), 'vs default', round(a['value'],3), '-> cost %.2f %%' % (100*(1-b['value']/a['value'])), '; step ms', round(b['unet_step_ms'],3), 'vs', round(a['unet_step_ms'],3))"
if [ "${NCU:-1}" = "1" ]; then
  SHORT="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-extras --no-yardstick"
  timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -s 61500 -c 1000 --csv --log-file $OUT/${T}_launches.csv $SHORT > $OUT/${T}_ncu_launch.log 2>&1; echo "ncu launch list rc=$?"; tail -2 $OUT/${T}_ncu_launch.log | cut -c1-300
  python tools/ncu_launch_summary.py $OUT/${T}_launches.csv --ours --traffic-json=$OUT/${T}_gemm_traffic.json > $OUT/${T}_ncu_launch_list.txt 2>&1; head -14 $OUT/${T}_ncu_launch_list.txt
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:gemm_tc_kernel -s 20000 -c 3 -o $OUT/${T}_prof_gemm $SHORT > $OUT/${T}_ncu_gemm.log 2>&1; echo "ncu gemm rc=$?"
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:attention_kernel -s 600 -c 2 -o $OUT/${T}_prof_attn $SHORT > $OUT/${T}_ncu_attn.log 2>&1; echo "ncu attention rc=$?"
  timeout 600 ncu --set full --clock-control none -k regex:gn_apply -s 2000 -c 3 -o $OUT/${T}_prof_gn $SHORT > $OUT/${T}_ncu_gn.log 2>&1; echo "ncu groupnorm rc=$?"
fi
