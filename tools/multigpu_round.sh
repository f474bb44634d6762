#!/bin/bash
# 2-GPU validation + measurements of one round (run with: gpurun --gpus 2 -- bash tools/multigpu_round.sh)
set -u
OUT=gpurun_out
mkdir -p $OUT
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
nvidia-smi topo -m > $OUT/r02_topo.txt 2>&1
timeout 900 python -m pytest tests/test_multigpu_gpu.py -q -s -p no:cacheprovider > $OUT/r02_pytest_multigpu.log 2>&1; echo "pytest multigpu rc=$?"; tail -3 $OUT/r02_pytest_multigpu.log
grep MULTIGPU $OUT/r02_pytest_multigpu.log
SWEEP_B=${SWEEP_B:-1,2,4,8,16} SWEEP_RES=${SWEEP_RES:-512,1024} timeout 900 $TR --master-port 29611 tools/sweep_config5.py > $OUT/r02_sweep_config5_2gpu.txt 2> $OUT/r02_sweep_config5_2gpu.err; echo "sweep rc=$?"; head -30 $OUT/r02_sweep_config5_2gpu.txt
for args in "--strong" "--pair-split --workload sd15_512_b1_euler20_cfg" "--pair-split" "--pair-split --pair-exchange nccl --workload sd15_512_b1_euler20_cfg"; do
  tag=$(echo $args | tr -d '-' | tr ' ' '_')
  timeout 400 $TR --master-port 29612 bench.py --gpus 2 --steps 3 --warmup 3 --no-extras $args > $OUT/r02_bench_2gpu_$tag.json 2> $OUT/r02_bench_2gpu_$tag.err; echo "bench $args rc=$?"
  python -c "import json,sys; d=json.load(open('$OUT/r02_bench_2gpu_$tag.json')); print(d['config']['parallelism'], d['scaling'], 'B/gpu', d['config']['batch_per_gpu'], 'images/s', round(d['value'],3), 'step ms', round(d['unet_step_ms'],3))"
done
