# Weak-scaling bench at N = 2, 4, 8 GPUs of one box (the driver's launch line); one JSON line per N in gpurun_out/.
for n in 2 4 8; do
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29520 + n)) bench.py --gpus $n --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/s12_scale_n$n.json 2> gpurun_out/s12_scale_n$n.err
  echo "N=$n rc=$? stdout lines: $(wc -l < gpurun_out/s12_scale_n$n.json)"; cut -c1-160 gpurun_out/s12_scale_n$n.json
done
