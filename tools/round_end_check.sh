set -x
python -m pytest tests -m gpu -x -q > gpurun_out/s12f_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/s12f_pytest.log
python bench.py > gpurun_out/s12f_bench.json 2> gpurun_out/s12f_bench.err; echo "bench rc=$?"; cat gpurun_out/s12f_bench.json
python bench.py --impl reference --steps 1 --warmup 0 > gpurun_out/s12f_ref.json 2> gpurun_out/s12f_ref.err; echo "ref rc=$?"; cat gpurun_out/s12f_ref.json
python tools/profile_step.py --batch 8 --out gpurun_out/s12f_step_b8.txt > /dev/null 2>&1; head -1 gpurun_out/s12f_step_b8.txt
python tools/profile_step.py --batch 1 --out gpurun_out/s12f_step_b1.txt > /dev/null 2>&1; head -1 gpurun_out/s12f_step_b1.txt
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 1330 --csv --log-file gpurun_out/s12f_launches.csv python bench.py --steps 1 --warmup 1 > gpurun_out/s12f_ncu1.log 2>&1; echo "ncu list rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:attention_kernel -c 1 -f -o gpurun_out/s12f_att python tools/bench_ops.py --only attn_self_4096_d40 --reps 1 > gpurun_out/s12f_ncu2.log 2>&1; echo "ncu att rc=$?"
ls -la gpurun_out | tail -12
