# Round-end verification on one B200: full GPU test suite, smoke, the three bench workloads, the reference arm and per-op step tables.
set -x
python -m pytest tests -m gpu -x -q > gpurun_out/s12j_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/s12j_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/s12j_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/s12j_smoke.log
python bench.py > gpurun_out/s12j_bench.json 2> gpurun_out/s12j_bench.err; echo "bench rc=$?"; cat gpurun_out/s12j_bench.json
python bench.py --workload sd15_512_b1_euler20_cfg --no-cpu-baseline > gpurun_out/s12j_bench_b1.json 2>/dev/null; echo "b1 rc=$?"; cat gpurun_out/s12j_bench_b1.json
python bench.py --workload sdxl_1024_b4_euler30_cfg --no-cpu-baseline > gpurun_out/s12j_bench_sdxl.json 2>/dev/null; echo "sdxl rc=$?"; cat gpurun_out/s12j_bench_sdxl.json
python tools/profile_step.py --batch 8 --out gpurun_out/s12j_step_b8.txt > /dev/null 2>&1; head -1 gpurun_out/s12j_step_b8.txt
python tools/profile_step.py --batch 1 --out gpurun_out/s12j_step_b1.txt > /dev/null 2>&1; head -1 gpurun_out/s12j_step_b1.txt
python tools/profile_step.py --model sdxl --res 1024 --batch 4 --out gpurun_out/s12j_step_sdxl.txt > /dev/null 2>&1; head -1 gpurun_out/s12j_step_sdxl.txt
