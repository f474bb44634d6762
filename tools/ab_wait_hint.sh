# A/B of the mbarrier.try_wait suspend-time hint on the sustained (power-capped) job rate; same box, back to back
for h in "-DB200SD_WAIT_HINT_NS=0" "-DB200SD_WAIT_HINT_NS=20000" "-DB200SD_WAIT_HINT_NS=0" "-DB200SD_WAIT_HINT_NS=1000000"; do
  B200SD_EXTRA_NVCC_FLAGS="$h" bash stablediffusion_b200/csrc/build.sh >/dev/null 2>&1
  echo "=== flags: $h"
  python bench.py --steps 4 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['clocks'])"
done
bash stablediffusion_b200/csrc/build.sh >/dev/null 2>&1
