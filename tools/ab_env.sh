# Same-box A/B of an environment switch on the sustained job rate: ab_env.sh VAR v1 v2 ... (each value twice, interleaved)
var=$1; shift
for rep in $(seq 1 ${REPS:-2}); do for v in "$@"; do
  echo "=== $var=$v"
  env $var=$v python bench.py --steps 4 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['clocks']['sm_mhz'])"
done; done
