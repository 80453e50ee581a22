# usage: bash tools/ab_env.sh VAR v1 v2 ...   (scratch: parity tests + stage times of bench config 3 per value of an env knob)
VAR=$1; shift
for v in "$@"; do
  echo "== $VAR=$v"
  env $VAR=$v python -m pytest tests/test_gpu_parity.py tests/test_gpu_fullsize.py tests/test_gpu_variants.py -m gpu -x -q 2>&1 | tail -2
  for i in 1 2; do
    env $VAR=$v python bench.py --steps 30 --warmup 5 --no-cpu-baseline --no-extras 2>/dev/null | python -c "
import sys,json; d=json.loads(sys.stdin.read()); print(round(d['ms_per_step'],4), {k:round(v['ms'],3) for k,v in d['stages'].items()})"
  done
done
