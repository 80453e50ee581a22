# usage: bash tools/ab_libs.sh lib1.so lib2.so ...   (scratch: stage times of bench config 3 per library variant)
for lib in "$@"; do
  for i in 1 2; do
    GSRAST_B200_LIB=$lib python bench.py --steps 30 --warmup 5 --no-cpu-baseline --no-extras 2>/dev/null | python -c "
import sys,json; d=json.loads(sys.stdin.read()); print('$lib', round(d['ms_per_step'],4), {k:round(v['ms'],3) for k,v in d['stages'].items()})"
  done
done
