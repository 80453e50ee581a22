#!/bin/bash
# usage: tools/gpu_env_ab.sh TAG "pytest targets|-" VAR v1 v2 ... — tests with the default environment, then config-3 stage times per value of VAR
TAG=$1; TESTS=$2; VAR=$3; shift 3
OUT=gpurun_out; mkdir -p $OUT
if [ "$TESTS" != "-" ]; then
  timeout 1500 python -m pytest $TESTS -m gpu -q -x --timeout 900 > $OUT/${TAG}_pytest.log 2>&1
  echo "pytest exit $?"; tail -15 $OUT/${TAG}_pytest.log
fi
for v in "$@"; do
  for i in 1 2; do
    env $VAR=$v timeout 300 python bench.py --steps 30 --warmup 5 --no-cpu-baseline --no-extras 2> $OUT/${TAG}_ab.err | python -c "
import sys,json; d=json.loads(sys.stdin.read()); print('$VAR=$v', round(d['ms_per_step'],4), 'e2e', round(d['e2e']['ms_per_step'],3), {k:round(v['ms'],3) for k,v in d['stages'].items()})" || tail -5 $OUT/${TAG}_ab.err
  done
done 2>&1 | tee $OUT/${TAG}_ab.txt
