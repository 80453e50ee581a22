#!/bin/bash
# diagnostic: e2e ms/step with pieces of the e2e loop removed (GSR_E2E_SKIP), and the blend-prefetch A/B
OUT=gpurun_out
for sk in "" h2d mul d2h h2d,mul,d2h; do
  GSR_E2E_SKIP=$sk timeout 300 python bench.py --steps 30 --warmup 5 --no-cpu-baseline --no-extras 2>/dev/null | python -c "
import sys,json; d=json.loads(sys.stdin.read()); print('skip=[$sk]', 'device', round(d['ms_per_step'],4), 'e2e', round(d['e2e']['ms_per_step'],3), 'h2d_ms', d['e2e'].get('h2d_ms_per_step'))"
done 2>&1 | tee $OUT/r02o_e2e_skip.txt
bash tools/gpu_ab.sh r02o - pf6 pf5
