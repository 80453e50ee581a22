#!/bin/bash
OUT=gpurun_out; mkdir -p $OUT; TAG=${1:-r02z}
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 \
    bench.py --gpus 8 --steps 20 --warmup 20 > $OUT/${TAG}_scale8.json 2> $OUT/${TAG}_scale8.err
python - <<PY
import json
try:
    d=json.loads(open("$OUT/${TAG}_scale8.json").read().strip().splitlines()[-1])
    print(8, "value", round(d["value"],1), "ms/step", round(d["ms_per_step"],3), "median", round(d["ms_per_step_median"],3), "e2e", round(d["e2e"]["value"],1), round(d["e2e"]["ms_per_step"],3))
    print("   ", d.get("ms_per_step_rank0"))
except Exception as e:
    print("parse fail", e); print(open("$OUT/${TAG}_scale8.err").read()[-1500:])
PY
