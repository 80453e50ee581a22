#!/bin/bash
# 8-GPU call, short: the configs[3] bench at N = 8 (twice) and N = 4
OUT=gpurun_out; mkdir -p $OUT; TAG=${1:-r02y}
run() {
  timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $1 --master-addr 127.0.0.1 --master-port 29511 \
      bench.py --gpus $1 --steps 20 --warmup 20 > $OUT/${TAG}_scale$1$2.json 2> $OUT/${TAG}_scale$1$2.err
  python - <<PY
import json
try:
    d=json.loads(open("$OUT/${TAG}_scale$1$2.json").read().strip().splitlines()[-1])
    print($1, "$2", "value", round(d["value"],1), "ms/step", round(d["ms_per_step"],3), "median", round(d["ms_per_step_median"],3), "e2e", round(d["e2e"]["value"],1), round(d["e2e"]["ms_per_step"],3))
    print("   ", {k:round(v["ms"],3) for k,v in d["stages"].items()}); print("   ", d.get("ms_per_step_rank0"))
except Exception as e:
    print("parse fail", e); print(open("$OUT/${TAG}_scale$1$2.err").read()[-1500:])
PY
}
run 8 ""
run 8 _b
run 4 ""
