#!/bin/bash
OUT=gpurun_out; mkdir -p $OUT; TAG=${1:-r02u}
timeout 1200 python -m pytest tests/test_gpu_sharded.py tests/test_gpu_multiview_kernels.py -m gpu -q -x --timeout 900 2>&1 | tail -4
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 20 > $OUT/${TAG}_scale2.json 2> $OUT/${TAG}_scale2.err
python - <<PY
import json
d=json.loads(open("$OUT/${TAG}_scale2.json").read().strip().splitlines()[-1])
print(2, "value", round(d["value"],1), "ms/step", round(d["ms_per_step"],3), "median", d.get("ms_per_step_median"), "e2e", round(d["e2e"]["value"],1), round(d["e2e"]["ms_per_step"],3))
print({k:round(v["ms"],3) for k,v in d["stages"].items()})
PY
timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-extras --views 8 --points 6000000 > $OUT/${TAG}_scale1.json 2> $OUT/${TAG}_scale1.err
python - <<PY
import json
d=json.loads(open("$OUT/${TAG}_scale1.json").read().strip().splitlines()[-1])
print(1, "value", round(d["value"],1), "ms/step", round(d["ms_per_step"],3), "e2e", round(d["e2e"]["value"],1), round(d["e2e"]["ms_per_step"],3))
print({k:round(v["ms"],3) for k,v in d["stages"].items()})
PY
