#!/bin/bash
# Round-2 multi-GPU call (N = 2, 4 or 8 GPUs): the sharded-step parity tests on real peer memory, then the
# configs[3] bench (6 M Gaussians, 8-view batch split over N ranks) for every world size up to N.
set -u
N=${1:-2}; TAG=${2:-r02c}
OUT=gpurun_out
mkdir -p $OUT
nvidia-smi topo -m > $OUT/${TAG}_topo.txt 2>&1
timeout 900 python -m pytest tests/test_gpu_sharded.py -m gpu -q --timeout 600 > $OUT/${TAG}_pytest_sharded.log 2>&1
echo "pytest exit $?"; tail -5 $OUT/${TAG}_pytest_sharded.log
for n in 2 4 8; do
  [ $n -le $N ] || continue
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 \
      bench.py --gpus $n --steps 20 --warmup 20 > $OUT/${TAG}_scale${n}.json 2> $OUT/${TAG}_scale${n}.err
  echo "bench N=$n exit $?"
  python - <<PY
import json
try:
    d=json.loads(open("$OUT/${TAG}_scale${n}.json").read().strip().splitlines()[-1])
    print($n, "value", round(d["value"],1), "ms/step", round(d["ms_per_step"],3), "median", d.get("ms_per_step_median"), "e2e", round(d["e2e"]["value"],1), round(d["e2e"]["ms_per_step"],3))
    print({k:round(v["ms"],3) for k,v in d["stages"].items()})
    print(d["config"]["workload"]); print(json.dumps(d.get("per_rank"))[:1500])
except Exception as e:
    print("parse fail", e); print(open("$OUT/${TAG}_scale${n}.err").read()[-2500:])
PY
done
timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-extras --views 8 --points 6000000 > $OUT/${TAG}_scale1.json 2> $OUT/${TAG}_scale1.err
echo "bench N=1 (8 views, 6M) exit $?"; head -c 700 $OUT/${TAG}_scale1.json; echo
