# usage: bash tools/scale_ab.sh N tag mode...   (scratch: N-GPU parity check + bench per multi-GPU mode)
N=$1; TAG=$2; shift 2
for ex in peer nccl; do
GSR_EXCHANGE=$ex timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 tools/sharded_check.py > gpurun_out/${TAG}_check${N}_$ex.log 2>&1; grep "sharded_check\|FAIL" gpurun_out/${TAG}_check${N}_$ex.log | head -12
done
for m in "$@"; do
if [ "$m" = "single" ]; then
python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/${TAG}_single.json 2> gpurun_out/${TAG}_single.err; f=gpurun_out/${TAG}_single
else
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 20 --warmup 5 --multi $m > gpurun_out/${TAG}_scale${N}_$m.json 2> gpurun_out/${TAG}_scale${N}_$m.err; f=gpurun_out/${TAG}_scale${N}_$m
fi
echo exit $?
python - <<PY
import json
try:
    d=json.loads(open("$f.json").read().strip().splitlines()[-1])
    print("$m", d["n_gpus"], "value", round(d["value"],1), "ms", round(d["ms_per_step"],3), "e2e", round(d["e2e"]["value"],1), round(d["e2e"]["ms_per_step"],3), {k:round(v["ms"],3) for k,v in d["stages"].items()})
except Exception as e:
    print("parse fail", e); print(open("$f.err").read()[-2500:])
PY
done
