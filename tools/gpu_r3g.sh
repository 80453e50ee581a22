#!/bin/bash
# Round 3, multi-GPU evidence on one >= 4-GPU lease: ALL of tests/test_gpu_sharded.py (world 1 / 2 / 4, the peer
# accumulator ring on whenever the accumulators sit on a peer), then bench.py at N = 2 and N = 4 with its in-run
# parity check (config.sharded_parity) and same-workload single-GPU point (config.n1_same_workload_views_per_s).
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/r3g_gpus.txt
timeout 1200 python -m pytest tests/test_gpu_sharded.py tests/test_gpu_multiview_kernels.py -q -m gpu -rs > gpurun_out/r3g_pytest_sharded.log 2>&1
echo "rc=$?" >> gpurun_out/r3g_pytest_sharded.log
tail -6 gpurun_out/r3g_pytest_sharded.log
for n in 2 4; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29600 + n)) \
      bench.py --gpus $n --steps 20 --warmup 5 > gpurun_out/r3g_bench_n$n.json 2> gpurun_out/r3g_bench_n$n.err
  echo "bench n=$n rc=$?"
  python - <<PY
import json
try:
    l = json.loads(open("gpurun_out/r3g_bench_n$n.json").read().strip().splitlines()[-1])
    c = l["config"]
    print("N=$n", round(l["value"], 1), "views/s", l["scaling"], "| e2e", round(l["e2e"]["value"], 1), "|", c.get("sharded_parity"),
          "| n1 same workload", c.get("n1_same_workload_views_per_s"), "eff", c.get("efficiency_vs_n1_same_workload"))
except Exception as e:
    print("N=$n: no line:", e)
    print(open("gpurun_out/r3g_bench_n$n.err").read()[-2500:])
PY
done
