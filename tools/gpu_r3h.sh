#!/bin/bash
# Round 3 final evidence on one 8-GPU lease: the whole GPU suite (no multi-GPU test skipped), smoke(), bench at
# N = 1 and N = 8 with the in-run parity check and the same-workload single-GPU point.
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/r3h_gpus.txt
rm -f gpurun_out/grad_report.txt
timeout 1500 python -m pytest tests -q -m gpu -rs > gpurun_out/r3h_pytest_gpu.log 2>&1
echo "suite rc=$?" >> gpurun_out/r3h_pytest_gpu.log
tail -5 gpurun_out/r3h_pytest_gpu.log
cp gpurun_out/grad_report.txt gpurun_out/r3h_grad_report.txt
python __graft_entry__.py smoke > gpurun_out/r3h_smoke.log 2>&1; tail -1 gpurun_out/r3h_smoke.log
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/r3h_bench_n1.json 2> gpurun_out/r3h_bench_n1.err
echo "bench n=1 rc=$?"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29618 \
    bench.py --gpus 8 --steps 20 --warmup 5 > gpurun_out/r3h_bench_n8.json 2> gpurun_out/r3h_bench_n8.err
echo "bench n=8 rc=$?"
python - <<'PY'
import json
for n in (1, 8):
    try:
        l = json.loads(open(f"gpurun_out/r3h_bench_n{n}.json").read().strip().splitlines()[-1])
        c = l["config"]
        print(f"N={n}", round(l["value"], 1), "views/s", l["scaling"], "| e2e", round(l["e2e"]["value"], 1), "|", c.get("sharded_parity"),
              "| n1 same workload", c.get("n1_same_workload_views_per_s"), "eff", c.get("efficiency_vs_n1_same_workload"),
              "e2e eff", c.get("e2e_efficiency_vs_n1_same_workload_device_timed"), "| fwd-only", c.get("forward_only_fps_sh3"), c.get("forward_only_fps_sh0"),
              "| n1 mvb", c.get("n1_multi_view_batch_views_per_s"))
    except Exception as e:
        print(f"N={n}: no line:", e)
        print(open(f"gpurun_out/r3h_bench_n{n}.err").read()[-2500:])
PY
