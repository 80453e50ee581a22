#!/bin/bash
OUT=gpurun_out; mkdir -p $OUT
timeout 600 python -m pytest tests/test_gpu_raw.py tests/test_gpu_train_iteration.py tests/test_gpu_parity.py -m gpu -q -x --timeout 500 2>&1 | tail -4
for v in 1 0; do
GSR_BACKWARD_TMA_RAW=$v timeout 400 python bench.py --steps 20 --warmup 5 --no-cpu-baseline 2>/dev/null | python -c "
import sys,json; d=json.loads(sys.stdin.read()); print('raw_tma=$v', round(d['ms_per_step'],4), 'raw_entry', d['raw_entry']['fused_raw_ms'], 'train_iteration', d['train_iteration']['fused_ms'], d['train_iteration']['torch_ms'])"
done
