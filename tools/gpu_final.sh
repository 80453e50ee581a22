#!/bin/bash
# Reduced evidence pass (no --set full capture): GPU tests, contract bench, ncu launch list.  usage: bash tools/gpu_final.sh TAG
TAG=${1:-final}; OUT=gpurun_out; mkdir -p $OUT
python -m pytest tests -m gpu -q > $OUT/${TAG}_pytest_gpu.log 2>&1; echo "pytest exit $?"; tail -2 $OUT/${TAG}_pytest_gpu.log
python bench.py --steps 20 --warmup 5 > $OUT/${TAG}_bench.json 2> $OUT/${TAG}_bench.err; echo "bench exit $?"
BENCH="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-extras"
ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file $OUT/${TAG}_launches.csv $BENCH > $OUT/${TAG}_ncu_launches.log 2>&1; echo "ncu exit $?"
