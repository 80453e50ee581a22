#!/bin/bash
# One gpurun call: GPU parity tests, the contract bench, then the two ncu passes of the profiling
# recipe (launch list + one --set full capture of every kernel of one step).  Outputs -> gpurun_out/.
#   gpurun --timeout 1500 -- 'bash tools/gpu_profile.sh [tag]'
set -u
TAG=${1:-r01}
LPS=${2:-16}   # kernel launches of libgsrast_b200.so per fwd+bwd step
OUT=gpurun_out
mkdir -p $OUT
BENCH="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-extras"

python -m pytest tests -m gpu -x -q > $OUT/${TAG}_pytest_gpu.log 2>&1
echo "pytest exit $?" | tee -a $OUT/${TAG}_pytest_gpu.log
tail -3 $OUT/${TAG}_pytest_gpu.log

python bench.py --steps 20 --warmup 5 > $OUT/${TAG}_bench.json 2> $OUT/${TAG}_bench.err
echo "bench exit $?"; tail -c 600 $OUT/${TAG}_bench.json

$BENCH > $OUT/${TAG}_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv \
    --log-file $OUT/${TAG}_launches.csv $BENCH > $OUT/${TAG}_ncu_launches.log 2>&1
echo "ncu launch list exit $?"

# kernels of OUR library only; skip the 3 warm-up steps (LPS matching launches per step)
$BENCH > $OUT/${TAG}_plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on \
    -k regex:'blend_|radix_|preprocess|emit_keys|tile_ranges|scan_|publish_' -s $((3*LPS)) -c $LPS \
    -f -o $OUT/${TAG}_full $BENCH > $OUT/${TAG}_ncu_full.log 2>&1
echo "ncu full exit $?"
ls -la $OUT
