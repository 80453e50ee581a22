#!/bin/bash
# usage: tools/gpu_ncu_bin.sh TAG REGEX COUNT [SKIP]  — one ncu --set full capture of the kernels matching REGEX in one bench step
TAG=$1; REGEX=$2; COUNT=$3; SKIP=${4:-$((3*COUNT))}
OUT=gpurun_out; mkdir -p $OUT
BENCH="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-extras"
$BENCH > $OUT/${TAG}_plain.log 2>&1 || { tail -5 $OUT/${TAG}_plain.log; exit 1; }
ncu --set full --clock-control none --import-source on -k regex:"$REGEX" -s $SKIP -c $COUNT -f -o $OUT/${TAG} $BENCH > $OUT/${TAG}_ncu.log 2>&1
echo "ncu exit $?"; tail -3 $OUT/${TAG}_ncu.log; ls -la $OUT/${TAG}.ncu-rep
