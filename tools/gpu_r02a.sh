#!/bin/bash
# Round-2 first GPU call: reference-CUDA probe, SIMT peak microbenchmark, bench line of the unchanged
# round-1 code, and the --set full capture of its final kernels (the capture VERDICT r01 found missing).
set -u
TAG=${1:-r02a}
OUT=gpurun_out
mkdir -p $OUT build/tools
python tools/probe_reference_cuda.py > $OUT/${TAG}_probe.json 2>&1; cat $OUT/${TAG}_probe.json
pip list 2>/dev/null | grep -i -E "gsplat|gaussian|raster|nerf|splat" > $OUT/${TAG}_pip_grep.txt; cat $OUT/${TAG}_pip_grep.txt
[ -x build/tools/fp32_peak ] || nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o build/tools/fp32_peak tools/fp32_peak.cu
nvidia-smi --query-gpu=clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active --format=csv -lms 100 > $OUT/${TAG}_fp32_clocks.csv &
SMI=$!
build/tools/fp32_peak > $OUT/${TAG}_fp32_peak.json 2>&1; cat $OUT/${TAG}_fp32_peak.json
kill $SMI
python bench.py --steps 20 --warmup 5 --no-extras > $OUT/${TAG}_bench.json 2> $OUT/${TAG}_bench.err
echo "bench exit $?"; head -c 900 $OUT/${TAG}_bench.json; echo
python bench.py --steps 10 --warmup 3 --no-extras --no-cpu-baseline --points 6000000 > $OUT/${TAG}_bench6m.json 2> $OUT/${TAG}_bench6m.err
echo "bench6m exit $?"; head -c 400 $OUT/${TAG}_bench6m.json; echo
BENCH="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-extras"
LPS=16
$BENCH > $OUT/${TAG}_plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on \
    -k regex:'blend_|radix_|preprocess|emit_keys|tile_ranges|scan_|publish_' -s $((3*LPS)) -c $LPS \
    -f -o $OUT/${TAG}_full $BENCH > $OUT/${TAG}_ncu_full.log 2>&1
echo "ncu full exit $?"
ls -la $OUT | tail -12
