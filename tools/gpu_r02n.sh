#!/bin/bash
# Full single-GPU evidence run: every GPU test, the default bench line, the ncu launch list and one --set full
# capture of every kernel of one step (recipe of /opt/skills/guides/B200_PROFILING.md).
set -u
TAG=${1:-r02n}; LPS=${2:-13}
OUT=gpurun_out; mkdir -p $OUT; rm -f $OUT/grad_report.txt
timeout 1500 python -m pytest tests -m gpu -q --timeout 900 > $OUT/${TAG}_pytest_gpu.log 2>&1
echo "pytest exit $?" | tee -a $OUT/${TAG}_pytest_gpu.log; tail -6 $OUT/${TAG}_pytest_gpu.log
cp $OUT/grad_report.txt $OUT/${TAG}_grad_report.txt 2>/dev/null
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
timeout 900 python bench.py --steps 20 --warmup 5 > $OUT/${TAG}_bench.json 2> $OUT/${TAG}_bench.err
echo "bench exit $?"; head -c 700 $OUT/${TAG}_bench.json; echo; tail -3 $OUT/${TAG}_bench.err
BENCH="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-extras"
$BENCH > $OUT/${TAG}_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/${TAG}_launches.csv $BENCH > $OUT/${TAG}_ncu_launches.log 2>&1
echo "ncu launch list exit $?"
$BENCH > $OUT/${TAG}_plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'blend_|radix_|preprocess|emit_|depth_hist|tile_offsets' \
    -s $((3*LPS)) -c $LPS -f -o $OUT/${TAG}_full $BENCH > $OUT/${TAG}_ncu_full.log 2>&1
echo "ncu full exit $?"; ls -la $OUT | grep ${TAG}
