#!/bin/bash
# Round-2 second GPU call (1 GPU): the whole GPU test suite with the new parity tests, the default bench line
# (with the configs[3]/[4] extras and both CPU baselines), the host-side probe of the e2e step.
set -u
TAG=${1:-r02b}
OUT=gpurun_out
mkdir -p $OUT
rm -f $OUT/grad_report.txt
timeout 1500 python -m pytest tests -m gpu -q --timeout 900 > $OUT/${TAG}_pytest_gpu.log 2>&1
echo "pytest exit $?" | tee -a $OUT/${TAG}_pytest_gpu.log
tail -25 $OUT/${TAG}_pytest_gpu.log
cp $OUT/grad_report.txt $OUT/${TAG}_grad_report.txt 2>/dev/null; cat $OUT/${TAG}_grad_report.txt 2>/dev/null
timeout 900 python bench.py --steps 20 --warmup 5 > $OUT/${TAG}_bench.json 2> $OUT/${TAG}_bench.err
echo "bench exit $?"; head -c 1500 $OUT/${TAG}_bench.json; echo; tail -5 $OUT/${TAG}_bench.err
timeout 300 python tools/host_probe.py > $OUT/${TAG}_host_probe.txt 2>&1; head -40 $OUT/${TAG}_host_probe.txt
