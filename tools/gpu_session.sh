#!/bin/bash
# One gpurun session: GPU tests, bench, launch list, full ncu capture of the fused kernel.
# usage: tools/gpu_session.sh <tag>
set -u
TAG=${1:-run}
OUT=gpurun_out
mkdir -p $OUT
nvidia-smi --query-gpu=index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active --format=csv -lms 500 > $OUT/${TAG}_clocks.csv &
SMI=$!
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/${TAG}_pytest.log 2>&1; echo "pytest exit $?" | tee -a $OUT/${TAG}_pytest.log
tail -5 $OUT/${TAG}_pytest.log
timeout 600 python bench.py > $OUT/${TAG}_bench.json 2> $OUT/${TAG}_bench.err; echo "bench exit $?"
cat $OUT/${TAG}_bench.json
kill $SMI
if [ "${SKIP_NCU:-0}" != "1" ]; then
CMD="python bench.py --steps 3 --warmup 3 --inner 50 --no-cpu --no-extra"
timeout 300 $CMD > $OUT/${TAG}_plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file $OUT/${TAG}_launches.csv $CMD > $OUT/${TAG}_ncu_l.log 2>&1
timeout 300 $CMD > $OUT/${TAG}_plain2.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:integrate_kernel -s 3 -c 1 -f -o $OUT/${TAG}_prof $CMD > $OUT/${TAG}_ncu_f.log 2>&1
echo "ncu exit $?"
fi
