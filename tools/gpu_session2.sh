#!/bin/bash
# Round-2 gpurun session: GPU tests, default bench, forces workload, small-R sweep, launch list, full ncu capture.
# usage: tools/gpu_session2.sh <tag> [pytest-args]
set -u
TAG=${1:-run}
OUT=gpurun_out
mkdir -p $OUT
nvidia-smi --query-gpu=index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active --format=csv -lms 500 > $OUT/${TAG}_clocks.csv &
SMI=$!
timeout 1500 python -m pytest tests -m gpu -q ${PYTEST_ARGS:--x} > $OUT/${TAG}_pytest.log 2>&1; echo "pytest exit $?" | tee -a $OUT/${TAG}_pytest.log
tail -15 $OUT/${TAG}_pytest.log
if [ "${SKIP_BENCH:-0}" != "1" ]; then
timeout 900 python bench.py > $OUT/${TAG}_bench.json 2> $OUT/${TAG}_bench.err; echo "bench exit $?"
head -c 3000 $OUT/${TAG}_bench.json; echo
for mode in plain rf; do
  timeout 300 python bench.py --workload forces --forces-mode $mode --no-cpu > $OUT/${TAG}_forces_$mode.json 2> $OUT/${TAG}_forces_$mode.err
  timeout 300 python bench.py --workload forces --forces-mode $mode --no-cpu --no-energies > $OUT/${TAG}_forces_${mode}_noene.json 2>> $OUT/${TAG}_forces_$mode.err
  cut -c1-400 $OUT/${TAG}_forces_$mode.json; cut -c1-300 $OUT/${TAG}_forces_${mode}_noene.json
done
for R in 1024 4096; do
  timeout 300 python bench.py --no-cpu --no-extra --strong-total 0 --steps 5 --warmup 3 --replicas $R > $OUT/${TAG}_sweep_$R.json 2> $OUT/${TAG}_sweep_$R.err
  cut -c1-200 $OUT/${TAG}_sweep_$R.json
done
fi
kill $SMI
if [ "${SKIP_NCU:-0}" != "1" ]; then
CMD="python bench.py --steps 3 --warmup 3 --inner 50 --no-cpu --no-extra --strong-total 0"
timeout 300 $CMD > $OUT/${TAG}_plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file $OUT/${TAG}_launches.csv $CMD > $OUT/${TAG}_ncu_l.log 2>&1
timeout 300 $CMD > $OUT/${TAG}_plain2.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:integrate_kernel -s 3 -c 1 -f -o $OUT/${TAG}_prof $CMD > $OUT/${TAG}_ncu_f.log 2>&1
echo "ncu exit $?"
fi
