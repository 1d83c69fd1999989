#!/bin/bash
# ncu evidence for profiles/: launch list of the bench command, full captures of the fused integrator (plain and cutoff
# mode), the single-shot force kernel and the FP64 energy kernel.  usage: tools/profile_session.sh <tag>
set -u
TAG=${1:-prof}
OUT=gpurun_out
mkdir -p $OUT
CMD="python bench.py --steps 3 --warmup 3 --inner 50 --no-cpu --no-extra --strong-total 0"
timeout 300 $CMD > $OUT/${TAG}_plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 80 --csv --log-file $OUT/${TAG}_launches.csv $CMD > $OUT/${TAG}_ncu_l.log 2>&1
timeout 300 $CMD > $OUT/${TAG}_plain2.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:integrate_kernel -s 3 -c 1 -f -o $OUT/${TAG}_integrate $CMD > $OUT/${TAG}_ncu_i.log 2>&1
echo "ncu integrate exit $?"
timeout 300 python tools/profile_forces.py > $OUT/${TAG}_pf.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:forces_kernel -s 4 -c 1 -f -o $OUT/${TAG}_forces python tools/profile_forces.py > $OUT/${TAG}_ncu_f.log 2>&1
echo "ncu forces exit $?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:energy_kernel -s 4 -c 1 -f -o $OUT/${TAG}_energy python tools/profile_forces.py > $OUT/${TAG}_ncu_e.log 2>&1
echo "ncu energy exit $?"
cat $OUT/${TAG}_pf.log
