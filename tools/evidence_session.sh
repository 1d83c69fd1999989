#!/bin/bash
# Round-2 evidence for profiles/ on one GPU: parity report, drift at the config's size, both bench arms, the forces
# workload, launch list + full ncu captures.   usage: tools/evidence_session.sh <tag>
set -u
TAG=${1:-ev}
OUT=gpurun_out
mkdir -p $OUT
python tools/parity_report.py > $OUT/${TAG}_parity.md 2> $OUT/${TAG}_parity.err; tail -3 $OUT/${TAG}_parity.md
python tools/nve_drift.py > $OUT/${TAG}_nve_drift.jsonl 2> $OUT/${TAG}_nve_drift.err; cut -c1-400 $OUT/${TAG}_nve_drift.jsonl
python bench.py --impl reference > $OUT/${TAG}_refarm.json 2> $OUT/${TAG}_refarm.err; cut -c1-160 $OUT/${TAG}_refarm.json
python bench.py > $OUT/${TAG}_bench.json 2> $OUT/${TAG}_bench.err; cut -c1-200 $OUT/${TAG}_bench.json
for mode in plain rf; do
  python bench.py --workload forces --forces-mode $mode > $OUT/${TAG}_forces_$mode.json 2> $OUT/${TAG}_forces_$mode.err
  python bench.py --workload forces --forces-mode $mode --no-cpu --no-energies > $OUT/${TAG}_forces_${mode}_noene.json 2>> $OUT/${TAG}_forces_$mode.err
  cut -c1-130 $OUT/${TAG}_forces_$mode.json; cut -c1-130 $OUT/${TAG}_forces_${mode}_noene.json
done
python bench.py --impl reference --workload forces > $OUT/${TAG}_refarm_forces.json 2>> $OUT/${TAG}_refarm.err
python tools/profile_energy.py > $OUT/${TAG}_energy_plain.log 2>&1; python tools/profile_energy.py rf > $OUT/${TAG}_energy_rf.log 2>&1
bash tools/profile_session.sh $TAG
