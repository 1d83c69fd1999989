#!/bin/bash
# One 8-GPU box: default bench (weak scaling + the strong-scaled leg) and the configs[4] replica sweep at N = 1, 2, 4, 8.
# N = 1, 2, 4 run side by side on disjoint GPUs (0 | 1-2 | 3-6), then N = 8 alone.   usage: tools/scale_session.sh <tag>
set -u
TAG=${1:-scale}
OUT=gpurun_out
mkdir -p $OUT
SWEEP=1024,4096,16384,65536,262144
ARGS="--steps 5 --warmup 3 --no-cpu --no-extra --sweep $SWEEP"
python -c "import torch" 2>/dev/null  # page the image in once
CUDA_VISIBLE_DEVICES=0 python bench.py --gpus 1 $ARGS > $OUT/${TAG}_1.json 2> $OUT/${TAG}_1.err &
P1=$!
CUDA_VISIBLE_DEVICES=1,2 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 $ARGS > $OUT/${TAG}_2.json 2> $OUT/${TAG}_2.err &
P2=$!
CUDA_VISIBLE_DEVICES=3,4,5,6 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29514 bench.py --gpus 4 $ARGS > $OUT/${TAG}_4.json 2> $OUT/${TAG}_4.err &
P4=$!
wait $P1 $P2 $P4
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29518 bench.py --gpus 8 $ARGS > $OUT/${TAG}_8.json 2> $OUT/${TAG}_8.err
for n in 1 2 4 8; do echo "== N=$n"; python - <<PY
import json
try:
    d=json.loads(open("$OUT/${TAG}_$n.json").read().strip().splitlines()[-1])
    print("weak", d["value"], "strong", d.get("extra",{}).get("strong_scaling",{}).get("value"))
    for s in d.get("extra",{}).get("replica_sweep",[]): print("  sweep", s["replicas_total"], s["replicas_per_gpu"], "%.3e"%s["value"])
except Exception as e: print("failed", e)
PY
done
