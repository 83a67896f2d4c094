#!/usr/bin/env bash
# Multi-GPU records on one box: tools/multi_gpu_runs.sh <max N> <tag>
# (the canvas-tiled cfg3 run at N = 2 .. max, and the default group-sharded
# cfg2 run at the same N), outputs under gpurun_out/.
set -u
MAXN="$1"; TAG="$2"
export XP_BENCH_CACHE=/tmp/xp_wl
mkdir -p gpurun_out
python - <<'PY'
import bench
for c, g in (("cfg3", [0]), ("cfg2", list(range(8)))):
    for grp in g:
        bench.make_workload(bench.CONFIGS[c], grp, "graphcut")
PY
port=29500
python bench.py --config cfg3 --tiled --steps 5 --warmup 3 --no-cpu \
  > gpurun_out/${TAG}_cfg3_tiled_n1.json 2> gpurun_out/${TAG}_cfg3_tiled_n1.err
for n in 2 4 8; do
  [ "$n" -le "$MAXN" ] || continue
  port=$((port + 1))
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $n \
    --master-addr 127.0.0.1 --master-port $port bench.py --gpus $n \
    --config cfg3 --tiled --steps 5 --warmup 3 \
    > gpurun_out/${TAG}_cfg3_tiled_n$n.json 2> gpurun_out/${TAG}_cfg3_tiled_n$n.err
  echo "tiled n=$n rc=$?"; tail -c 300 gpurun_out/${TAG}_cfg3_tiled_n$n.err
  port=$((port + 1))
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $n \
    --master-addr 127.0.0.1 --master-port $port bench.py --gpus $n \
    --steps 10 --warmup 3 \
    > gpurun_out/${TAG}_cfg2_groups_n$n.json 2> gpurun_out/${TAG}_cfg2_groups_n$n.err
  echo "groups n=$n rc=$?"; tail -c 200 gpurun_out/${TAG}_cfg2_groups_n$n.err
done
