#!/bin/bash
# The reference's launch mode on one machine: one parameter-server task and N worker tasks of examples/distribute.py,
# each started with --job_name / --task_index like on a real cluster (no launcher). Workers take GPU task_index when
# --gpu_num_per_worker=1 is passed through; without it everything runs on the CPU.
#   examples/launch_local_cluster.sh 2                              # ps + 2 CPU workers
#   examples/launch_local_cluster.sh 8 --gpu_num_per_worker=1       # ps + 8 workers, one B200 each
set -e
N=${1:-2}; shift || true
cd "$(dirname "$0")/.."
BASE=${DTF_BASE_PORT:-23450}
PS="127.0.0.1:$BASE"
WORKERS=$(for i in $(seq 1 "$N"); do printf "127.0.0.1:%d," $((BASE + i)); done | sed 's/,$//')
MODEL_DIR=${DTF_MODEL_DIR:-$(mktemp -d /tmp/dtf_cluster_XXXXXX)}
COMMON="--ps_hosts=$PS --worker_hosts=$WORKERS --model_dir=$MODEL_DIR $*"
python examples/distribute.py --job_name=ps --task_index=0 $COMMON &
PIDS=$!
for i in $(seq 0 $((N - 1))); do
  python examples/distribute.py --job_name=worker --task_index="$i" $COMMON &
  PIDS="$PIDS $!"
done
RC=0
for p in $PIDS; do wait "$p" || RC=$?; done
echo "cluster finished (exit $RC); checkpoints in $MODEL_DIR"
exit $RC
