#!/bin/bash
# Where does the N-GPU step lose time against the 1-GPU step? Collective variants / bucket sizes / CTA counts at N ranks.
N=${1:-2}
run() { # tag, env..., extra args after --
  tag=$1; shift
  envs=(); while [ "$1" != "--" ]; do envs+=("$1"); shift; done; shift
  env "${envs[@]}" timeout 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29600 + RANDOM % 300)) bench.py --gpus $N --steps 20 --warmup 5 "$@" > gpurun_out/sweep_n${N}_$tag.json 2> gpurun_out/sweep_n${N}_$tag.err
  python - <<PY
import json
try:
    d=json.loads([l for l in open("gpurun_out/sweep_n${N}_$tag.json") if l.startswith("{")][-1])
    print("%-22s %8.3f ms/step  e2e %9.1f img/s  %s  equal=%s" % ("$tag", d["ms_per_step"], d["e2e"]["value"], d["config"]["collective"], d.get("replicas_equal")))
except Exception as e:
    print("$tag", "FAILED", e)
PY
}
if [ "${SWEEP_FULL:-1}" = "1" ]; then
run default X=1 --
run bucket16 DTF_BUCKET_MB=16 --
run bucket128 DTF_BUCKET_MB=128 --
run blocks16 DTF_COMM_BLOCKS=16 --
run blocks32_t256 DTF_COMM_BLOCKS=32 DTF_COMM_THREADS=256 --
run p2p DTF_NVLS=0 --
run nccl X=1 -- --algo nccl
fi
run pull X=1 -- --algo twoshot_nvls_pull
run oneshot X=1 -- --algo oneshot
run cifar X=1 -- --config cifar10 --steps 50
run dehaze X=1 -- --config dehaze --steps 30
run r18_async X=1 -- --config resnet18_async --steps 30
run r18_sync X=1 -- --config resnet18_async --steps 30 --algo auto_sync
