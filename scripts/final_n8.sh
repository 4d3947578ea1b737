#!/bin/bash
# Round-end records on N GPUs of one box: multi-rank collective tests at world N, then bench.py for every GPU config.
N=${1:-8}
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_comm.py -m gpu -q -k "collectives and $N" 2>&1 | tail -3
run() { tag=$1; shift
  timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29600 + RANDOM % 300)) bench.py --gpus $N --warmup 5 "$@" > gpurun_out/final_n${N}_$tag.json 2> gpurun_out/final_n${N}_$tag.err
  python - <<PY
import json
try:
    d=json.loads([l for l in open("gpurun_out/final_n${N}_$tag.json") if l.startswith("{")][-1])
    ga=d.get("grad_allreduce") or {}
    print("%-14s %9.3f ms/step %11.1f img/s e2e %11.1f  %s equal=%s vs_baseline=%s rs_opt=%.3f ms (%.0f%% of 770 GB/s link roofline) nccl_ar=%.3f ms" % (
        "$tag", d["ms_per_step"], d["value"], d["e2e"]["value"], d["config"]["collective"], d.get("replicas_equal"),
        ("%.3f" % d["vs_baseline"]) if d.get("vs_baseline") else None, ga.get("k_rs_opt_ms", float("nan")),
        100 * ga.get("k_rs_opt_frac_of_770", float("nan")), ga.get("nccl_ms", float("nan"))))
except Exception as e:
    print("$tag", "FAILED", e)
PY
}
run resnet50 --steps 20
run cifar10 --config cifar10 --steps 50
run dehaze --config dehaze --steps 30
run resnet18_async --config resnet18_async --steps 30
run resnet18_sync --config resnet18_async --steps 30 --algo auto_sync
run resnet50_pull --steps 20 --algo twoshot_nvls_pull
