"""Headline benchmark: training images/sec (device-timed, max over ranks) of a BASELINE.json config on N B200s.

    python bench.py --gpus 1 --steps 30 --warmup 5                      # single GPU
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29501 \
        bench.py --gpus 8 --steps 30 --warmup 5                          # one rank per GPU
    python bench.py --impl reference ...                                 # unmodified reference (unavailable: TF1)
    python bench.py --impl torch_nccl ...                                # B-NCCL-AR stand-in baseline

Prints ONE JSON line on rank 0. Timing: W warm-up steps, then K steps each bracketed by CUDA events on the launching
stream; an untimed 256 MiB write flushes L2 between timed steps; barrier + synchronize on both sides; the reported
time is the max over ranks of the summed step times. `e2e` re-measures the same K steps through the public
Trainer.step() API including the per-step H2D copy of uint8 images + labels from pinned memory and the D2H read of
the loss. Weak scaling: the per-GPU batch is fixed.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

CONFIGS = {
    # name: (model, per-GPU batch, image shape, classes (0 = image target), loss, optimizer)
    "cifar10": dict(model="Cifar10CNN", batch=128, shape=(32, 32, 3), classes=10, loss="xent",
                    desc="CIFAR-10-tutorial CNN multi-tower sync gradient averaging"),
    "resnet50": dict(model="ResNet50", batch=256, shape=(224, 224, 3), classes=1000, loss="xent",
                     desc="ResNet-50 synthetic 224x224 sync data-parallel (ps/worker-equivalent) bf16"),
    "resnet18_async": dict(model="ResNet18", batch=64, shape=(224, 224, 3), classes=1000, loss="xent",
                           desc="ResNet-18 async parameter-server mode with sharded owner-side updates"),
    "dehaze": dict(model="DehazeFCN", batch=8, shape=(256, 256, 3), classes=0, loss="mse",
                   desc="fully-convolutional image-to-image net (dehazing-style, 256x256, MSE loss) multi-tower"),
}


class ClockSampler(object):
    """nvidia-smi clocks / throttle reasons sampled during the timed region (B200_PROFILING.md recipe)."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.gpu_index = gpu_index
        self.lines = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                                          "-lms", "20", "-i", str(self.gpu_index)], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        t0 = time.time()
        while not self.lines and time.time() - t0 < 1.0:  # very short timed regions: wait for the first sample
            time.sleep(0.02)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            pass
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1]))
                mx.append(float(f[2]))
            except ValueError:
                continue
            for nm, val in zip(names, f[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(nm)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def dist_setup(n_gpus):
    import torch
    import torch.distributed as dist
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    group = None
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        group = dist.group.WORLD
    return rank, world, local, group


def max_over_ranks(x, world, device):
    import torch
    import torch.distributed as dist
    if world == 1:
        return x
    t = torch.tensor([x], device=device, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def make_batch(cfg, batch, seed):
    import torch
    g = torch.Generator().manual_seed(seed)
    x = torch.randint(0, 256, (batch,) + tuple(cfg["shape"]), generator=g, dtype=torch.uint8).pin_memory()
    if cfg["classes"]:
        y = torch.randint(0, cfg["classes"], (batch,), generator=g).pin_memory()
    else:
        y = torch.rand((batch,) + tuple(cfg["shape"]), generator=g).pin_memory()
    return x, y


def run_ours(args, cfg):
    import torch
    import distributed_tensorflow_framework_b200 as dtf
    from distributed_tensorflow_framework_b200 import models
    from distributed_tensorflow_framework_b200.ops import _lib
    rank, world, local, group = dist_setup(args.gpus)
    device = torch.device("cuda", local)
    _lib.lib()
    batch = args.batch or cfg["batch"]
    loss = dtf.SoftmaxCrossEntropyLoss() if cfg["loss"] == "xent" else dtf.MSELoss()
    algo = args.algo
    if args.config == "resnet18_async" and world > 1 and algo == "auto":
        algo = "async"
    if algo == "auto_sync":  # the async config's model with the synchronous collective (for comparison)
        algo = "auto"
    # asynchronous parameter server: every rank's push is its own update, so the rate is scaled by 1/N as usual
    opt = dtf.MomentumOptimizer(0.01 / world if algo == "async" else 0.01, 0.9)
    trainer = dtf.Trainer(dtf.Net(getattr(models, cfg["model"])()), loss, opt, device=device, rank=rank, world=world,
                          group=group, algo=algo, use_cuda_graph=not args.no_graph, seed=0)
    pool = [make_batch(cfg, batch, 100 + rank * 10 + i) for i in range(2)]
    dev_pool = [(x.to(device), y.to(device)) for x, y in pool]
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=device)

    def barrier():
        torch.cuda.synchronize(device)
        if world > 1:
            import torch.distributed as dist
            dist.barrier()
        torch.cuda.synchronize(device)

    # ---- warm-up (also builds variables, arenas, the communicator and captures the CUDA graph)
    for i in range(max(args.warmup, 3)):
        trainer.step_device([dev_pool[i % 2]])
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    # ---- device-timed region
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    launches0 = _lib.launch_count
    graph_launches = getattr(trainer, "_graph_launches", 0)
    barrier()
    for i in range(args.steps):
        flush.fill_(i & 0xff)  # untimed L2 flush: 256 MiB > 126 MB L2
        ev[i][0].record()
        loss_t = trainer.step_device([dev_pool[i % 2]])
        ev[i][1].record()
    barrier()
    dev_ms_local = sum(a.elapsed_time(b) for a, b in ev)
    dev_ms = max_over_ranks(dev_ms_local, world, device)
    per_rank = [dev_ms_local / args.steps]
    if world > 1:
        import torch.distributed as dist
        t = torch.tensor([dev_ms_local / args.steps], device=device, dtype=torch.float64)
        allt = [torch.zeros_like(t) for _ in range(world)]
        dist.all_gather(allt, t)
        per_rank = [float(v.item()) for v in allt]
    replicas_equal = trainer.engine.replicas_equal() if (world > 1 and trainer.algo != "async") else None
    launches = (_lib.launch_count - launches0) + graph_launches * args.steps
    final_loss = float(loss_t)
    # ---- end-to-end region through the public API: H2D of inputs from pinned memory + D2H of the loss every step
    barrier()
    # the loop a user writes: start the H2D copy of the next pinned batch (copy stream) while this step runs, then read
    # this step's loss; every step's inputs cross PCIe and every step's loss comes back inside the timed region
    trainer.prefetch(pool[0])
    torch.cuda.synchronize(device)
    t0 = time.perf_counter()
    for i in range(args.steps):
        loss_i = trainer.step(pool[i % 2])
        trainer.prefetch(pool[(i + 1) % 2])
        l = float(loss_i)
    torch.cuda.synchronize(device)
    e2e_s = time.perf_counter() - t0
    e2e_s = max_over_ranks(e2e_s, world, device)
    clocks = sampler.stop() if rank == 0 else None
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
    allreduce = measure_allreduce(trainer, world, device, barrier) if world > 1 else None
    gb = batch * world
    x0, y0 = pool[0]
    # BASELINE.md has no published reference number (the TF1 reference cannot run); the number to beat recorded there
    # is the strongest stock-PyTorch stand-in measured on this hardware (baseline/standin_numbers.json), scaled
    # ideally (x N) for N > 1
    vs_baseline, baseline_note = None, None
    try:
        st = json.load(open(os.path.join(ROOT, "baseline", "standin_numbers.json")))[args.config]["1"]
        vs_baseline = (gb * args.steps / (dev_ms / 1e3)) / (st["images_per_s"] * world)
        baseline_note = ("stock-PyTorch stand-in (%s mode) %.1f images/s on 1 GPU, scaled x%d; not the reference"
                         % (st["mode"], st["images_per_s"], world))
    except Exception:
        pass
    res = {
        "metric": "images/sec (device-timed, max over ranks)", "value": gb * args.steps / (dev_ms / 1e3),
        "unit": "images/s", "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
        "ms_per_step": dev_ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": vs_baseline,
        "baseline_note": baseline_note,
        "dtype": "bf16", "data": "synthetic (uint8 images of the named shape, random labels; random-init weights)",
        "impl": "ours",
        "config": {"model": cfg["model"], "desc": cfg["desc"], "global_batch": gb, "per_gpu_batch": batch,
                   "image": list(cfg["shape"]), "parallelism": "dp%d" % world,
                   "collective": trainer.algo + ("_nvls" if getattr(trainer.engine, "nvls", False) else "") +
                   ("_pull" if getattr(trainer.engine, "nopush", None) else ""),
                   "optimizer": "momentum", "cuda_graph": bool(trainer._graph is not None),
                   "l2": "256 MiB write between timed steps (outside the event pairs)"},
        "e2e": {"value": gb * args.steps / e2e_s, "unit": "images/s",
                "h2d_bytes_per_step": x0.numel() * x0.element_size() + y0.numel() * y0.element_size(),
                "d2h_bytes_per_step": 4},
        "gpu_launches": int(launches), "final_loss": final_loss, "clocks": clocks,
        "per_rank_ms_per_step": {"min": min(per_rank), "median": sorted(per_rank)[len(per_rank) // 2],
                                 "max": max(per_rank), "all": [round(v, 4) for v in per_rank]},
        "replicas_equal": replicas_equal,
        "grad_allreduce": allreduce,
    }
    if rank == 0:
        print(json.dumps(res), flush=True)
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
        if trainer.algo == "nccl":
            # NCCL collectives captured in the step's CUDA graph leave ProcessGroupNCCL's watchdog waiting at
            # teardown (observed: destroy_process_group never returns); the result is out, leave without it
            sys.stdout.flush()
            os._exit(0)
        dist.destroy_process_group()


def measure_allreduce(trainer, world, device, barrier, iters=10):
    """Second half of the BASELINE.json metric: bus bandwidth of the gradient collective on this model's gradient
    arena. K-AR (one-shot peer-load allreduce, fp32) and NCCL allreduce on the same buffer, CUDA events, max over
    ranks; bus GB/s = 2 (N-1)/N * bytes / t (nccl-tests convention); fraction of the measured 770 GB/s per-direction
    peer-copy bandwidth (B200_PROFILING.md)."""
    import torch
    import torch.distributed as dist
    eng = trainer.engine
    if eng is None or eng.comm is None:
        return None
    st = trainer.store
    n = st.n_params // 4 * 4
    nbytes = n * 4
    out = torch.empty(n, device=device, dtype=torch.float32)
    goff = eng.heap.offset_of("grad")

    def timed(fn):
        for _ in range(3):
            fn()
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            fn()
        e1.record()
        barrier()
        return max_over_ranks(e0.elapsed_time(e1) / iters, world, device)

    cur = torch.cuda.current_stream(device)
    t_ar = timed(lambda: eng.comm.allreduce_mean(goff, 0, n, out, stream=cur))
    buf = st.grad_arena[:n].clone()
    t_nccl = timed(lambda: dist.all_reduce(buf))
    import distributed_tensorflow_framework_b200 as dtf
    frozen = dtf.GradientDescentOptimizer(0.0)  # lr 0: the owner update leaves the weights unchanged
    t_rs = timed(lambda: eng.comm.rs_optim_ag(goff, st.owner_bounds, frozen, 1, st.master_arena, None, None,
                                              eng.heap.offset_of("shadow"), stream=cur))
    # B-NCCL-PS: the parameter-server data path on library collectives -- ncclReduceScatter of the fp32 gradients,
    # owner-shard SGD update (one elementwise kernel), ncclAllGather of the bf16 weights
    t_ps = None
    try:
        shard = (n // world) // 8 * 8
        if shard > 0:
            g_in = st.grad_arena[:shard * world].clone()
            g_out = torch.empty(shard, device=device, dtype=torch.float32)
            w_master = torch.zeros(shard, device=device, dtype=torch.float32)
            w_bf16 = torch.empty(shard, device=device, dtype=torch.bfloat16)
            w_all = torch.empty(shard * world, device=device, dtype=torch.bfloat16)

            def nccl_ps():
                dist.reduce_scatter_tensor(g_out, g_in)
                w_master.add_(g_out, alpha=-0.0)
                w_bf16.copy_(w_master)
                dist.all_gather_into_tensor(w_all, w_bf16)
            t_ps = timed(nccl_ps)
    except Exception as e:  # an older torch without the tensor collectives: leave the entry empty
        t_ps = None
    f = 2.0 * (world - 1) / world
    return {"bytes": nbytes, "k_rs_opt_ms": t_rs, "nccl_rs_opt_ag_ms": t_ps,
            "k_rs_opt_bus_gbs": (world - 1) / world * (nbytes + nbytes / 2) / t_rs / 1e6, "k_ar_ms": t_ar, "k_ar_bus_gbs": f * nbytes / t_ar / 1e6,
            "nccl_ms": t_nccl, "nccl_bus_gbs": f * nbytes / t_nccl / 1e6,
            "k_ar_frac_of_770": f * nbytes / t_ar / 1e6 / 770.0,
            # link roofline of K-RS-OPT: every GPU must send (N-1)/N of its fp32 gradient arena towards the owners
            # (the bf16 all-gather is S/2N per owner on NVLS); achieved fraction of the measured 770 GB/s per direction
            "k_rs_opt_link_bytes_per_gpu": (world - 1) / world * nbytes,
            "k_rs_opt_frac_of_770": ((world - 1) / world * nbytes / 770e9) / (t_rs / 1e3),
            "note": "one-shot K-AR reads (N-1) remote copies per rank: its wire bytes per GPU are (N-1)*S, so the "
                    "nccl-tests bus figure understates link utilisation for N > 2"}


def run_reference(args, cfg):
    """The unmodified reference needs TensorFlow 1.x (tf.app.flags, tf.contrib, SyncReplicasOptimizer); it has no
    setup.py / pyproject, `pip install --no-index --target baseline/_ref /root/reference` fails ("not installable"),
    and no TF wheel exists for this Python."""
    if int(os.environ.get("RANK", "0")) == 0:
        print(json.dumps({"impl": "reference", "unavailable":
                          "reference is TF1-only pure-Python code without setup.py; TensorFlow 1.x is not installed "
                          "and not installable offline on Python 3.12 (pip: 'Directory /root/reference is not "
                          "installable')"}), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference", "torch_nccl"])
    ap.add_argument("--config", default=os.environ.get("DTF_BENCH_CONFIG", "resnet50"), choices=sorted(CONFIGS))
    ap.add_argument("--batch", type=int, default=0, help="per-GPU batch (default: the config's)")
    ap.add_argument("--algo", default="auto")
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--baseline-mode", default=None, choices=[None, "eager", "graph", "compile"],
                    help="--impl torch_nccl only: see baseline/torch_baseline.py")
    args = ap.parse_args()
    cfg = CONFIGS[args.config]
    if args.impl == "reference":
        return run_reference(args, cfg)
    if args.impl == "torch_nccl":
        sys.path.insert(0, os.path.join(ROOT, "baseline"))
        import torch_baseline
        return torch_baseline.run(args, cfg, dist_setup, max_over_ranks, make_batch, ClockSampler)
    return run_ours(args, cfg)


if __name__ == "__main__":
    main()
