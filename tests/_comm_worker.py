"""Multi-rank worker for tests/test_gpu_comm.py (launched with torchrun, one rank per GPU)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist

import distributed_tensorflow_framework_b200 as dtf
from distributed_tensorflow_framework_b200 import models
from distributed_tensorflow_framework_b200.parallel.comm import Communicator, SymmetricHeap


def check(cond, msg):
    if not cond:
        print("RANK %s FAIL: %s" % (os.environ.get("RANK"), msg), flush=True)
        sys.exit(1)


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    n = 1 << 20
    heap = SymmetricHeap(n * 4 * 8 + (8 << 20), dev)
    grad = heap.alloc("grad", n, torch.float32)
    shadow = heap.alloc("shadow", n, torch.bfloat16)
    master_h = heap.alloc("master", n, torch.float32)
    comm = Communicator(heap, rank, world)
    torch.manual_seed(1234)
    base = torch.randn(world, n, device=dev)  # same on all ranks
    grad.copy_(base[rank])
    want_mean = base.sum(0) / world  # reference reduces in rank order too
    want_mean = sum(base[r] for r in range(world)) / world
    torch.cuda.synchronize()
    dist.barrier()

    # ---- K-AR as plain allreduce-mean
    out = torch.empty(n, device=dev)
    with torch.cuda.stream(comm.stream):
        comm.allreduce_mean(heap.offset_of("grad"), 0, n, out, stream=comm.stream)
    comm.stream.synchronize()
    check(torch.allclose(out, want_mean, atol=1e-6, rtol=1e-6), "allreduce_mean value")
    gathered = [torch.empty_like(out) for _ in range(world)]
    dist.all_gather(gathered, out)
    check(all(torch.equal(gathered[0], g) for g in gathered), "allreduce_mean not bitwise identical across ranks")

    # ---- K-AR fused with momentum
    opt = dtf.MomentumOptimizer(0.1, 0.9)
    p0 = torch.randn(n, device=dev, generator=torch.Generator(device=dev).manual_seed(7))
    master = p0.clone()
    m = torch.zeros(n, device=dev)
    v = torch.zeros(n, device=dev)
    sh = torch.empty(n, device=dev, dtype=torch.bfloat16)
    ref_p, ref_m = p0.clone(), torch.zeros(n, device=dev)
    for step in (1, 2):
        with torch.cuda.stream(comm.stream):
            comm.allreduce_optim(heap.offset_of("grad"), 0, n, opt, step, master, m, v, sh, stream=comm.stream)
        comm.stream.synchronize()
        ref_m = 0.9 * ref_m + want_mean
        ref_p = ref_p - 0.1 * ref_m
    check(torch.allclose(master, ref_p, atol=1e-5, rtol=1e-5), "allreduce_optim master")
    check(torch.equal(sh, master.to(torch.bfloat16)), "allreduce_optim shadow")

    # ---- K-RS-OPT + all-gather by stores (SGD), fp32 masters broadcast too
    sgd = dtf.GradientDescentOptimizer(0.5)
    per = (n // world + 7) // 8 * 8
    bounds = [min(i * per, n) for i in range(world)] + [n]
    master_h.copy_(p0)
    torch.cuda.synchronize()
    dist.barrier()
    with torch.cuda.stream(comm.stream):
        comm.rs_optim_ag(heap.offset_of("grad"), bounds, sgd, 1, master_h, None, None, heap.offset_of("shadow"),
                         master_off=heap.offset_of("master"), stream=comm.stream)
    comm.stream.synchronize()
    dist.barrier()
    want = p0 - 0.5 * want_mean
    check(torch.allclose(master_h, want, atol=1e-5, rtol=1e-5), "rs_optim_ag masters (all shards, via broadcast)")
    check(torch.equal(shadow, want.to(torch.bfloat16)) or
          torch.allclose(shadow.float(), want, atol=2e-2, rtol=2e-2), "rs_optim_ag shadow")

    # ---- barrier + heartbeat
    comm.barrier()
    seen = comm.heartbeat(5)
    torch.cuda.synchronize()
    dist.barrier()
    seen = comm.heartbeat(6).tolist()
    check(all(s >= 5 for s in seen[:world]), "heartbeat %s" % seen)

    # ---- async push / apply: after everybody pushed and (later) applied, each owner has applied `world` gradients
    mailbox_cap = (max(bounds[i + 1] - bounds[i] for i in range(world)) + 3) // 4 * 4
    mailbox = heap.alloc("mailbox", mailbox_cap * world, torch.float32)
    master_h.copy_(p0)
    torch.cuda.synchronize()
    dist.barrier()
    comm.async_push(heap.offset_of("grad"), heap.offset_of("mailbox"), mailbox_cap, bounds)
    torch.cuda.synchronize()
    dist.barrier()  # all pushes have landed
    snap = comm.async_apply(heap.offset_of("mailbox"), mailbox_cap, bounds, sgd, 1, master_h, None, None,
                            heap.offset_of("shadow")).tolist()
    torch.cuda.synchronize()
    dist.barrier()
    check(all(s == 1 for s in snap[:world]), "async snapshot %s" % snap)
    lo, hi = bounds[rank], bounds[rank + 1]
    want_own = p0[lo:hi] - 0.5 * sum(base[r][lo:hi] for r in range(world))
    check(torch.allclose(master_h[lo:hi], want_own, atol=1e-4, rtol=1e-4), "async owner shard")
    check(torch.allclose(shadow.float(), (p0 - 0.5 * base.sum(0)), atol=3e-2, rtol=3e-2), "async shadows everywhere")
    snap2 = comm.async_apply(heap.offset_of("mailbox"), mailbox_cap, bounds, sgd, 2, master_h, None, None,
                             heap.offset_of("shadow")).tolist()
    torch.cuda.synchronize()
    check(all(s == 0 for s in snap2[:world]), "async re-apply must be a no-op %s" % snap2)
    dist.barrier()

    # ---- Trainer level: data-parallel CIFAR CNN, every algorithm
    curves = {}
    for algo in ("oneshot", "twoshot", "twoshot_pull", "twoshot_nvls", "nccl", "async"):
        torch.manual_seed(100 + rank)
        x = torch.rand(32, 32, 32, 3)
        y = torch.randint(0, 10, (32,))
        # asynchronous mode applies every rank's push as its own update: the usual 1/N learning-rate scaling
        lr = 0.01 / world if algo == "async" else 0.01
        tr = dtf.Trainer(dtf.Net(models.Cifar10CNN()), dtf.SoftmaxCrossEntropyLoss(), dtf.MomentumOptimizer(lr, 0.9),
                         device=dev, rank=rank, world=world, group=dist.group.WORLD, algo=algo, seed=5,
                         use_cuda_graph=(algo == "oneshot"))
        losses = [float(tr.step((x, y))) for _ in range(12)]
        check(all(l == l for l in losses), "%s: NaN loss" % algo)
        if algo == "async":
            # every push is its own optimizer update at the full learning rate (N updates per step, each from another
            # rank's batch): the loss on this rank's batch is noisy, it only has to get below its start at some point
            check(min(losses[1:]) < losses[0], "%s: loss never fell below its start %s" % (algo, losses))
        else:
            check(losses[-1] < losses[0], "%s: loss did not decrease %s" % (algo, losses))
        curves[algo] = losses
        if algo == "twoshot_nvls":
            check(tr.engine.heap.multicast_ptr != 0, "no NVLS multicast mapping")
            diff = max(abs(a - b) for a, b in zip(losses, curves["twoshot"]))
            check(diff < 2e-2, "NVLS path deviates from the P2P path: %s vs %s" % (losses, curves["twoshot"]))
        if algo == "twoshot_pull":
            check(len(tr.engine.nopush) == 2, "pull mode should cover local3 and local4 weights")
            diff = max(abs(a - b) for a, b in zip(losses, curves["twoshot"]))
            check(diff < 2e-2, "K-AG-GEMM pull path deviates from the push path: %s vs %s" % (losses, curves["twoshot"]))
        torch.cuda.synchronize()
        dist.barrier()
        if algo not in ("async", "twoshot_pull"):
            sh = tr.store.shadow_arena
            g = [torch.empty_like(sh) for _ in range(world)]
            dist.all_gather(g, sh)
            check(all(torch.equal(g[0], t) for t in g), "%s: replicas diverged" % algo)
        if rank == 0:
            print("trainer %-8s ok: %.4f -> %.4f" % (algo, losses[0], losses[-1]), flush=True)
    dist.barrier()
    if rank == 0:
        print("COMM_WORKER_OK", flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
