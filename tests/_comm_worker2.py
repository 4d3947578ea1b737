"""Round-2 multi-rank checks (torchrun, one rank per GPU): fp32 masters of biases / BN parameters on NON-owner ranks
in the owner-sharded modes, backup workers with an injected straggler, Adam + learning-rate schedule under CUDA-graph
replay across ranks, async mode under graph replay (versions keep advancing, staleness bound), a barrier that times
out (error record instead of a trap) and re-synchronises."""
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


def gather_equal(t):
    g = [torch.empty_like(t) for _ in range(dist.get_world_size())]
    dist.all_gather(g, t.contiguous())
    return all(torch.equal(g[0], x) for x in g)


def make_trainer(rank, world, dev, algo, model=None, opt=None, graph=False, **kw):
    return dtf.Trainer(dtf.Net(model or models.Cifar10CNN()), dtf.SoftmaxCrossEntropyLoss(),
                       opt or dtf.MomentumOptimizer(0.01, 0.9), device=dev, rank=rank, world=world,
                       group=dist.group.WORLD, algo=algo, seed=5, use_cuda_graph=graph, **kw)


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    torch.manual_seed(100 + rank)
    x = torch.rand(32, 32, 32, 3)
    y = torch.randint(0, 10, (32,))

    # ---- 1. fp32-read parameters (biases, BN gamma/beta) on non-owner ranks: every owner-sharded mode must leave
    # them identical on all ranks and equal to the replicated-master one-shot run
    ref = {}
    for algo in ("oneshot", "twoshot", "twoshot_nvls"):
        tr = make_trainer(rank, world, dev, algo)
        losses = [float(tr.step((x, y))) for _ in range(6)]
        torch.cuda.synchronize()
        dist.barrier()
        st = tr.store
        lo, hi = st.f32_ranges[rank]
        check(hi > lo, "%s: empty fp32 region" % algo)
        f32 = st.master_arena[lo:hi].clone()
        check(gather_equal(f32), "%s: fp32 region (biases) differs between ranks" % algo)
        check(gather_equal(st.shadow_arena), "%s: bf16 replicas diverged" % algo)
        bias = st.vars["conv2/biases"].master
        check(float((bias - 0.1).abs().max()) > 0 or float(bias.abs().max()) > 0, "%s: bias never updated" % algo)
        if algo == "oneshot":
            ref["f32"], ref["losses"] = f32, losses
        else:
            # same layout (bytes placement): compare with the replicated-master run
            err = float((f32 - ref["f32"]).abs().max())
            check(err < 5e-4, "%s: biases deviate from the one-shot run by %g" % (algo, err))
            dl = max(abs(a - b) for a, b in zip(losses, ref["losses"]))
            check(dl < 2e-2, "%s: loss curve deviates from one-shot: %s vs %s" % (algo, losses, ref["losses"]))
        if rank == 0:
            print("f32-region %-13s ok  %.4f -> %.4f" % (algo, losses[0], losses[-1]), flush=True)
        del tr
    # ResNet-18 (batch norm): gamma / beta on non-owners under NVLS two-shot
    xr = torch.randint(0, 256, (8, 64, 64, 3), dtype=torch.uint8)
    yr = torch.randint(0, 1000, (8,))
    tr = make_trainer(rank, world, dev, "twoshot_nvls", model=models.ResNet18())
    l = [float(tr.step((xr, yr))) for _ in range(4)]
    torch.cuda.synchronize()
    dist.barrier()
    g = tr.store.vars["conv1/bn/gamma"].master
    check(gather_equal(g) and float((g - 1.0).abs().max()) > 0, "BN gamma stale on a non-owner rank")
    check(gather_equal(tr.store.shadow_arena), "resnet18 nvls: replicas diverged")
    if rank == 0:
        print("resnet18 gamma/beta replicated ok", l[0], l[-1], flush=True)
    del tr

    # ---- 2. backup workers at the collective level: rank (world-1) is late, R = world-1 -> every owner aggregates
    # the other ranks only, scale 1/R
    n = 1 << 18
    heap = SymmetricHeap(n * 4 * 6 + (4 << 20), dev)
    grad = heap.alloc("grad", n, torch.float32)
    shadow = heap.alloc("shadow", n, torch.bfloat16)
    master = heap.alloc("master", n, torch.float32)
    comm = Communicator(heap, rank, world)
    base = torch.randn(world, n, device=dev, generator=torch.Generator(device=dev).manual_seed(3))
    per = (n // world + 7) // 8 * 8
    bounds = [min(i * per, n) for i in range(world)] + [n]
    sgd = dtf.GradientDescentOptimizer(1.0)
    R = world - 1
    for rnd in range(2):
        grad.copy_(base[rank] * (rnd + 1))
        master.zero_()
        torch.cuda.synchronize()
        dist.barrier()
        if rank == world - 1:
            comm.sleep(0.05)  # the straggler: its collective starts 50 ms after everybody else's
        comm.rs_optim_ag(heap.offset_of("grad"), bounds, sgd, 1, master, None, None, heap.offset_of("shadow"),
                         master_off=heap.offset_of("master"), backup_r=R)
        torch.cuda.synchronize()
        dist.barrier()
        want = -(rnd + 1) * sum(base[r] for r in range(world - 1)) / R
        check(comm.error() is None, "backup: error record %s" % (comm.error(),))
        check(torch.allclose(master, want, atol=1e-5, rtol=1e-5), "backup round %d: late rank was not dropped" % rnd)
    if rank == 0:
        print("backup-worker collective ok (R=%d of %d)" % (R, world), flush=True)

    # ---- 3. barrier time-out: the last rank skips a barrier; the others get an error record, nobody traps, and the
    # group re-synchronises afterwards
    comm.set_timeout_ms(300)
    dist.barrier()
    if rank != world - 1:
        comm.barrier()
        torch.cuda.synchronize()
        e = comm.error()
        check(e is not None and e["code"] == 1 and e["peer"] == world - 1, "expected a barrier time-out, got %s" % (e,))
        try:
            comm.check()
            check(False, "check() did not raise")
        except dtf.WorkerLostError:
            pass
        comm.clear_error()
        torch.cuda.synchronize()
    dist.barrier()
    if rank == world - 1:
        comm.barrier()  # the late barrier pairs with the epochs the others already published
        torch.cuda.synchronize()
    dist.barrier()
    comm.set_timeout_ms(30000)
    comm.barrier()
    torch.cuda.synchronize()
    check(comm.error() is None, "communicator did not recover after the time-out")
    if rank == 0:
        print("barrier time-out -> error record -> recovery ok", flush=True)
    dist.barrier()

    # ---- 3b. protocol check: ranks that enter DIFFERENT collectives at the same barrier epoch are told so (this is
    # what a chief-only checkpoint gather used to do silently: pair with the peers' next gradient collective)
    dist.barrier()
    if rank == 0:
        comm.allgather_slices(heap.offset_of("shadow"), [0] + [8] * world)  # 128 bytes from rank 0: one block
    else:
        comm.barrier()
    torch.cuda.synchronize()
    e = comm.error()
    check(e is not None and e["code"] == 3, "expected a collective-mismatch record, got %s" % (e,))
    dist.barrier()
    comm.clear_error()
    torch.cuda.synchronize()
    dist.barrier()
    comm.barrier()
    torch.cuda.synchronize()
    check(comm.error() is None, "communicator unusable after the mismatch was cleared: %s" % (comm.error(),))
    if rank == 0:
        print("collective-kind mismatch detected and cleared ok", flush=True)
    dist.barrier()

    # ---- 4. backup mode through the Trainer (captured graph) with the straggler injected on the last rank
    # (every rank trains on the SAME batch here: whichever subset an owner aggregates, the update is the same, so the
    # loss must fall on every rank; that late gradients are really dropped is what section 2 verified exactly)
    gsame = torch.Generator().manual_seed(77)
    xs, ys = torch.rand(32, 32, 32, 3, generator=gsame), torch.randint(0, 10, (32,), generator=gsame)
    tr = make_trainer(rank, world, dev, "auto", graph=True, replicas_to_aggregate=R)
    check(tr.replicas_to_aggregate == R, "trainer did not keep replicas_to_aggregate")
    ls = []
    for i in range(8):
        if rank == world - 1 and i % 2 == 1:
            tr.engine.comm.sleep(0.02) if tr.engine is not None else None
        ls.append(float(tr.step((xs, ys))))
    check(tr.engine.comm.error() is None, "backup trainer: error record %s" % (tr.engine.comm.error(),))
    torch.cuda.synchronize()
    dist.barrier()
    check(tr.engine.backup_r == R and tr.engine.algo == "twoshot" and not tr.engine.nvls, "backup engine setup")
    check(all(v == v for v in ls) and ls[-1] < ls[0], "backup trainer: loss %s" % ls)
    check(gather_equal(tr.store.shadow_arena), "backup trainer: replicas diverged")
    if rank == 0:
        print("backup-worker trainer ok %.4f -> %.4f" % (ls[0], ls[-1]), flush=True)
    del tr

    # ---- 5. Adam + a learning-rate schedule under CUDA-graph replay == the same run without a graph
    class Sched(object):
        def __init__(self):
            self.learning_rate = 1e-3
    curves = {}
    for graph in (False, True):
        sched = Sched()
        tr = make_trainer(rank, world, dev, "oneshot", opt=dtf.AdamOptimizer(sched), graph=graph)
        ls = []
        for i in range(8):
            if i == 4:
                sched.learning_rate = 1e-4
            ls.append(float(tr.step((x, y))))
        torch.cuda.synchronize()
        dist.barrier()
        check(tr.store.global_step == 8, "global_step %d after 8 steps (graph=%s)" % (tr.store.global_step, graph))
        check((tr._graph is not None) == graph, "graph capture state")
        curves[graph] = (ls, tr.store.master_arena.clone())
        check(gather_equal(tr.store.master_arena), "adam graph=%s: masters differ between ranks" % graph)
        del tr
    dl = max(abs(a - b) for a, b in zip(curves[False][0], curves[True][0]))
    dm = float((curves[False][1] - curves[True][1]).abs().mean())  # red.add filter gradients are not bit-stable
    check(dl < 2e-2 and dm < 2e-4, "Adam under graph replay deviates: dloss %g dmaster %g\n%s\n%s"
          % (dl, dm, curves[False][0], curves[True][0]))
    if rank == 0:
        print("adam + lr schedule under graph replay ok (dloss %.2e, dmaster %.2e)" % (dl, dm), flush=True)

    # ---- 6. async mode under graph replay: the owner keeps applying (versions advance in-kernel)
    tr = make_trainer(rank, world, dev, "async", graph=True)
    ls = [float(tr.step((x, y))) for _ in range(10)]
    torch.cuda.synchronize()
    dist.barrier()
    stats = tr.engine.comm.async_stats()
    check(tr._graph is not None, "async: no graph")
    check(stats["pushed"] >= 10 and stats["applied"] >= 10, "async under graph: owner stopped applying %s" % stats)
    check(all(v == v for v in ls) and min(ls[5:]) < ls[0], "async graph: loss %s" % ls)
    lo, hi = tr.store.f32_ranges[rank]
    if rank == 0:
        print("async under graph replay ok: %s, %.4f -> %.4f" % (stats, ls[0], ls[-1]), flush=True)
    del tr
    # staleness bound at the collective level (deterministic): a push stamped with an old owner step is dropped
    cap = (max(bounds[i + 1] - bounds[i] for i in range(world)) + 3) // 4 * 4
    heap.alloc("mailbox", cap * world, torch.float32)
    goff, moff, soff = heap.offset_of("grad"), heap.offset_of("mailbox"), heap.offset_of("shadow")
    grad.copy_(base[rank])
    master.zero_()

    def settle():
        torch.cuda.synchronize()
        dist.barrier()

    settle()
    comm.async_begin()
    comm.async_push(goff, moff, cap, bounds)
    settle()
    snap = comm.async_apply(moff, cap, bounds, sgd, 1, master, None, None, soff, max_staleness=0).tolist()
    settle()
    check(all(v != 0 for v in snap[:world]), "fresh pushes must be applied: %s" % snap)
    s1 = comm.async_stats()
    check(s1["applied"] == world and s1["dropped"] == 0 and s1["pushed"] == 1, "async stats %s" % s1)
    comm.async_push(goff, moff, cap, bounds)  # no async_begin: still stamped with owner step 0 -> stale by `world`
    settle()
    snap = comm.async_apply(moff, cap, bounds, sgd, 2, master, None, None, soff, max_staleness=0).tolist()
    settle()
    s2 = comm.async_stats()
    check(all(v == 0 for v in snap[:world]) and s2["dropped"] == world and s2["applied"] == world,
          "stale pushes must be dropped: snap %s stats %s" % (snap, s2))
    comm.async_begin()
    comm.async_push(goff, moff, cap, bounds)
    settle()
    snap = comm.async_apply(moff, cap, bounds, sgd, 3, master, None, None, soff, max_staleness=0).tolist()
    settle()
    s3 = comm.async_stats()
    check(all(v != 0 for v in snap[:world]) and s3["applied"] == 2 * world and s3["pushed"] == 3,
          "re-stamped pushes must be applied: snap %s stats %s" % (snap, s3))
    lo, hi = bounds[rank], bounds[rank + 1]
    want_own = -2.0 * sum(base[r][lo:hi] for r in range(world))
    check(torch.allclose(master[lo:hi], want_own, atol=1e-4, rtol=1e-4), "async owner shard after 2 applied rounds")
    if rank == 0:
        print("async staleness bound ok: %s" % s3, flush=True)
    dist.barrier()
    if rank == 0:
        print("COMM_WORKER2_OK", flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
