"""B-NCCL-AR stand-in baseline: the same model / config in stock PyTorch -- cuDNN convolutions and
cuBLAS GEMMs, channels_last, NCCL gradient all-reduce, torch.optim.SGD(momentum). This is NOT the reference (which is
TF1-only and cannot run here); it is the "library path" the hand-written kernels are compared against. None of this
repo's kernels are used on this path.

Modes (DTF_BASELINE_MODE or --baseline-mode):
  eager    round-1 stand-in: bf16 autocast over fp32 parameters, eager launches, flat ncclAllReduce after backward.
  graph    (default) the strongest plain-library configuration: parameters and activations in bf16 channels_last (no
           autocast casts, no fp32 master copy -- less work than this repo's fp32-master path does), the whole step
           (forward, backward, fused SGD) captured in ONE CUDA graph at N = 1; at N > 1 DistributedDataParallel
           (bucketed NCCL all-reduce overlapped with backward), eager.
  compile  as graph, with the model compiled by torch.compile (inductor fuses batch norm / ReLU / residual adds into
           Triton kernels; mode="max-autotune-no-cudagraphs" + our own graph capture at N = 1)."""
import os
import json
import time

import torch
import torch.nn as nn
import torch.nn.functional as F


class CifarCNN(nn.Module):
    def __init__(self):
        super().__init__()
        self.c1 = nn.Conv2d(3, 64, 5, padding=2)
        self.c2 = nn.Conv2d(64, 64, 5, padding=2)
        self.f1 = nn.Linear(8 * 8 * 64, 384)
        self.f2 = nn.Linear(384, 192)
        self.f3 = nn.Linear(192, 10)

    def forward(self, x):
        x = F.max_pool2d(F.relu(self.c1(x)), 3, 2, 1)
        x = F.local_response_norm(x, 9, alpha=0.001, beta=0.75, k=1.0)
        x = F.local_response_norm(F.relu(self.c2(x)), 9, alpha=0.001, beta=0.75, k=1.0)
        x = F.max_pool2d(x, 3, 2, 1)
        x = x.permute(0, 2, 3, 1).flatten(1)
        return self.f3(F.relu(self.f2(F.relu(self.f1(x)))))


def _cbr(cin, cout, k, s):
    return nn.Sequential(nn.Conv2d(cin, cout, k, s, k // 2, bias=False), nn.BatchNorm2d(cout, eps=1e-3))


class Bottleneck(nn.Module):
    def __init__(self, cin, mid, stride):
        super().__init__()
        cout = mid * 4
        self.a, self.b, self.c = _cbr(cin, mid, 1, 1), _cbr(mid, mid, 3, stride), _cbr(mid, cout, 1, 1)
        self.sc = _cbr(cin, cout, 1, stride) if (stride != 1 or cin != cout) else None

    def forward(self, x):
        s = x if self.sc is None else self.sc(x)
        return F.relu(self.c(F.relu(self.b(F.relu(self.a(x))))) + s)


class Basic(nn.Module):
    def __init__(self, cin, cout, stride):
        super().__init__()
        self.a, self.b = _cbr(cin, cout, 3, stride), _cbr(cout, cout, 3, 1)
        self.sc = _cbr(cin, cout, 1, stride) if (stride != 1 or cin != cout) else None

    def forward(self, x):
        s = x if self.sc is None else self.sc(x)
        return F.relu(self.b(F.relu(self.a(x))) + s)


class ResNet(nn.Module):
    def __init__(self, blocks, bottleneck, classes=1000):
        super().__init__()
        self.stem = _cbr(3, 64, 7, 2)
        layers, cin = [], 64
        for st, n in enumerate(blocks):
            w = 64 * 2 ** st
            for b in range(n):
                stride = 2 if (b == 0 and st > 0) else 1
                if bottleneck:
                    layers.append(Bottleneck(cin, w, stride)); cin = w * 4
                else:
                    layers.append(Basic(cin, w, stride)); cin = w
        self.layers = nn.Sequential(*layers)
        self.fc = nn.Linear(cin, classes)

    def forward(self, x):
        x = F.max_pool2d(F.relu(self.stem(x)), 3, 2, 1)
        return self.fc(self.layers(x).mean((2, 3)))


class Dehaze(nn.Module):
    def __init__(self, c=64, blocks=4):
        super().__init__()
        cv = lambda i, o, s=1: nn.Conv2d(i, o, 3, s, 1)
        self.c1, self.c2, self.d1, self.d2 = cv(3, c), cv(c, c), cv(c, 2 * c, 2), cv(2 * c, 4 * c, 2)
        self.res = nn.ModuleList([nn.ModuleList([cv(4 * c, 4 * c), cv(4 * c, 4 * c)]) for _ in range(blocks)])
        self.u1 = nn.ConvTranspose2d(4 * c, 2 * c, 3, 2, 1, output_padding=1, bias=False)
        self.u2 = nn.ConvTranspose2d(2 * c, c, 3, 2, 1, output_padding=1, bias=False)
        self.c3, self.out = cv(c, c), cv(c, 3)

    def forward(self, x):
        e1 = F.relu(self.c2(F.relu(self.c1(x))))
        x = F.relu(self.d2(F.relu(self.d1(e1))))
        for a, b in self.res:
            x = F.relu(x + b(F.relu(a(x))))
        x = F.relu(self.u2(F.relu(self.u1(x)))) + e1
        return self.out(F.relu(self.c3(x)))


def build(name):
    return {"Cifar10CNN": CifarCNN, "ResNet50": lambda: ResNet((3, 4, 6, 3), True),
            "ResNet18": lambda: ResNet((2, 2, 2, 2), False), "DehazeFCN": Dehaze}[name]()


def run(args, cfg, dist_setup, max_over_ranks, make_batch, ClockSampler):
    rank, world, local, group = dist_setup(args.gpus)
    device = torch.device("cuda", local)
    torch.backends.cudnn.benchmark = True
    mode = getattr(args, "baseline_mode", None) or os.environ.get("DTF_BASELINE_MODE", "graph")
    pure_bf16 = mode in ("graph", "compile")
    model = build(cfg["model"]).to(device).to(memory_format=torch.channels_last)
    if pure_bf16:
        model = model.to(torch.bfloat16)
    params = [p for p in model.parameters()]
    try:
        opt = torch.optim.SGD(params, lr=0.01, momentum=0.9, fused=pure_bf16)
    except (TypeError, RuntimeError):
        opt = torch.optim.SGD(params, lr=0.01, momentum=0.9)
    batch = args.batch or cfg["batch"]
    pool = [make_batch(cfg, batch, 100 + rank * 10 + i) for i in range(2)]
    fwd = model
    if world > 1 and pure_bf16:
        from torch.nn.parallel import DistributedDataParallel as DDP
        fwd = DDP(model, device_ids=[local], gradient_as_bucket_view=True, bucket_cap_mb=25)
    if mode == "compile":
        fwd = torch.compile(fwd, mode="max-autotune-no-cudagraphs")
    act_dtype = torch.bfloat16 if pure_bf16 else torch.float32

    def compute(x_u8, y):
        x = x_u8.permute(0, 3, 1, 2).to(act_dtype).div_(255.).contiguous(memory_format=torch.channels_last)
        if pure_bf16:
            out = fwd(x)
        else:
            with torch.autocast("cuda", dtype=torch.bfloat16):
                out = fwd(x)
        if cfg["classes"]:
            loss = F.cross_entropy(out.float(), y)
        else:
            loss = F.mse_loss(out.float(), y.permute(0, 3, 1, 2))
        opt.zero_grad(set_to_none=False)
        loss.backward()
        if world > 1 and not pure_bf16:
            import torch.distributed as dist
            flat = torch.cat([p.grad.reshape(-1) for p in params])
            dist.all_reduce(flat)
            flat.div_(world)
            off = 0
            for p in params:
                n = p.numel()
                p.grad.copy_(flat[off:off + n].view_as(p.grad))
                off += n
        opt.step()
        return loss

    graph = {"g": None}
    static = {}

    def step(xh, yh):
        if graph["g"] is not None:  # replay: copy the inputs into the captured step's static buffers
            static["x"].copy_(xh, non_blocking=True)
            static["y"].copy_(yh, non_blocking=True)
            graph["g"].replay()
            return static["loss"]
        return compute(xh.to(device, non_blocking=True), yh.to(device, non_blocking=True))

    def capture():
        x0, y0 = pool[0]
        static["x"], static["y"] = x0.to(device), y0.to(device)
        s = torch.cuda.Stream(device=device)
        s.wait_stream(torch.cuda.current_stream(device))
        with torch.cuda.stream(s):
            for _ in range(3):
                compute(static["x"], static["y"])
        torch.cuda.current_stream(device).wait_stream(s)
        torch.cuda.synchronize(device)
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            static["loss"] = compute(static["x"], static["y"])
        graph["g"] = g

    def barrier():
        torch.cuda.synchronize(device)
        if world > 1:
            import torch.distributed as dist
            dist.barrier()

    dev_pool = [(x.to(device), y.to(device)) for x, y in pool]
    for i in range(max(args.warmup, 3)):
        step(*dev_pool[i % 2])
    graphed = False
    if pure_bf16 and world == 1:
        try:
            capture()
            graphed = True
            for i in range(2):
                step(*dev_pool[i % 2])
        except Exception as e:  # capture is an optimisation of the stand-in, never a reason to lose the number
            graph["g"] = None
            print("baseline: CUDA graph capture failed (%s); running eagerly" % str(e)[:200], flush=True)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=device)
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    barrier()
    for i in range(args.steps):
        flush.fill_(i & 0xff)
        ev[i][0].record()
        loss = step(*dev_pool[i % 2])
        ev[i][1].record()
    barrier()
    dev_ms = max_over_ranks(sum(a.elapsed_time(b) for a, b in ev), world, device)
    barrier()
    t0 = time.perf_counter()
    for i in range(args.steps):
        l = float(step(*pool[i % 2]))
    torch.cuda.synchronize(device)
    e2e_s = max_over_ranks(time.perf_counter() - t0, world, device)
    clocks = sampler.stop() if rank == 0 else None
    gb = batch * world
    x0, y0 = pool[0]
    if rank == 0:
        print(json.dumps({
            "metric": "images/sec (device-timed, max over ranks)", "value": gb * args.steps / (dev_ms / 1e3),
            "unit": "images/s", "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": dev_ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "bf16" if pure_bf16 else "bf16 autocast", "data": "synthetic",
            "impl": "torch_nccl (cuDNN/cuBLAS + NCCL stand-in, mode %s)" % mode,
            "config": {"model": cfg["model"], "global_batch": gb, "parallelism": "dp%d" % world, "mode": mode,
                       "cuda_graph": graphed, "ddp_overlap": bool(world > 1 and pure_bf16)},
            "e2e": {"value": gb * args.steps / e2e_s, "unit": "images/s",
                    "h2d_bytes_per_step": x0.numel() * x0.element_size() + y0.numel() * y0.element_size(),
                    "d2h_bytes_per_step": 4},
            "gpu_launches": 0, "final_loss": float(loss), "clocks": clocks}), flush=True)
    if world > 1:
        import torch.distributed as dist
        dist.destroy_process_group()
