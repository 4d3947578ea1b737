import os, torch, torch.distributed as dist
import torch.distributed._symmetric_memory as symm
rank = int(os.environ["RANK"]); local = int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
try:
    t = symm.empty(1 << 20, dtype=torch.uint8, device=torch.device("cuda", local))
    h = symm.rendezvous(t, dist.group.WORLD.group_name)
    print(rank, "buffer_ptrs", [hex(p) for p in h.buffer_ptrs][:3], "multicast_ptr", hex(h.multicast_ptr) if h.multicast_ptr else h.multicast_ptr, "signal_pad", len(h.signal_pad_ptrs), flush=True)
except Exception as e:
    print(rank, "ERR", repr(e)[:300], flush=True)
dist.barrier(); dist.destroy_process_group()
