import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from distributed_tensorflow_framework_b200.ops import raw, _lib
torch.manual_seed(0)
for (B, H, W, Ci, Co, k) in [(2, 14, 14, 64, 64, 3), (4, 56, 56, 64, 64, 3), (3, 28, 28, 128, 128, 3), (2, 14, 14, 256, 256, 3), (2, 40, 150, 64, 128, 3), (64, 56, 56, 64, 64, 3)]:
    x = torch.randn(B, H, W, Ci, device="cuda").to(torch.bfloat16)
    dy = torch.randn(B, H, W, Co, device="cuda").to(torch.bfloat16)
    raw.USE_HALO = False
    ref = raw.conv_wgrad(x, dy, k, k, 1)
    raw.USE_HALO = True
    out = torch.zeros_like(ref)
    pad = (k - 1) // 2
    rc = _lib.call_raw("dtf_conv_wgrad_halo_bf16", x, dy, out, B, H, W, Ci, k, k, Co, pad, pad, _lib.stream_ptr())
    torch.cuda.synchronize()
    err = (out - ref).abs().max().item()
    print("B%d %dx%d %d->%d k%d rc=%d err=%.4f ref_max=%.2f nan=%s" % (B, H, W, Ci, Co, k, rc, err, ref.abs().max().item(), bool(torch.isnan(out).any())), flush=True)
