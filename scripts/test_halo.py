import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from distributed_tensorflow_framework_b200.ops import raw, _lib
torch.manual_seed(0)
def run(B, H, W, Ci, Co, k, mode):
    x = torch.randn(B, H, W, Ci, device="cuda").to(torch.bfloat16)
    w = (torch.randn(k, k, Ci, Co, device="cuda") * 0.05).to(torch.bfloat16)
    w2 = w.reshape(k * k * Ci, Co)
    bias = torch.randn(Co, device="cuda")
    ref = raw.conv_fprop(x, w2, bias, k, k, 1, relu=True)       # gather kernel (validated against torch)
    out = torch.empty_like(ref)
    pad = (k - 1) // 2
    rc = _lib.call_raw("dtf_conv_halo_bf16", x, w2, out, bias, B, H, W, Ci, k, k, Co, pad, pad, 0, 1 | (mode << 8), _lib.stream_ptr())
    torch.cuda.synchronize()
    err = (out.float() - ref.float()).abs().max().item()
    # data gradient
    dy = torch.randn(B, H, W, Co, device="cuda").to(torch.bfloat16)
    wt = w.permute(0, 1, 3, 2).contiguous().reshape(k * k * Co, Ci)
    dref = raw.conv_dgrad(dy, wt, (H, W), k, k, 1)
    dout = torch.empty_like(dref)
    rc2 = _lib.call_raw("dtf_conv_halo_bf16", dy, wt, dout, None, B, H, W, Co, k, k, Ci, pad, pad, 1, (mode << 8), _lib.stream_ptr())
    torch.cuda.synchronize()
    derr = (dout.float() - dref.float()).abs().max().item()
    print("B%d %dx%d %d->%d k%d base_off_mode=%d rc=%d/%d fprop_err=%.4f (ref max %.2f) dgrad_err=%.4f (ref max %.2f)" % (
        B, H, W, Ci, Co, k, mode, rc, rc2, err, ref.float().abs().max().item(), derr, dref.float().abs().max().item()), flush=True)
for mode in (0, 1):
    for cfg in [(2, 14, 14, 64, 64, 3), (2, 56, 56, 64, 64, 3), (3, 28, 28, 128, 128, 3), (2, 14, 14, 256, 256, 3), (2, 16, 16, 64, 64, 5), (2, 17, 19, 64, 128, 3)]:
        try:
            run(*cfg, mode)
        except Exception as e:
            print("EXC", cfg, mode, str(e)[:100]); break
