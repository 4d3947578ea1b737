import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from distributed_tensorflow_framework_b200.ops import raw
B, H, Ci, Co, k, s = 64, 56, 64, 64, 3, 1
if len(sys.argv) > 1:
    B, H, Ci, Co, k, s = [int(v) for v in sys.argv[1:7]]
mode = sys.argv[7] if len(sys.argv) > 7 else "fprop"
x = torch.randn(B, H, H, Ci, device="cuda").to(torch.bfloat16)
w = (torch.randn(k, k, Ci, Co, device="cuda") * 0.05).to(torch.bfloat16)
w2 = w.reshape(k * k * Ci, Co); wt = w.permute(0, 1, 3, 2).contiguous().reshape(k * k * Co, Ci)
y = raw.conv_fprop(x, w2, None, k, k, s); dy = torch.randn_like(y); dw = torch.zeros(k * k * Ci, Co, device="cuda")
for _ in range(4):
    if mode == "fprop": raw.conv_fprop(x, w2, None, k, k, s)
    elif mode == "dgrad": raw.conv_dgrad(dy, wt, (H, H), k, k, s)
    else: raw.conv_wgrad(x, dy, k, k, s, out=dw)
torch.cuda.synchronize()
print("done")
