"""Roofline position of every ResNet-50 convolution (fprop / dgrad / wgrad) from the measured per-layer table
(profiles/r2/per_layer_vs_cudnn.txt, batch 256) and the measured peaks (MEASURED_PEAKS.json): ideal time =
max(flops / cuBLAS bf16 rate, algorithmic bytes / copy bandwidth). Runs on the CPU box."""
import json
import os
import re
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
flat = json.dumps(peaks)
bw = float(re.search(r'"(?:hbm|copy)[^"]*gbs?[^"]*":\s*([0-9.]+)', flat, re.I).group(1)) * 1e9
tf = float(re.search(r'"[^"]*(?:tflops|bf16)[^"]*":\s*([0-9.]+)', flat, re.I).group(1)) * 1e12
B = 256
rows = []
for line in open(os.path.join(ROOT, "profiles", "r2", "per_layer_vs_cudnn.txt")):
    m = re.match(r"\s*(\d+)x(\d+)\s+(\d+)->(\d+)\s+k(\d) s(\d) x(\d+)\s+([\d.]+)\s+([\d.]+)\s+([\d.]+) \|\s+([\d.]+)\s+([\d.]+)\s+([\d.]+)", line)
    if not m:
        continue
    H, _, Ci, Co, k, s, cnt = (int(m.group(i)) for i in range(1, 8))
    ours = [float(m.group(i)) for i in (8, 9, 10)]
    cud = [float(m.group(i)) for i in (11, 12, 13)]
    OH = -(-H // s)
    flops = 2.0 * B * OH * OH * Co * k * k * Ci
    xb, yb, wb = 2.0 * B * H * H * Ci, 2.0 * B * OH * OH * Co, 2.0 * k * k * Ci * Co
    bytes_ = [xb + yb + wb, xb + yb + wb, xb + yb + 2 * wb]  # wgrad writes fp32 filters
    ideal = [max(flops / tf, b / bw) * 1e3 for b in bytes_]
    rows.append((H, Ci, Co, k, s, cnt, ours, cud, ideal, ["mem" if b / bw > flops / tf else "math" for b in bytes_]))
print("# Roofline position of the ResNet-50 convolutions (batch 256, bf16, 1x B200)\n")
print("Ideal = max(flops / %.0f TFLOP/s measured cuBLAS bf16, algorithmic bytes / %.2f TB/s measured copy bandwidth);" % (tf / 1e12, bw / 1e12))
print("times in ms from `per_layer_vs_cudnn.txt` (CUDA events, cold L2). `x` = layers of that shape in the net.\n")
print("| layer | x | bound | ideal f/d/w | ours f/d/w | ours / ideal | cuDNN / ideal |")
print("|---|---|---|---|---|---|---|")
tot = {"i": 0.0, "o": 0.0, "c": 0.0}
for (H, Ci, Co, k, s, cnt, ours, cud, ideal, bound) in rows:
    passes = (0, 1, 2) if not (H == 224) else (0, 2)  # the stem has no data gradient in training
    for p in passes:
        tot["i"] += ideal[p] * cnt
        tot["o"] += ours[p] * cnt
        tot["c"] += cud[p] * cnt
    print("| %dx%d %d->%d k%d s%d | %d | %s | %s | %s | %s | %s |" % (
        H, H, Ci, Co, k, s, cnt, "/".join(bound), " / ".join("%.3f" % v for v in ideal), " / ".join("%.3f" % v for v in ours),
        " / ".join("%.1f" % (o / i) for o, i in zip(ours, ideal)), " / ".join("%.1f" % (c / i) for c, i in zip(cud, ideal))))
print("\nWhole net (weighted by layer count, stem data gradient excluded): ideal %.2f ms, ours %.2f ms (%.1fx), cuDNN %.2f ms (%.1fx)."
      % (tot["i"], tot["o"], tot["o"] / tot["i"], tot["c"], tot["c"] / tot["i"]))
