"""Every layer of the library on CUDA behind a 1x1 convolution that produces C channels (C not a multiple of 8) on odd
spatial sizes: forward, loss and the update of the 1x1 filter (i.e. the layer's backward) against the CPU path from the
same variables (tests/scenarios.layer_case_parity). One line per case; failures are reported, not raised."""
import os
import sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import scenarios  # noqa: E402

bad = 0
for (n, h, w, C) in ((4, 9, 13, 20), (3, 8, 8, 3), (2, 7, 7, 36), (5, 12, 10, 64)):
    for name in scenarios.LAYER_CASES:
        try:
            k, shape = scenarios.layer_case_parity(name, n, h, w, C)
            print("%-20s C=%-3d %dx%d ok   (%d variables compared, out %s)" % (name, C, h, w, k, shape), flush=True)
        except Exception as e:  # noqa
            bad += 1
            print("%-20s C=%-3d %dx%d FAIL %s: %s" % (name, C, h, w, type(e).__name__,
                                                       str(e)[:170].replace("\n", " ")), flush=True)
print("bad cases:", bad)
