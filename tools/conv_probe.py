#!/usr/bin/env python
"""Per-layer micro-benchmark of the dense-contraction kernel through masg_conv1d (C-ABI) at the config-3 shapes of
the HiFi-GAN generator and the FastSpeech2 decoder.  Prints ms, algorithmic TFLOP/s and the algorithmic HBM GB/s of
each layer class.  `--only i` runs a single case (for ncu)."""
import argparse, ctypes, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import masg_b200  # noqa
from masg_b200 import _lib

# name, B, T, Cin, Cout, k, dil, ups, slope, flags(2=bf16 acts,4=residual,1=split,8=fp16 single pass)
CASES = [
    ("hg s4 c1 k3 d1", 64, 153600, 32, 32, 3, 1, 0, 0.1, 2),
    ("hg s4 c2 k3 +res", 64, 153600, 32, 32, 3, 1, 0, 0.1, 6),
    ("hg s4 c1 k11 d5", 64, 153600, 32, 32, 11, 5, 0, 0.1, 2),
    ("hg s3 c1 k7 d3", 64, 76800, 64, 64, 7, 3, 0, 0.1, 2),
    ("hg s3 c2 k11 +res", 64, 76800, 64, 64, 11, 1, 0, 0.1, 6),
    ("hg s2 c1 k3 d1", 64, 38400, 128, 128, 3, 1, 0, 0.1, 2),
    ("hg s2 c2 k7 +res", 64, 38400, 128, 128, 7, 1, 0, 0.1, 6),
    ("hg s2 c1 k11 d5", 64, 38400, 128, 128, 11, 5, 0, 0.1, 2),
    ("hg s1 c1 k11 d5", 64, 4800, 256, 256, 11, 5, 0, 0.1, 2),
    ("hg up1 256->128 x8", 64, 4800, 256, 128, 16, 1, 8, 0.1, 2),
    ("fs2 dec w1 k9 (split)", 64, 600, 256, 1024, 9, 1, 0, 0.0, 1),
    ("fs2 dec w2 k1 (split)", 64, 600, 1024, 256, 1, 1, 0, 0.0, 1),
    ("fs2 dec qkv (split)", 64, 600, 256, 768, 1, 1, 0, 0.0, 1),
    ("postnet 512 k5 (split)", 64, 600, 512, 512, 5, 1, 0, 0.0, 1),
    ("fs2 dec w1 k9 (fp16)", 64, 600, 256, 1024, 9, 1, 0, 0.0, 8),
    ("fs2 dec w2 k1 (fp16)", 64, 600, 1024, 256, 1, 1, 0, 0.0, 8),
    ("fs2 dec qkv (fp16)", 64, 600, 256, 768, 1, 1, 0, 0.0, 8),
    ("fs2 dec fc (fp16)", 64, 600, 256, 256, 1, 1, 0, 0.0, 8),
    ("postnet 512 k5 (fp16)", 64, 600, 512, 512, 5, 1, 0, 0.0, 8),
    ("hg s2 c2 k3 +res", 64, 38400, 128, 128, 3, 1, 0, 0.1, 6),
    ("hg s3 c1 k11 d5", 64, 76800, 64, 64, 11, 5, 0, 0.1, 2),
    ("hg s3 c1 k11 d1", 64, 76800, 64, 64, 11, 1, 0, 0.1, 2),
    ("hg s3 c2 k11 +res (own buffer)", 64, 76800, 64, 64, 11, 1, 0, 0.1, 22),
    ("hg s2 c2 k7 +res (own buffer)", 64, 38400, 128, 128, 7, 1, 0, 0.1, 22),
    ("hg s2 c2 k3 +res (own buffer)", 64, 38400, 128, 128, 3, 1, 0, 0.1, 22),
]

def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--only", type=int, default=-1)
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--precision", default="bf16")
    a = ap.parse_args()
    cx = _lib.default_context("cuda:0", a.precision)
    for i, (name, B, T, Cin, Cout, k, dil, ups, slope, flags) in enumerate(CASES):
        if a.only >= 0 and i != a.only:
            continue
        g = torch.Generator().manual_seed(i)
        x = torch.randn(B, T, Cin, generator=g).cuda()
        w = (torch.randn((Cin, Cout, k) if ups else (Cout, Cin, k), generator=g) / (Cin * k) ** 0.5).cuda()
        bias = torch.randn(Cout, generator=g).cuda()
        So = T * ups if ups else T
        y = torch.empty(B, So, Cout, device="cuda")
        ms = ctypes.c_float(0)
        cx.check(cx.lib.masg_conv1d(cx.h, cx.stream(), _lib.ptr(x), B, T, Cin, _lib.ptr(w), _lib.ptr(bias), Cout, k, dil, ups, slope, 0,
                                    flags if a.precision == "bf16" else 0, a.reps, _lib.ptr(y), ctypes.byref(ms)), "conv1d")
        taps = 2 if ups else k
        N = ups * Cout if ups else Cout
        flops = 2.0 * B * T * taps * Cin * N
        esz = 2 if (flags & 2) else 4
        byts = B * T * Cin * esz * (2 if flags & 4 else 1) + B * So * Cout * esz
        print("%2d %-32s %8.3f ms  %7.1f TFLOP/s  %7.1f GB/s" % (i, name, ms.value, flops / ms.value / 1e9, byts / ms.value / 1e6), flush=True)
        del x, y

if __name__ == "__main__":
    main()
