#!/usr/bin/env python
"""Micro-benchmark of the fused ResBlock1 pair kernel (masg_respair) at the config-3 shapes of the narrow HiFi-GAN
stages, next to the two separate conv launches it replaces (masg_conv1d x 2).  Prints ms and algorithmic TFLOP/s / GB/s."""
import argparse, ctypes, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import masg_b200  # noqa
from masg_b200 import _lib

CASES = [  # name, B, T, C, k, dil
    ("s4 k3 d1", 64, 153600, 32, 3, 1), ("s4 k3 d5", 64, 153600, 32, 3, 5), ("s4 k7 d3", 64, 153600, 32, 7, 3),
    ("s4 k11 d1", 64, 153600, 32, 11, 1), ("s4 k11 d5", 64, 153600, 32, 11, 5),
    ("s3 k3 d1", 64, 76800, 64, 3, 1), ("s3 k3 d5", 64, 76800, 64, 3, 5), ("s3 k7 d3", 64, 76800, 64, 7, 3),
    # CTA-pair kernel (respair2_tc.cu)
    ("s3 k11 d1", 64, 76800, 64, 11, 1), ("s3 k11 d5", 64, 76800, 64, 11, 5), ("s2 k3 d1", 64, 38400, 128, 3, 1), ("s2 k3 d5", 64, 38400, 128, 3, 5),
]

def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--only", type=int, default=-1)
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--unfused", action="store_true", help="also time the two separate conv launches")
    a = ap.parse_args()
    cx = _lib.default_context("cuda:0", "bf16")
    P = _lib.ptr
    for i, (name, B, T, C, k, dil) in enumerate(CASES):
        if a.only >= 0 and i != a.only:
            continue
        g = torch.Generator().manual_seed(i)
        x = torch.randn(B, T, C, generator=g).cuda()
        w1 = (torch.randn(C, C, k, generator=g) / (C * k) ** 0.5).cuda(); w2 = (torch.randn(C, C, k, generator=g) / (C * k) ** 0.5).cuda()
        b1 = torch.randn(C, generator=g).cuda(); b2 = torch.randn(C, generator=g).cuda()
        y = torch.empty(B, T, C, device="cuda")
        ms = ctypes.c_float(0)
        cx.check(cx.lib.masg_respair(cx.h, cx.stream(), P(x), None, B, T, C, P(w1), P(b1), P(w2), P(b2), k, dil, 1.0, 1.0, a.reps, P(y),
                                     ctypes.byref(ms)), "respair")
        flops = 2 * 2.0 * B * T * k * C * C
        byts = 2.0 * B * T * C * 2
        line = "%2d %-10s fused %7.3f ms  %7.1f TFLOP/s  %7.1f GB/s (x in + y out)" % (i, name, ms.value, flops / ms.value / 1e9, byts / ms.value / 1e6)
        if a.unfused:
            m1, m2 = ctypes.c_float(0), ctypes.c_float(0)
            cx.check(cx.lib.masg_conv1d(cx.h, cx.stream(), P(x), B, T, C, P(w1), P(b1), C, k, dil, 0, 0.1, 0, 2, a.reps, P(y), ctypes.byref(m1)), "c1")
            cx.check(cx.lib.masg_conv1d(cx.h, cx.stream(), P(x), B, T, C, P(w2), P(b2), C, k, 1, 0, 0.0, 0, 6, a.reps, P(y), ctypes.byref(m2)), "c2")
            line += " | separate %.3f + %.3f = %.3f ms" % (m1.value, m2.value, m1.value + m2.value)
        print(line, flush=True)
        del x, y

if __name__ == "__main__":
    main()
