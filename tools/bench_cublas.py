#!/usr/bin/env python
"""Yardstick only (never on the product path): cuBLAS/cuBLASLt through torch for the encoder GEMM shapes, timed like
tools/bench_kernels.py (CUDA events, operand sets rotated).  Bias-only epilogue (F.linear)."""
import json
import sys

import torch

M = int(sys.argv[1]) if len(sys.argv) > 1 else 25216
dev = torch.device("cuda", 0)
ns, iters = 3, 24


def timed(fn):
    for i in range(3):
        fn(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(iters):
        fn(i)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters * 1e3


out = {}
for name, N, K in (("qkv", 1536, 512), ("o", 512, 512), ("ffn1", 1024, 512), ("ffn2", 512, 1024)):
    a = [torch.randn(M, K, device=dev, dtype=torch.float16) for _ in range(ns)]
    w = torch.randn(N, K, device=dev, dtype=torch.float16) * 0.04
    b = torch.randn(N, device=dev, dtype=torch.float16)
    o = [torch.empty(M, N, device=dev, dtype=torch.float16) for _ in range(ns)]
    us = timed(lambda i: torch.nn.functional.linear(a[i % ns], w, b))
    out[name] = us
    print("cuBLAS linear %-5s M=%d N=%d K=%d: %7.1f us  %7.1f TFLOP/s" % (name, M, N, K, us, 2.0 * M * N * K / us / 1e6))
print(json.dumps(out))
