#!/usr/bin/env python
"""Timeline of one CTA of the fused feed-forward kernel (gd_ffn_trace): prints, per role, the SM-clock deltas
between the tagged points of ffn_ln.cu (tags: producer 1/2 = GEMM1/GEMM2 loads start; MMA 10/11/12 = GEMM1
wait/start/issued, 20/21/22 = GEMM2 wait/start/issued, 30 = tile done; epilogue 40..45 = H chunk: wait, accumulator
ready, TMEM drained, arithmetic done, slot free, written + copy issued; 50..55 = LayerNorm tail: wait, Y ready,
pass A done, statistics in, pass B done, staging read)."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from gesturediffusion_b200 import _cabi, kernels as K  # noqa: E402


def main():
    dev = torch.device("cuda", 0)
    M, d, F = 25216, 512, 1024
    g = torch.Generator(device=dev).manual_seed(0)
    rnd = lambda *s, dt=torch.float16, sc=1.0: (torch.randn(*s, device=dev, generator=g) * sc).to(dt)
    hi, lo = rnd(M, d), rnd(M, d, sc=1e-3)
    w1, w2 = rnd(F, d, sc=0.04), rnd(d, F, sc=0.03)
    b1, b2, gam, bet = (rnd(n, dt=torch.float32) for n in (F, d, d, d))
    oh, ol = torch.empty_like(hi), torch.empty_like(lo)

    def run():
        K.ffn_ln_split(hi, lo, w1, b1, w2, b2, gam, bet, out_hi=oh, out_lo=ol)
    for _ in range(3):
        run()
    roles = ("epilogue warp 2", "mma", "producer")
    buf = torch.zeros(len(roles) * 4096, dtype=torch.int64, device=dev)
    lib = _cabi.load()
    lib.gd_ffn_trace(buf.data_ptr())
    run()
    torch.cuda.synchronize()
    lib.gd_ffn_trace(None)
    b = buf.cpu().tolist()
    t0 = min((v & ((1 << 48) - 1)) for v in b if v)
    for role, name in enumerate(roles):
        rec = [(v >> 48, (v & ((1 << 48) - 1)) - t0) for v in b[role * 4096:(role + 1) * 4096] if v]
        print("== %s: %d records" % (name, len(rec)))
        prev = 0
        for tag, c in rec:
            print("  tag %2d  t=%7d  (+%d)" % (tag, c, c - prev))
            prev = c


if __name__ == "__main__":
    main()
