#!/usr/bin/env python
"""The cluster / tcgen05 kernels at one-tile and multi-round shapes, for `compute-sanitizer` (race detection):

    compute-sanitizer --tool memcheck  python tools/sanitize_kernels.py      (one tool per gpurun call)
    compute-sanitizer --tool racecheck python tools/sanitize_kernels.py

Runs gd_gemm_h16 (three epilogues), gd_gemm_h16_ln_split, gd_ffn_h16_ln_split, gd_attention_h16 (short and block-wise
kernels), gd_local_attention and gd_posterior_step once each per shape and checks the results against torch, so a
sanitizer run that changes the numerics is caught too.  Shapes: one row tile (M = 100), one full round (M = 74 x 128),
and several rounds with a ragged last tile (M = 20 000 + 37 for the GEMMs).
"""
import math
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from gesturediffusion_b200 import kernels as K  # noqa: E402


def rel(a, b):
    return float((a.double() - b.double()).norm() / b.double().norm())


def main():
    big = "--big" in sys.argv
    dev = torch.device("cuda", 0)
    g = torch.Generator(device=dev).manual_seed(0)
    dt = torch.float16
    rnd = lambda *s, sc=1.0: torch.randn(*s, device=dev, generator=g) * sc
    d, F = 512, 1024
    for M in ((100, 74 * 128, 20037) if big else (100, 74 * 128 + 5)):
        x = rnd(M, d)
        hi = x.to(dt)
        lo = (x - hi.float()).to(dt)
        w_qkv, b_qkv = rnd(3 * d, d, sc=d ** -0.5).to(dt), rnd(3 * d)
        qkv = torch.empty(M, 3 * d, device=dev, dtype=dt)
        K.gemm(hi, w_qkv, K.epilogue(bias=b_qkv, out_h16=qkv))
        assert rel(qkv.float(), hi.double() @ w_qkv.double().T + b_qkv.double()) < 1e-3
        w1, b1 = rnd(F, d, sc=d ** -0.5).to(dt), rnd(F, sc=0.5)
        w2, b2 = rnd(d, F, sc=F ** -0.5).to(dt), rnd(d)
        gam, bet = 1 + 0.1 * rnd(d), 0.1 * rnd(d)
        h = torch.empty(M, F, device=dev, dtype=dt)
        K.gemm(hi, w1, K.epilogue(bias=b1, act=K.GD_ACT_GELU_ERF, out_h16=h))
        oh, ol = torch.empty_like(hi), torch.empty_like(lo)
        K.gemm_ln_split(h, w2, b2, gam, bet, out_hi=oh, out_lo=ol, res_hi=hi, res_lo=lo, K=F)
        fh, fl = torch.empty_like(hi), torch.empty_like(lo)
        K.ffn_ln_split(hi, lo, w1, b1, w2, b2, gam, bet, out_hi=fh, out_lo=fl)
        ref = torch.nn.functional.layer_norm(
            hi.double() + lo.double() + torch.nn.functional.gelu(hi.double() @ w1.double().T + b1.double()).to(dt).double()
            @ w2.double().T + b2.double(), (d,), gam.double(), bet.double(), 1e-5)
        assert rel(oh.float() + ol.float(), ref) < 2e-4 and rel(fh.float() + fl.float(), ref) < 2e-4
        o32 = torch.empty(M, d, device=dev)
        K.gemm(hi, w2[:, :d].contiguous(), K.epilogue(bias=b2, residual=x, out_f32=o32))
        print("M=%d gemm / gemm_ln / ffn_ln ok" % M, flush=True)
    for n_seq, S in ((2, 197), (40, 197), (2, 121), (2, 301), (3, 601)):
        qkv = rnd(n_seq * S, 3 * d).to(dt)
        out = torch.empty(n_seq * S, d, device=dev, dtype=dt)
        K.attention(qkv, out, n_seq, S, 4, 128)
        q, k, v = qkv.double().view(n_seq, S, 3, 4, 128).permute(2, 0, 3, 1, 4)
        ref = (torch.softmax(q @ k.transpose(-1, -2) / math.sqrt(128), -1) @ v).permute(0, 2, 1, 3).reshape(n_seq * S, d)
        assert rel(out.float(), ref) < 1e-3
        print("attention n_seq=%d S=%d ok" % (n_seq, S), flush=True)
    torch.cuda.synchronize()
    print("SANITIZE-RUN-OK")


if __name__ == "__main__":
    main()
