#!/usr/bin/env python
"""Per-kernel timing of the encoder-layer kernels at a bench workload's shapes (CUDA events, warm, looped).

    python tools/bench_kernels.py [--rows 25216] [--seq 197] [--iters 20] [--fmt fp16]

Rotates over several buffer sets larger than L2 in total so that no launch finds its operands L2-resident from
its own previous launch (weights stay hot, as in the real step).  Prints microseconds per launch and the
achieved TFLOP/s or GB/s against MEASURED_PEAKS.json.
"""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from gesturediffusion_b200 import kernels as K  # noqa: E402


GRAPH = "--graph" in sys.argv        # time `iters` launches replayed from ONE CUDA graph: no host launch cost in the figure
                                     # (a ~25 us kernel whose wrapper costs ~25 us of ctypes marshalling is host-bound otherwise)


def timed(fn, iters):
    for i in range(3):
        fn(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    if GRAPH:
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            for i in range(iters):
                fn(i)
        g.replay()
        torch.cuda.synchronize()
        e0.record()
        g.replay()
        e1.record()
    else:
        e0.record()
        for i in range(iters):
            fn(i)
        e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters * 1e3        # us


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=0)
    ap.add_argument("--seq", type=int, default=197)
    ap.add_argument("--nseq", type=int, default=128)
    ap.add_argument("--iters", type=int, default=24)
    ap.add_argument("--fmt", default="fp16")
    ap.add_argument("--sets", type=int, default=3)
    ap.add_argument("--only", default="")
    ap.add_argument("--graph", action="store_true")
    a = ap.parse_args()
    dev = torch.device("cuda", 0)
    act = torch.float16 if a.fmt == "fp16" else torch.bfloat16
    d, F, H = 512, 1024, 4
    S, nseq = a.seq, a.nseq
    M = a.rows or S * nseq
    peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(
        os.path.join(ROOT, "MEASURED_PEAKS.json")) else {"bf16_tflops": 1590.0, "hbm_gbs": 6650.0}
    g = torch.Generator(device=dev).manual_seed(0)

    def rnd(*shape, dtype=torch.float32, scale=1.0):
        return (torch.randn(*shape, device=dev, generator=g) * scale).to(dtype)

    ns = a.sets
    Xa = [rnd(M, d, dtype=act) for _ in range(ns)]
    X32 = [rnd(M, d) for _ in range(ns)]
    Y32 = [torch.empty(M, d, device=dev) for _ in range(ns)]
    QKV = [rnd(M, 3 * d, dtype=act) for _ in range(ns)]
    ATT = [rnd(M, d, dtype=act) for _ in range(ns)]
    Hh = [rnd(M, F, dtype=act) for _ in range(ns)]
    w_qkv, b_qkv = rnd(3 * d, d, dtype=act, scale=0.04), rnd(3 * d)
    w_o, b_o = rnd(d, d, dtype=act, scale=0.04), rnd(d)
    w_1, b_1 = rnd(F, d, dtype=act, scale=0.04), rnd(F)
    w_2, b_2 = rnd(d, F, dtype=act, scale=0.03), rnd(d)
    gam, bet = rnd(d), rnd(d)
    out = {}

    def rec(name, us, flops=None, nbytes=None):
        line = "%-34s %8.1f us" % (name, us)
        if flops:
            tf = flops / us / 1e6
            line += "  %7.1f TFLOP/s (%.2f of burst peak)" % (tf, tf / peaks["bf16_tflops"])
        if nbytes:
            gb = nbytes / us / 1e3
            line += "  %7.1f GB/s algorithmic (%.2f of HBM peak)" % (gb, gb / peaks["hbm_gbs"])
        print(line)
        out[name] = us

    esz = 2
    want = set(a.only.split(",")) if a.only else None

    def on(n):
        return want is None or n in want

    if on("qkv"):
        rec("gemm qkv  N=1536 K=512 ->h16", timed(lambda i: K.gemm(Xa[i % ns], w_qkv, K.epilogue(bias=b_qkv, out_h16=QKV[i % ns])), a.iters),
            flops=2.0 * M * 3 * d * d, nbytes=M * d * esz + M * 3 * d * esz)
    if on("attn"):
        rec("attention S=%d H=4 dh=128" % S, timed(lambda i: K.attention(QKV[i % ns], ATT[i % ns], M // S, S, H, d // H), a.iters),
            flops=4.0 * (M // S) * H * S * S * (d // H), nbytes=M * 4 * d * esz)
    if on("oproj"):
        rec("gemm o    N=512 K=512 +res ->f32", timed(lambda i: K.gemm(ATT[i % ns], w_o, K.epilogue(bias=b_o, residual=X32[i % ns], out_f32=Y32[i % ns])), a.iters),
            flops=2.0 * M * d * d, nbytes=M * d * (esz + 4 + 4))
    if on("ln"):
        rec("layernorm d=512 ->f32+h16", timed(lambda i: K.layernorm(Y32[i % ns], gam, bet, out_f32=X32[i % ns], out_h16=Xa[i % ns]), a.iters),
            nbytes=M * d * (4 + 4 + esz))
    if on("ffn1"):
        rec("gemm ffn1 N=1024 K=512 gelu ->h16", timed(lambda i: K.gemm(Xa[i % ns], w_1, K.epilogue(bias=b_1, act=K.GD_ACT_GELU_ERF, out_h16=Hh[i % ns])), a.iters),
            flops=2.0 * M * F * d, nbytes=M * d * esz + M * F * esz)
    if on("ffn2"):
        rec("gemm ffn2 N=512 K=1024 +res ->f32", timed(lambda i: K.gemm(Hh[i % ns], w_2, K.epilogue(bias=b_2, residual=X32[i % ns], out_f32=Y32[i % ns])), a.iters),
            flops=2.0 * M * d * F, nbytes=M * F * esz + M * d * 8)
    if on("oln"):
        rec("gemm o+LN   N=512 K=512 ->f32+h16", timed(lambda i: K.gemm_ln(ATT[i % ns], w_o, b_o, X32[i % ns], gam, bet, out_f32=X32[i % ns], out_h16=Xa[i % ns]), a.iters),
            flops=2.0 * M * d * d, nbytes=M * d * (esz + 4 + 4 + esz))
    if on("f2ln"):
        rec("gemm ffn2+LN N=512 K=1024 ->f32+h16", timed(lambda i: K.gemm_ln(Hh[i % ns], w_2, b_2, X32[i % ns], gam, bet, out_f32=X32[i % ns], out_h16=Xa[i % ns]), a.iters),
            flops=2.0 * M * d * F, nbytes=M * F * esz + M * d * (4 + 4 + esz))
    Xlo = [rnd(M, d, dtype=act, scale=1e-3) for _ in range(ns)]
    if on("olns"):
        rec("gemm o+LN   split residual (hi,lo)", timed(lambda i: K.gemm_ln_split(ATT[i % ns], w_o, b_o, gam, bet, out_hi=Xa[i % ns], out_lo=Xlo[i % ns], res_hi=Xa[i % ns], res_lo=Xlo[i % ns]), a.iters),
            flops=2.0 * M * d * d, nbytes=M * d * esz + M * d * 8)
    if on("f2lns"):
        rec("gemm ffn2+LN split residual (hi,lo)", timed(lambda i: K.gemm_ln_split(Hh[i % ns], w_2, b_2, gam, bet, out_hi=Xa[i % ns], out_lo=Xlo[i % ns], res_hi=Xa[i % ns], res_lo=Xlo[i % ns]), a.iters),
            flops=2.0 * M * d * F, nbytes=M * F * esz + M * d * 8)
    if on("ffnln"):
        rec("ffn1+gelu+ffn2+LN fused (hi,lo)", timed(lambda i: K.ffn_ln_split(Xa[i % ns], Xlo[i % ns], w_1, b_1, w_2, b_2, gam, bet, out_hi=Xa[i % ns], out_lo=Xlo[i % ns]), a.iters),
            flops=4.0 * M * d * F, nbytes=M * d * 8)
    if on("seed"):
        for (mb, kk) in ((128, 4980), (64, 2630), (1, 2630)):   # EmbedSeedPose: [B, 10 J] x [512, 10 J]^T, once per sampling call
            sa, sw, sb = rnd(mb, kk), rnd(d, kk, scale=0.02), rnd(d)
            so = torch.empty(mb, d, device=dev)
            rec("gemm_f32 seed pose M=%d K=%d" % (mb, kk), timed(lambda i: K.gemm(sa, sw, K.epilogue(bias=sb, out_f32=so)), a.iters),
                flops=2.0 * mb * d * kk, nbytes=4.0 * (mb * kk + d * kk + mb * d))
    if on("inproj") or on("outproj"):
        B, T, nb, J, kin = nseq // 2, S - 1, 2, 263, 296
        a_in = [rnd(B * T, kin, dtype=act) for _ in range(ns)]
        w_in, b_in, pe = rnd(d, kin, dtype=act, scale=0.05), rnd(d), rnd(5000, d)
        w_out, b_out = rnd(J, d, dtype=act, scale=0.04), rnd(J)
        OUT = [torch.empty(nb * B, J, T, device=dev) for _ in range(ns)]
        if on("inproj"):
            rec("gemm in   N=512 K=289 pe dup2 ->f32+h16", timed(lambda i: K.gemm(
                a_in[i % ns], w_in, K.epilogue(bias=b_in, rowtab=pe, tab_period=T, tab_offset=1, rowmap=(T, S, 1), dup=nb,
                                               dup_row_stride=B * S, out_f32=X32[i % ns], out_h16=Xa[i % ns]), K=289), a.iters),
                flops=2.0 * B * T * d * 289, nbytes=B * T * kin * esz + nb * B * T * d * (4 + esz))
        if on("outproj"):
            rec("gemm out  N=263 K=512 ->f32 [B,J,T]", timed(lambda i: K.gemm(
                Xa[i % ns], w_out, K.epilogue(bias=b_out, out_t=OUT[i % ns], t_s=S, t_t=T, t_skip=1), K=d), a.iters),
                flops=2.0 * M * J * d, nbytes=M * d * esz + nb * B * J * T * 4)
    if on("inproj2"):
        B2, T2, nb2, kin2 = 128, 120, 2, 504
        a2 = [rnd(B2 * T2, kin2, dtype=act) for _ in range(ns)]
        w2 = rnd(d, kin2, dtype=act, scale=0.05)
        AM = rnd(B2 * T2, d)
        co = rnd(nb2 * B2, d)
        XL = [torch.empty(nb2 * B2 * T2, d, device=dev) for _ in range(ns)]
        rec("gemm in V2 N=512 K=498 AM+coarse dup2 ->f32", timed(lambda i: K.gemm(
            a2[i % ns], w2, K.epilogue(rowtab=AM, tab_period=B2 * T2, tab_offset=0, seqbias=co, seq_rows=T2, seq_count=B2,
                                       dup=nb2, dup_row_stride=B2 * T2, out_f32=XL[i % ns]), K=498), a.iters),
            flops=2.0 * B2 * T2 * d * 498, nbytes=B2 * T2 * (kin2 * esz + d * 4) + nb2 * B2 * T2 * d * 4)
    if on("post"):
        for (Bp, Jp, Tp) in ((64, 263, 196), (128, 498, 120)):
            OUTp = [rnd(2 * Bp, Jp, Tp) for _ in range(ns)]
            Xp = [rnd(Bp, Jp, Tp) for _ in range(ns)]
            kin = (Jp + 26 + 7) // 8 * 8
            Ap = [torch.zeros(Bp * Tp, kin, device=dev, dtype=act) for _ in range(ns)]
            coef = torch.rand(6, 50, device=dev) * 0.5
            stepp = torch.full((1,), 45, dtype=torch.int32, device=dev)
            scale = torch.full((Bp,), 2.5, device=dev)
            tmap = torch.arange(50, device=dev, dtype=torch.int64) * 20
            nxt = torch.zeros(Bp, dtype=torch.int64, device=dev)

            def run(i):
                K.posterior_step(OUTp[i % ns][:Bp], Xp[i % ns], coef, stepp, Bp, Jp, Tp, uncond=OUTp[i % ns][Bp:],
                                 scale=scale, x_prev=Xp[i % ns], a_next=Ap[i % ns], timestep_map=tmap, next_t=nxt,
                                 seed=3, sample_base=0)
            rec("posterior B=%d J=%d T=%d cfg philox" % (Bp, Jp, Tp), timed(run, a.iters),
                nbytes=Bp * Jp * Tp * (4 * 4 + esz))
    if on("lattn"):
        nsq, T2, hd = 256, 120, 8
        e2 = d // hd
        pos = torch.arange(T2 + 1, device=dev, dtype=torch.float32)[:, None]
        inv = 1.0 / (10000 ** (torch.arange(0, e2, 2, device=dev, dtype=torch.float32) / e2))
        ang = torch.cat([pos * inv, pos * inv], dim=1)
        rc, rs = ang.cos().contiguous(), ang.sin().contiguous()
        ZL = [rnd(nsq * T2, d) for _ in range(ns)]
        XO = [torch.zeros(nsq * (T2 + 1), d, device=dev) for _ in range(ns)]
        XH = [torch.zeros(nsq * (T2 + 1), d, device=dev, dtype=act) for _ in range(ns)]
        rec("local attention T=120 e=64 w=10", timed(lambda i: K.local_attention(ZL[i % ns], rc, rs, XO[i % ns], XH[i % ns], nsq, T2, d, hd, 10), a.iters),
            nbytes=nsq * T2 * d * (4 + 4 + esz))
    print(json.dumps({"rows": M, "S": S, "us": out}))


if __name__ == "__main__":
    main()
