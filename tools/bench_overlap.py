#!/usr/bin/env python
"""Upper bound for overlapping the tail of one layer kernel with the start of the next: the pair is captured once as
two dependent launches on one stream and once as two INDEPENDENT launches on two streams (fork / join inside the graph;
different buffers, so the result is meaningless -- only the time counts).  With one CTA per SM the second kernel's CTAs
get SMs as the first kernel's CTAs retire, which is what a tile-level dependency between the two kernels would allow.

    python tools/bench_overlap.py [--rows 25216] [--reps 8]
"""
import argparse
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from gesturediffusion_b200 import kernels as K  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=25216)
    ap.add_argument("--reps", type=int, default=8)
    a = ap.parse_args()
    dev = torch.device("cuda", 0)
    M, d, F, S, H = a.rows, 512, 1024, 197, 4
    g = torch.Generator(device=dev).manual_seed(0)
    rnd = lambda *s, dt=torch.float16, sc=1.0: (torch.randn(*s, device=dev, generator=g) * sc).to(dt)
    Xa, Xlo, Xb = rnd(M, d), rnd(M, d, sc=1e-3), rnd(M, d)
    ATT, QKV = rnd(M, d), torch.empty(M, 3 * d, device=dev, dtype=torch.float16)
    Oh, Ol = torch.empty_like(Xa), torch.empty_like(Xa)
    w1, b1, w2, b2 = rnd(F, d, sc=0.04), rnd(F, dt=torch.float32), rnd(d, F, sc=0.03), rnd(d, dt=torch.float32)
    wq, bq = rnd(3 * d, d, sc=0.04), rnd(3 * d, dt=torch.float32)
    wo, bo = rnd(d, d, sc=0.04), rnd(d, dt=torch.float32)
    gam, bet = rnd(d, dt=torch.float32), rnd(d, dt=torch.float32)
    ffn = lambda: K.ffn_ln_split(Xa, Xlo, w1, b1, w2, b2, gam, bet, out_hi=Oh, out_lo=Ol)
    qkv = lambda: K.gemm(Xb, wq, K.epilogue(bias=bq, out_h16=QKV))
    oln = lambda: K.gemm_ln_split(ATT, wo, bo, gam, bet, out_hi=Oh, out_lo=Ol, res_hi=Xa, res_lo=Xlo)
    att = lambda: K.attention(QKV, ATT, M // S, S, H, d // H)
    pairs = {"ffn -> qkv": (ffn, qkv), "out_proj+LN -> ffn": (oln, ffn), "qkv -> attention": (qkv, att),
             "attention -> out_proj+LN": (att, oln)}
    for f in (ffn, qkv, oln, att):
        f()
    torch.cuda.synchronize()
    side = torch.cuda.Stream(device=dev)

    def timed(graph):
        graph.replay()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        graph.replay()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) * 1e3 / a.reps

    for name, (f1, f2) in pairs.items():
        gs = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gs):
            for _ in range(a.reps):
                f1()
                f2()
        gp = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gp):
            cur = torch.cuda.current_stream()
            for _ in range(a.reps):
                side.wait_stream(cur)
                f1()
                with torch.cuda.stream(side):
                    f2()
                cur.wait_stream(side)
        print("%-28s dependent %7.1f us   independent (two streams) %7.1f us" % (name, timed(gs), timed(gp)))


if __name__ == "__main__":
    main()
