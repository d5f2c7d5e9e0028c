import sys, torch
sys.path.insert(0, "/root/repo")
from gesturediffusion_b200 import kernels as K
dev = torch.device("cuda", 0)
def timed(fn, iters=200):
    for _ in range(5): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters * 1e3
def graph_timed(fn, n=50, iters=20):
    fn(); torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(n): fn()
    g.replay(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters): g.replay()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters / n * 1e3
for (M, N, Kk) in [(128, 256, 64), (128*148, 256, 64), (128*148, 256, 512), (25216, 1536, 64), (25216, 1536, 512), (25216, 1024, 512)]:
    a = torch.randn(M, Kk, device=dev).half(); w = torch.randn(N, Kk, device=dev).half(); b = torch.zeros(N, device=dev)
    o = torch.empty(M, N, device=dev, dtype=torch.float16)
    ep = K.epilogue(bias=b, out_h16=o)
    f = lambda: K.gemm(a, w, ep)
    print("gemm M=%d N=%d K=%d: stream %.1f us, graph %.1f us" % (M, N, Kk, timed(f, 50), graph_timed(f)))
