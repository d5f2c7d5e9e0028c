#!/usr/bin/env python
"""Throughput sweep of the sampling loop over batch, sequence length, schedule length and GPU count
(BASELINE.json configs[4]):

    python tools/sweep.py [--arch v1] [--batches 1,4,16,64,256,1024,4096] [--frames 60,120,200,300,400,600]
                          [--steps 50] [--steps-long 1000 --long-batches 64] [--cpu-ref]
    python -m torch.distributed.run --nproc-per-node N ... tools/sweep.py --gpus N --batches 1024,4096 ...

`--batches` are GLOBAL batches; under torchrun each point is sharded over the ranks (dist.shard_range), the weights /
seed are broadcast once and the motions are all-gathered inside the timed region.  Per point: one warm-up call (builds
the step graph), then `--calls` timed p_sample_loop calls with CFG 2.5 and in-kernel Philox noise, CUDA-event timed,
barrier on both sides, max over ranks.  `--cpu-ref` adds, per (arch, frames), the UNMODIFIED reference on this box's
host cores (oracle/_ref, else the oracle port) at batch min(B, 16) x 2 denoise steps -- the CPU rate per frame does
not depend on the batch beyond that -- as frames/s of a `--steps`-step chain.  Prints one JSON line per point and a
markdown table.
"""
import argparse
import json
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def cpu_reference_rate(cfg, sd, T, D, batch=16, steps=2):
    """frames/s of a D-step guided chain on the host cores (bench.py's CPU arm on a bounded sample)."""
    import bench
    fps, ms_step, kind, desc = bench.cpu_rate(cfg, sd, T, D, True, batch, steps, os.cpu_count() or 1, min_seconds=2.0)
    return fps, kind


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--arch", default="v1", choices=["v1", "v2"])
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--batches", default="1,4,16,64,256,1024,4096")
    ap.add_argument("--frames", default="60,120,200,300,400,600")
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--steps-long", type=int, default=0, help="also time this schedule length (e.g. 1000) ...")
    ap.add_argument("--long-batches", default="64", help="... at these batches")
    ap.add_argument("--calls", type=int, default=2)
    ap.add_argument("--precision", default="fp16")
    ap.add_argument("--cpu-ref", action="store_true")
    ap.add_argument("--max-rows", type=int, default=5200000, help="skip points with more token rows per GPU than this")
    a = ap.parse_args()
    import torch.distributed as dist
    from gesturediffusion_b200 import ClassifierFreeSampleModel, create_gaussian_diffusion
    from gesturediffusion_b200 import dist as gdist
    from gesturediffusion_b200.synthetic import build_model, make_inputs, make_state_dict, model_config
    rank, world = gdist.init_process_group()
    if world != a.gpus:
        raise SystemExit("--gpus %d but WORLD_SIZE=%d" % (a.gpus, world))
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    J = 263 if a.arch == "v1" else 498
    cfg = model_config(arch=a.arch, njoints=J)
    sd = make_state_dict(cfg, seed=0)
    net = build_model(cfg, sd, dev, a.precision)
    gdist.broadcast_engine(net.engine(), src=0)
    seed = gdist.broadcast_seed(1, dev)
    model = ClassifierFreeSampleModel(net)
    long_batches = set(int(v) for v in a.long_batches.split(",")) if a.steps_long else set()
    rows = []
    for T in [int(v) for v in a.frames.split(",")]:
        cpu = None
        if a.cpu_ref and rank == 0:
            cpu = cpu_reference_rate(cfg, sd, T, a.steps)
        for B in [int(v) for v in a.batches.split(",")]:
            lo, hi = gdist.shard_range(B, rank, world)
            b_loc = hi - lo
            if b_loc == 0 or 2 * b_loc * (T + 1) > a.max_rows:
                continue
            for D in [a.steps] + ([a.steps_long] if B in long_batches else []):
                dif = create_gaussian_diffusion(timestep_respacing=[D])
                dif.rng, dif.seed, dif.sample_base = "philox", seed, lo
                _, y = make_inputs(cfg, b_loc, T, seed=1234 + lo, device=dev)
                shape = (b_loc, J, 1, T)
                calls = a.calls if D <= 100 else 1

                def call(skip=0):
                    out = dif.p_sample_loop(model, shape, clip_denoised=False, model_kwargs={"y": y}, skip_timesteps=skip)
                    return gdist.gather_motions(out, B) if world > 1 else out
                out = call(skip=max(D - 10, 0) if D > 100 else 0)          # warm-up: builds the step graph
                torch.cuda.synchronize()
                if world > 1:
                    dist.barrier()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                for _ in range(calls):
                    out = call()
                e1.record()
                torch.cuda.synchronize()
                ms = torch.tensor([e0.elapsed_time(e1) / calls], device=dev)
                if world > 1:
                    dist.barrier()
                    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
                ms = float(ms.item())
                rec = dict(arch=a.arch, batch=B, per_gpu_batch=b_loc, n_gpus=world, frames=T, steps=D, ms_per_call=ms,
                           ms_per_denoise_step=ms / D, frames_per_s=B * T / (ms * 1e-3),
                           finite=bool(torch.isfinite(out).all()))
                if cpu is not None and D == a.steps:
                    rec["cpu_ref_frames_per_s"], rec["cpu_ref_kind"] = cpu
                rows.append(rec)
                if rank == 0:
                    print(json.dumps(rec), flush=True)
                del out, y
            net.engine()._ws.clear()                        # release this shape's workspace before the next one
            torch.cuda.empty_cache()
    if rank == 0:
        print("\n| frames | global batch | GPUs | denoise steps | ms / denoise step | frames/s | CPU reference frames/s | x |")
        print("|---|---|---|---|---|---|---|---|")
        for r in rows:
            c = r.get("cpu_ref_frames_per_s")
            print("| %d | %d | %d | %d | %.3f | %.0f | %s | %s |"
                  % (r["frames"], r["batch"], r["n_gpus"], r["steps"], r["ms_per_denoise_step"], r["frames_per_s"],
                     ("%.0f" % c) if c else "", ("%.0f" % (r["frames_per_s"] / c)) if c else ""))
    if dist.is_initialized():
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
