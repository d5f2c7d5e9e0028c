#!/usr/bin/env python
"""Short run of the bench workload for `ncu`: a few denoise steps of one p_sample_loop call.

    python tools/profile_step.py --workload cfg2a --denoise-steps 3 [--graph]

Same model, shapes, precision and kernels as bench.py's timed region, but only the LAST `--denoise-steps`
steps of the 50-step chain (skip_timesteps), eager launches by default so that every kernel shows up as its
own row in ncu's launch list.  Prints one JSON line with the wall time; numbers printed under ncu are not
bench values.
"""
import argparse
import json
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import bench  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="cfg2a", choices=sorted(bench.WORKLOADS))
    ap.add_argument("--denoise-steps", type=int, default=3)
    ap.add_argument("--chain", type=int, default=50)
    ap.add_argument("--precision", default="fp16")
    ap.add_argument("--batch", type=int, default=0)
    ap.add_argument("--graph", action="store_true")
    ap.add_argument("--no-cfg", action="store_true")
    a = ap.parse_args()
    from gesturediffusion_b200 import ClassifierFreeSampleModel, _cabi, create_gaussian_diffusion
    from gesturediffusion_b200.synthetic import build_model, make_inputs
    _cabi.load()
    dev = torch.device("cuda", 0)
    cfg, sd, B, T = bench.build_case(a.workload, a.batch)
    den = build_model(cfg, sd, dev, a.precision)
    model = den if a.no_cfg else ClassifierFreeSampleModel(den)
    dif = create_gaussian_diffusion(timestep_respacing=[a.chain])
    dif.rng, dif.seed, dif.use_cuda_graph = "philox", 1, bool(a.graph)
    _, y = make_inputs(cfg, B, T, guidance=bench.GUIDANCE, device=dev)
    shape = (B, cfg["njoints"], 1, T)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    out = dif.p_sample_loop(model, shape, clip_denoised=False, model_kwargs={"y": y},
                            skip_timesteps=a.chain - a.denoise_steps)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    print(json.dumps({"workload": a.workload, "B": B, "T": T, "denoise_steps": a.denoise_steps,
                      "graph": bool(a.graph), "wall_s": dt, "launches": _cabi.launch_count(),
                      "finite": bool(torch.isfinite(out).all())}))


if __name__ == "__main__":
    main()
