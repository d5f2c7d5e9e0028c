#!/usr/bin/env python
"""Throughput sweep of the sampling loop over batch and sequence length (BASELINE.json configs[4]):

    python tools/sweep.py [--arch v1] [--batches 1,8,64,256,1024] [--frames 60,196,600] [--steps 50]

Per point: one warm-up call (builds the step graph), then `--calls` timed p_sample_loop calls with CFG 2.5 and
in-kernel Philox noise, CUDA-event timed.  Prints a markdown table and one JSON line per point.
"""
import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--arch", default="v1", choices=["v1", "v2"])
    ap.add_argument("--batches", default="1,8,64,256,1024")
    ap.add_argument("--frames", default="60,196,600")
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--calls", type=int, default=2)
    ap.add_argument("--precision", default="fp16")
    ap.add_argument("--max-rows", type=int, default=1400000, help="skip points with more than this many token rows")
    a = ap.parse_args()
    from gesturediffusion_b200 import ClassifierFreeSampleModel, create_gaussian_diffusion
    from gesturediffusion_b200.synthetic import build_model, make_inputs, make_state_dict, model_config
    dev = torch.device("cuda", 0)
    J = 263 if a.arch == "v1" else 498
    cfg = model_config(arch=a.arch, njoints=J)
    model = ClassifierFreeSampleModel(build_model(cfg, make_state_dict(cfg, seed=0), dev, a.precision))
    rows = []
    for T in [int(v) for v in a.frames.split(",")]:
        for B in [int(v) for v in a.batches.split(",")]:
            if 2 * B * (T + 1) > a.max_rows:
                continue
            dif = create_gaussian_diffusion(timestep_respacing=[a.steps])
            dif.rng, dif.seed = "philox", 1
            _, y = make_inputs(cfg, B, T, device=dev)
            shape = (B, J, 1, T)
            out = dif.p_sample_loop(model, shape, clip_denoised=False, model_kwargs={"y": y})
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(a.calls):
                out = dif.p_sample_loop(model, shape, clip_denoised=False, model_kwargs={"y": y})
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / a.calls
            rec = dict(arch=a.arch, batch=B, frames=T, steps=a.steps, ms_per_call=ms, ms_per_denoise_step=ms / a.steps,
                       frames_per_s=B * T / (ms * 1e-3), finite=bool(torch.isfinite(out).all()))
            rows.append(rec)
            print(json.dumps(rec), flush=True)
            model.model.engine()._ws.clear()            # release this shape's workspace before the next one
            del out, y
            torch.cuda.empty_cache()
    print("\n| frames | batch | ms / denoise step | frames/s |\n|---|---|---|---|")
    for r in rows:
        print("| %d | %d | %.3f | %.0f |" % (r["frames"], r["batch"], r["ms_per_denoise_step"], r["frames_per_s"]))


if __name__ == "__main__":
    main()
