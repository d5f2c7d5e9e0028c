#!/usr/bin/env python
"""bench.py -- throughput of the B200-native denoising sampling loop on BASELINE.json's metric.

    python bench.py --gpus N --steps K --warmup W            (N > 1: launched under torch.distributed.run)
    python bench.py --impl reference --gpus N --steps K --warmup W

Metric: motion frames/sec of classifier-free-guided DDPM sampling (and ms per denoise step).
One bench "step" = one full `p_sample_loop` call (D denoise steps) over one batch of synthetic
conditioning.  Workload at every N: BASELINE.json configs[1] per GPU -- MDM trans_enc (`MDM_Old`) d512 /
8 layers / 4 heads, text-to-motion shape 263 x 196, batch 64 per GPU, CFG 2.5, random-init weights,
synthetic conditioning; D = 50 respaced steps (the metric's "50-step CFG DDPM"); one extra
1000-step call is timed and reported under "ddpm1000" (configs[1] quotes 1000 steps), a 50-step DDIM call under
"ddim50" and, for the GENEA feature layout (--workload cfg3), a 4-chunk take under "take4".
Weak scaling: per-GPU batch fixed, global batch = 64 * N, weights/seed broadcast once, motions
all-gathered at the end of every call, no per-step collective.

The JSON line carries: value (device-resident inputs), e2e (host pinned inputs -> H2D, result -> D2H inside
the timed region), roofline (dominant kernel, CUDA-event timed in this process), cpu_baseline (the CPU oracle
port on this box's cores, bounded sample), clocks (nvidia-smi during the timed region), gpu_launches.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    # name: (arch, njoints, frames, batch per GPU, use_text)
    "cfg2a": ("v1", 263, 196, 64, False),      # BASELINE configs[1] on the reference's runnable V1 path
    "cfg3": ("v2", 498, 120, 128, False),      # BASELINE configs[2]: gesture model with local attention
    "cfg4": ("v1", 251, 196, 128, False),      # BASELINE configs[3] per-GPU shard at 8 GPUs
    "cfg1": ("v1", 263, 196, 1, False),        # BASELINE configs[0] shape (B=1)
}
GUIDANCE = 2.5


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="cfg2a", choices=sorted(WORKLOADS))
    ap.add_argument("--denoise-steps", type=int, default=50)
    ap.add_argument("--precision", default="fp16", choices=["fp16", "bf16", "fp32"])
    ap.add_argument("--batch", type=int, default=0, help="override per-GPU batch")
    ap.add_argument("--no-cfg", action="store_true")
    ap.add_argument("--skip-cpu-baseline", action="store_true")
    ap.add_argument("--skip-1000", action="store_true")
    ap.add_argument("--profile-out", default="", help="write the per-kernel CUDA-event breakdown to this JSON file")
    return ap.parse_args()


# ------------------------------------------------------------------------------------------ helpers
class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms during the timed region."""
    FIELDS = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.gpu_index = gpu_index
        self.rows = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "--query-gpu=" + self.FIELDS, "--format=csv,noheader,nounits",
                                          "-lms", "200", "-i", str(self.gpu_index)],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            parts = [p.strip() for p in line.split(",")]
            if len(parts) >= 8:
                self.rows.append(parts)

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm = sorted(float(r[1]) for r in self.rows if r[1].replace(".", "").isdigit())
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in self.rows for i in range(4) if r[4 + i].lower().startswith("active")})
        pw = [float(r[3]) for r in self.rows if r[3].replace(".", "").isdigit()]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None,
                "sm_max_mhz": float(self.rows[0][2]) if self.rows else None,
                "power_w_max": max(pw) if pw else None, "samples": len(self.rows), "reasons": reasons}


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        p = json.load(open(path))
        return dict(tflops_burst=p["bf16_tflops"], tflops_sustained=p["bf16_tflops_sustained"], hbm_gbs=p["hbm_gbs"],
                    source="measured (MEASURED_PEAKS.json)")
    return dict(tflops_burst=1590.0, tflops_sustained=1400.0, hbm_gbs=6650.0, source="fallback (B200_PROFILING.md)")


def build_case(args, device=None):
    from gesturediffusion_b200.synthetic import make_inputs, make_state_dict, model_config
    arch, J, T, B, use_text = WORKLOADS[args.workload]
    if args.batch:
        B = args.batch
    cfg = model_config(arch=arch, njoints=J, use_text=use_text)
    sd = make_state_dict(cfg, seed=0)
    return cfg, sd, B, T


def workload_name(args, cfg, B, T, D, use_cfg):
    return ("%s: %s d%d L%d H%d, %dx%d, batch %d/GPU, %s, %d-step respaced DDPM (cosine, START_X, FIXED_SMALL)"
            % (args.workload, "MDM_Old (V1 trans_enc)" if cfg["arch"] == "v1" else "MDM (V2 local-attn+RoPE+trans_enc)",
               cfg["latent_dim"], cfg["num_layers"], cfg["num_heads"], cfg["njoints"], T, B,
               "CFG %.1f" % GUIDANCE if use_cfg else "no CFG", D))


# ------------------------------------------------------------------------------------------ CPU oracle timing
def cpu_oracle_rate(cfg, sd, T, D, use_cfg, sample_b, sample_steps, threads, min_seconds=12.0):
    """Times the CPU oracle port (oracle/: the reference's algorithm restated on torch CPU ops, pinned to the
    reference by tests/golden) on a bounded sample; returns (frames/s for a D-step chain, ms/denoise-step,
    description).  Linear extrapolation in batch and steps is stated in the description."""
    from gesturediffusion_b200.synthetic import make_inputs, make_step_noise
    from oracle.diffusion_oracle import OracleDiffusion, oracle_cfg
    from oracle.model_oracle import make_model_fn
    torch.set_num_threads(threads)
    fn = make_model_fn(sd, cfg)
    model = oracle_cfg(fn) if use_cfg else fn
    od = OracleDiffusion(respacing=[D])
    x, y = make_inputs(cfg, sample_b, T, guidance=GUIDANCE)
    noises = make_step_noise(tuple(x.shape), sample_steps)
    od.p_sample_loop(model, x, y, noises[:1], skip_timesteps=D - 1)           # warm-up: one step
    reps, dt = 0, 0.0
    while dt < min_seconds or reps == 0:                                       # bounded sample: >= min_seconds of CPU work
        t0 = time.perf_counter()
        od.p_sample_loop(model, x, y, noises, skip_timesteps=D - sample_steps)
        dt += time.perf_counter() - t0
        reps += 1
    ms_step = dt / (reps * sample_steps) * 1e3
    fps = sample_b * T / (ms_step * 1e-3 * D)
    desc = ("oracle port (torch CPU fp32, %d threads), batch %d x %d denoise steps x %d repeats timed (%.1f s), %s; "
            "frames/s = batch*frames / (ms_per_denoise_step * %d steps)"
            % (threads, sample_b, sample_steps, reps, dt, "CFG" if use_cfg else "no CFG", D))
    return fps, ms_step, desc


def run_reference(args):
    """--impl reference: the reference's CPU implementation of the path on this box's host cores.  The
    reference is Python and does not travel to the GPU box, so this is the oracle port (kind 'port')."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cfg, sd, B, T = build_case(args)
    D = args.denoise_steps
    use_cfg = not args.no_cfg
    threads = os.cpu_count() or 1
    sample_b, sample_steps = (min(B, 16), 10) if B > 1 else (1, D)
    vals, ms = [], []
    for i in range(args.warmup + args.steps):
        fps, ms_step, desc = cpu_oracle_rate(cfg, sd, T, D, use_cfg, sample_b, sample_steps, threads)
        if i >= args.warmup:
            vals.append(fps)
            ms.append(ms_step)
    v = sum(vals) / len(vals)
    line = {"metric": "motion frames/sec (CFG DDPM sampling)", "value": v, "unit": "frames/s", "impl": "reference",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": sum(ms) / len(ms) * D, "ms_per_denoise_step": sum(ms) / len(ms),
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": workload_name(args, cfg, B, T, D, use_cfg), "denoise_steps": D,
                       "parallelism": "cpu"},
            "cpu_baseline": {"value": v, "unit": "frames/s", "cores": threads, "kind": "port", "sample": desc},
            "e2e": {"value": v, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))


# ------------------------------------------------------------------------------------------ B200 arm
def run_b200(args):
    import torch.distributed as dist
    from gesturediffusion_b200 import ClassifierFreeSampleModel, _cabi, create_gaussian_diffusion
    from gesturediffusion_b200 import dist as gdist
    from gesturediffusion_b200 import kernels as K
    from gesturediffusion_b200 import sampler as S
    from gesturediffusion_b200.synthetic import make_inputs

    rank, world = gdist.init_process_group()
    if world != args.gpus:
        raise SystemExit("--gpus %d but WORLD_SIZE=%d (launch N>1 under torch.distributed.run)" % (args.gpus, world))
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    _cabi.load()                                     # fail loudly if the CUDA library is missing

    cfg, sd, B, T = build_case(args)
    D = args.denoise_steps
    use_cfg = not args.no_cfg
    J = cfg["njoints"]

    # model: rank 0's weights are the job's weights (broadcast of the packed arenas), seed broadcast
    from gesturediffusion_b200.synthetic import build_model
    denoiser = build_model(cfg, sd, dev, args.precision)
    eng = denoiser.engine()
    gdist.broadcast_engine(eng, src=0)
    model = ClassifierFreeSampleModel(denoiser) if use_cfg else denoiser
    dif = create_gaussian_diffusion(timestep_respacing=[D])
    dif.rng = "philox"
    dif.seed = gdist.broadcast_seed(20261104, dev)
    dif.sample_base = rank * B
    shape = (B, J, 1, T)

    # synthetic conditioning: pinned host copies (e2e) and device-resident copies (value)
    _, y_host = make_inputs(cfg, B, T, seed=1234 + rank, guidance=GUIDANCE)
    y_pin = {k: v.pin_memory() for k, v in y_host.items()}
    y_dev = {k: v.to(dev) for k, v in y_host.items()}
    out_pin = torch.empty(B * world if world > 1 else B, J, 1, T).pin_memory()
    h2d = sum(v.numel() * v.element_size() for v in y_pin.values())
    d2h = out_pin.numel() * 4

    def one_call(e2e, diffusion=dif, skip=0):
        if e2e:
            y = {k: v.to(dev, non_blocking=True) for k, v in y_pin.items()}
        else:
            y = y_dev
        out = diffusion.p_sample_loop(model, shape, clip_denoised=False, model_kwargs={"y": y}, skip_timesteps=skip)
        if world > 1:
            out = gdist.gather_motions(out, B * world)
        if e2e:
            out_pin.copy_(out, non_blocking=True)
            torch.cuda.current_stream().synchronize()
        return out

    def timed(n, e2e, diffusion=dif):
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(n):
            one_call(e2e, diffusion)
        e1.record()
        torch.cuda.synchronize()
        ms = torch.tensor([e0.elapsed_time(e1)], device=dev)
        if world > 1:
            dist.barrier()
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item())

    for _ in range(max(args.warmup, 3)):
        one_call(False)
    one_call(True)

    sampler = ClockSampler(local) if rank == 0 else None
    if sampler:
        sampler.start()
    l0 = _cabi.launch_count() + S.STATS["graph_kernels"]
    ms_total = timed(args.steps, False)
    launches = _cabi.launch_count() + S.STATS["graph_kernels"] - l0
    ms_e2e = timed(args.steps, True)
    clocks = sampler.stop() if sampler else None

    frames = B * world * T * args.steps
    value = frames / (ms_total * 1e-3)
    ms_step = ms_total / args.steps
    e2e_value = frames / (ms_e2e * 1e-3)

    extra = {}
    if not args.skip_1000:
        dif1000 = create_gaussian_diffusion(timestep_respacing="")
        dif1000.rng, dif1000.seed, dif1000.sample_base = "philox", dif.seed, dif.sample_base
        one_call(False, dif1000, skip=990)                     # builds this schedule's step graph (10 steps)
        ms1000 = timed(1, False, dif1000)
        extra["ddpm1000"] = {"value": B * world * T / (ms1000 * 1e-3), "unit": "frames/s",
                             "ms_per_denoise_step": ms1000 / 1000.0, "calls": 1}

    # SURVEY 8.6 rows on the same workload, rank-local (no gather): 50-step DDIM (eta 0) and, for the 6-interleaved
    # GENEA layout (cfg3), a 4-chunk take through generate_take (seed hand-off + de-normalise/split on the device)
    if not args.skip_1000:
        def timed_local(fn):
            fn()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            fn()
            e1.record()
            torch.cuda.synchronize()
            return e0.elapsed_time(e1)
        dif_ddim = create_gaussian_diffusion(timestep_respacing="ddim%d" % D)
        dif_ddim.rng, dif_ddim.seed, dif_ddim.sample_base = "philox", dif.seed, dif.sample_base
        ms_ddim = timed_local(lambda: dif_ddim.ddim_sample_loop(model, shape, clip_denoised=False,
                                                                model_kwargs={"y": y_dev}, eta=0.0))
        extra["ddim%d" % D] = {"value": B * T / (ms_ddim * 1e-3), "unit": "frames/s per GPU",
                               "ms_per_denoise_step": ms_ddim / D}
        if J % 6 == 0 and use_cfg:
            from gesturediffusion_b200.generate import generate_take
            chunks = [y_dev] * 4
            mean, std = torch.zeros(J, device=dev), torch.ones(J, device=dev)
            ms_take = timed_local(lambda: generate_take(model, dif, chunks, B, T, seed_poses=cfg["seed_poses"],
                                                        guidance_param=GUIDANCE, mean=mean, std=std))
            extra["take4"] = {"value": 4 * B * T / (ms_take * 1e-3), "unit": "frames/s per GPU", "chunks": 4,
                              "ms_per_denoise_step": ms_take / (4 * D)}

    # ---- roofline of the dominant kernel: CUDA events around every launch of one eager denoise pass
    peaks = measured_peaks()
    roof, breakdown = None, None
    if rank == 0:
        dif.use_cuda_graph = False
        dif_p = create_gaussian_diffusion(timestep_respacing=[D])
        dif_p.rng, dif_p.seed, dif_p.use_cuda_graph = "philox", dif.seed, False
        K.PROFILE = []
        dif_p.p_sample_loop(model, shape, clip_denoised=False, model_kwargs={"y": y_dev}, skip_timesteps=D - 6)
        summ = K.profile_summary(K.PROFILE)
        K.PROFILE = None
        dif.use_cuda_graph = True
        tot = sum(t for _, t in summ.values())
        breakdown = sorted(((n, dict(m), c, t) for (n, m), (c, t) in summ.items()), key=lambda r: -r[3])
        # launches of one kernel at one (M, N, K) are one entry, whatever their residual / output byte variant (rw)
        merged = {}
        for n, m, c, t in breakdown:
            if n not in ("gemm_h16", "gemm_ln_h16", "ffn_ln_h16"):
                continue
            # the fused feed-forward kernel is two chained GEMMs: N = d, K = F stands for both (flops 4 M d F)
            key = (n, m["M"], m["N"], m["F"] if n == "ffn_ln_h16" else m["K"])
            c0_, t0_, b0_ = merged.get(key, (0, 0.0, 0.0))
            merged[key] = (c0_ + c, t0_ + t, b0_ + c * float(m.get("rw", 10)))
        gemm = sorted(((n, {"M": M_, "N": N_, "K": K_, "rw": b / c}, c, t) for (n, M_, N_, K_), (c, t, b) in merged.items()),
                      key=lambda r: -r[3])
        if gemm:
            # algorithmic work per launch: 2MNK flops; bytes = A (16 bit) + outputs (+ fp32 residual for
            # the GEMM+LayerNorm kernel); the weights stay L2-resident and are not counted
            def algo(n, m):
                if n == "ffn_ln_h16":
                    # linear1 + linear2: 4 M d F flops; bytes = the residual pair in (it is also the A operand) and
                    # the pair (or its hi half) out, m["rw"] per element; H never leaves the SM
                    return 4.0 * m["M"] * m["N"] * m["K"], m["M"] * m["N"] * float(m["rw"])
                fl = 2.0 * m["M"] * m["N"] * m["K"]
                # GEMM+LN: fp32 residual in, fp32 + 16-bit out = 10 B per element; with the residual stream carried
                # as a 16-bit (hi, lo) pair the wrapper records its own figure (m["rw"]: 4 in + 4 or 2 out)
                by = m["M"] * m["K"] * 2.0 + (m["M"] * m["N"] * float(m.get("rw", 10)) if n == "gemm_ln_h16"
                                              else m["M"] * m["N"] * 2.0)
                return fl, by
            fl_all = sum(algo(n, m)[0] * c for n, m, c, t in gemm)
            tt = sum(t for n, m, c, t in gemm)
            n0, m0, c0, t0 = gemm[0]                           # the kernel with the largest share of the step
            fl0, by0 = algo(n0, m0)
            sec = t0 / c0 * 1e-3
            hbm_bound = by0 / (peaks["hbm_gbs"] * 1e9) > fl0 / (peaks["tflops_sustained"] * 1e12)
            if n0 == "ffn_ln_h16":
                label = "ffn_ln_tcgen05_kernel M=%d d=%d F=%d" % (m0["M"], m0["N"], m0["K"])
            else:
                label = "%s M=%d N=%d K=%d" % ("gemm_ln_tcgen05_kernel" if n0 == "gemm_ln_h16"
                                               else "gemm_h16_tcgen05_kernel", m0["M"], m0["N"], m0["K"])
            traffic = None
            tpath = os.path.join(ROOT, "profiles", "traffic.json")
            if os.path.exists(tpath):
                traffic = json.load(open(tpath)).get(label, {}).get("dram_bytes_per_launch")
            if hbm_bound:
                ach, peak, unit = by0 / sec / 1e9, peaks["hbm_gbs"], "GB/s"
            else:
                ach, peak, unit = fl0 / sec / 1e12, peaks["tflops_sustained"], "TFLOP/s"
            roof = {"bound": "hbm" if hbm_bound else "tensor", "kernel": label, "achieved": ach, "peak": peak,
                    "unit": unit, "frac": ach / peak, "traffic": traffic, "peak_source": peaks["source"],
                    "launch_ms": t0 / c0, "share_of_step": t0 / tot, "algorithmic_bytes_per_launch": by0,
                    "algorithmic_flops_per_launch": fl0,
                    "tensor_tflops_same_kernel": fl0 / sec / 1e12,
                    "all_gemms": {"achieved": fl_all / (tt * 1e-3) / 1e12, "unit": "TFLOP/s",
                                  "frac_of_tensor_peak": fl_all / (tt * 1e-3) / 1e12 / peaks["tflops_sustained"],
                                  "share_of_step": tt / tot}}
        if args.profile_out:
            with open(args.profile_out, "w") as fh:
                json.dump({"workload": workload_name(args, cfg, B, T, D, use_cfg), "eager_denoise_steps": 6,
                           "total_ms": tot, "kernels": [{"name": n, "meta": m, "launches": c, "total_ms": t,
                                                         "share": t / tot} for n, m, c, t in breakdown]}, fh, indent=1)

    if rank != 0:
        return
    flops_step = eng.flops_per_step(B, T, 2 if use_cfg else 1)
    step_tflops = flops_step / (ms_step / D * 1e-3) / 1e12
    line = {"metric": "motion frames/sec (CFG DDPM sampling)", "value": value, "unit": "frames/s",
            "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms_step,
            "ms_per_denoise_step": ms_step / D, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": {"fp16": "f16 operands / f32 accumulate", "bf16": "bf16 operands / f32 accumulate",
                      "fp32": "f32"}[args.precision],
            "data": "synthetic (random-init weights of the reference architecture, N(0,1) seed poses / MFCC)",
            "config": {"workload": workload_name(args, cfg, B, T, D, use_cfg), "global_batch": B * world,
                       "frames": T, "denoise_steps": D, "guidance": GUIDANCE if use_cfg else None,
                       "parallelism": "dp%d (batch-sharded, no per-step collective)" % world,
                       "noise": "in-kernel Philox4x32-10 keyed by (seed, global sample, step)",
                       "l2": "per-denoise-step working set ~%d MB of activations > 126 MB L2; weights (%d MB) stay "
                             "L2-resident as in production" % (2 * B * (T + 1) * 12 * cfg["latent_dim"] * 2 // 2 ** 20,
                                                               sum(a.numel() * a.element_size() for a in eng.arenas.values()) // 2 ** 20)},
            "e2e": {"value": e2e_value, "unit": "frames/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "ms_per_step": ms_e2e / args.steps},
            "gpu_launches": int(launches),
            "step_roofline": {"bound": "tensor", "achieved": step_tflops, "peak": peaks["tflops_sustained"],
                              "unit": "TFLOP/s", "frac": step_tflops / peaks["tflops_sustained"],
                              "algorithmic_gflop_per_denoise_step": flops_step / 1e9},
            "roofline": roof, "clocks": clocks}
    line.update(extra)
    if world == 1 and not args.skip_cpu_baseline:
        threads = os.cpu_count() or 1
        sb, ss = (min(B, 16), 10) if B > 1 else (1, D)
        fps, ms_c, desc = cpu_oracle_rate(cfg, sd, T, D, use_cfg, sb, ss, threads)
        line["cpu_baseline"] = {"value": fps, "unit": "frames/s", "cores": threads, "kind": "port", "sample": desc,
                                "ms_per_denoise_step_at_sample_batch": ms_c}
    print(json.dumps(line))


if __name__ == "__main__":
    a = parse_args()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_b200(a)
        import torch.distributed as _dist
        if _dist.is_available() and _dist.is_initialized():
            _dist.destroy_process_group()
