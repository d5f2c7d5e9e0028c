#!/usr/bin/env python
"""bench.py -- throughput of the B200-native denoising sampling loop on BASELINE.json's metric.

    python bench.py --gpus N --steps K --warmup W            (N > 1: launched under torch.distributed.run)
    python bench.py --impl reference --gpus N --steps K --warmup W

Metric: motion frames/sec of classifier-free-guided DDPM sampling (and ms per denoise step).
One bench "step" = one full `p_sample_loop` call (D denoise steps) over one batch of synthetic
conditioning.  Headline workload at every N (kept from round 1 so rounds stay comparable): BASELINE.json
configs[1] per GPU -- MDM trans_enc (`MDM_Old`) d512 / 8 layers / 4 heads, text-to-motion shape 263 x 196,
batch 64 per GPU, CFG 2.5, random-init weights, synthetic conditioning; D = 50 respaced steps (the metric's
"50-step CFG DDPM").  Weak scaling: per-GPU batch fixed, global batch = 64 * N, weights/seed broadcast once,
motions all-gathered at the end of every call, no per-step collective.

Extra keys of the same JSON line (each a full measurement of another BASELINE config, `--skip-extras` drops them):
  cfg4      configs[3]: KIT 251 x 196, CFG, GLOBAL batch 1024 split over the N ranks (strong scaling), with the
            start-up broadcast and the final gather timed separately
  cfg3      configs[2]: the gesture model (V2: local attention + RoPE + encoder), 498 x 120, batch 128 per GPU
  cfg2b     configs[1] as quoted: V2 with a 512-d text embedding, 263 x 200, batch 64, CFG, 1000-step DDPM
  ddpm1000  the headline workload on the full 1000-step schedule;  ddim50 / take4: SURVEY 8.6 rows

The JSON line carries: value (device-resident inputs), e2e (host pinned inputs -> H2D, every rank's shard of the
result -> D2H inside the timed region), roofline (dominant kernel, CUDA-event timed in this process), cpu_baseline
(the UNMODIFIED reference on this box's cores when oracle/_ref is staged, else the oracle port; bounded sample),
clocks (nvidia-smi during the timed region), gpu_launches.
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
    "cfg2b": ("v2", 263, 200, 64, True),       # BASELINE configs[1] as quoted: V2 + 512-d text embedding (T 196 -> 200)
    "cfg3": ("v2", 498, 120, 128, False),      # BASELINE configs[2]: gesture model with local attention
    "cfg4": ("v1", 251, 196, 128, False),      # BASELINE configs[3] per-GPU shard at 8 GPUs (global batch 1024)
    "cfg1": ("v1", 263, 196, 1, False),        # BASELINE configs[0] shape (B=1)
}
CFG4_GLOBAL_BATCH = 1024
GUIDANCE = 2.5
PRECISION_NOTE = ("16-bit mode = fp16 operands / fp32 accumulate (north_star names bf16): bf16's 8-bit mantissa gives "
                  "3-4.5e-3 per-step rel-L2 by operand rounding alone (tools/sim_16bit_rounding.py), above the 2e-3 "
                  "bound; fp16 measures 5-8e-4 at identical speed (same tcgen05 kind::f16 path); --precision bf16 runs "
                  "the bf16 variant")


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
    ap.add_argument("--frames", type=int, default=0, help="override the number of frames")
    ap.add_argument("--no-cfg", action="store_true")
    ap.add_argument("--skip-cpu-baseline", action="store_true")
    ap.add_argument("--skip-extras", "--skip-1000", dest="skip_extras", action="store_true",
                    help="headline workload only (no cfg4 / cfg3 / cfg2b / ddpm1000 / ddim / take legs)")
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


def build_case(workload, batch=0, frames=0):
    from gesturediffusion_b200.synthetic import make_state_dict, model_config
    arch, J, T, B, use_text = WORKLOADS[workload]
    if batch:
        B = batch
    if frames:
        T = frames
    cfg = model_config(arch=arch, njoints=J, use_text=use_text)
    sd = make_state_dict(cfg, seed=0)
    return cfg, sd, B, T


def workload_name(workload, cfg, B, T, D, use_cfg):
    return ("%s: %s d%d L%d H%d%s, %dx%d, batch %d/GPU, %s, %d-step respaced DDPM (cosine, START_X, FIXED_SMALL)"
            % (workload, "MDM_Old (V1 trans_enc)" if cfg["arch"] == "v1" else "MDM (V2 local-attn+RoPE+trans_enc)",
               cfg["latent_dim"], cfg["num_layers"], cfg["num_heads"], " +text" if cfg.get("use_text") else "",
               cfg["njoints"], T, B, "CFG %.1f" % GUIDANCE if use_cfg else "no CFG", D))


# ------------------------------------------------------------------------------------------ CPU arm
def cpu_rate(cfg, sd, T, D, use_cfg, sample_b, sample_steps, threads, min_seconds=12.0):
    """Times the reference's CPU implementation of the path on a bounded sample: the UNMODIFIED reference
    (`SpacedDiffusion.p_sample_loop` over `MDM` / `MDM_Old` under `ClassifierFreeSampleModel`, fp32, CPU) when it
    is available -- /root/reference in the build container, the byte-identical archive oracle/_ref/ that
    oracle/stage_ref.py staged on the GPU box -- (kind 'reference'), else the oracle port (kind 'port').
    Returns (frames/s for a D-step chain, ms per denoise step, kind, description).  The sample runs the LAST
    `sample_steps` of the D-step schedule (every step costs the same); frames/s = batch*frames / (ms * D)."""
    from gesturediffusion_b200.synthetic import make_inputs, make_step_noise
    torch.set_num_threads(threads)
    x, y = make_inputs(cfg, sample_b, T, guidance=GUIDANCE)
    kind = "port"
    try:
        from oracle import ref_shims
        if ref_shims.reference_available():
            kind = "reference"
    except Exception:                                   # noqa: BLE001 -- any import problem -> the port
        kind = "port"
    if kind == "reference":
        ns = ref_shims.load_reference()
        net = ref_shims.build_reference_model(cfg["arch"], cfg["njoints"], latent_dim=cfg["latent_dim"],
                                              num_layers=cfg["num_layers"], num_heads=cfg["num_heads"],
                                              ff_size=cfg["ff_size"], seed_poses=cfg["seed_poses"],
                                              use_text=cfg["use_text"])
        missing, unexpected = net.load_state_dict(sd, strict=False)
        assert not missing and not unexpected
        model = ns.ClassifierFreeSampleModel(net) if use_cfg else net
        dif = ref_shims.build_reference_diffusion([D])

        def run(n_steps):
            with torch.no_grad():
                dif.p_sample_loop(model, tuple(x.shape), noise=x, clip_denoised=False, model_kwargs={"y": dict(y)},
                                  skip_timesteps=D - n_steps, init_image=None, progress=False, dump_steps=None,
                                  const_noise=False)
        what = "unmodified reference (%s, torch CPU fp32, %d threads)" % (ref_shims.reference_kind(), threads)
    else:
        from oracle.diffusion_oracle import OracleDiffusion, oracle_cfg
        from oracle.model_oracle import make_model_fn
        fn = make_model_fn(sd, cfg)
        model = oracle_cfg(fn) if use_cfg else fn
        od = OracleDiffusion(respacing=[D])
        noises = make_step_noise(tuple(x.shape), sample_steps)

        def run(n_steps):
            od.p_sample_loop(model, x, y, noises[:n_steps], skip_timesteps=D - n_steps)
        what = "oracle port (torch CPU fp32, %d threads)" % threads
    run(1)                                                                     # warm-up: one step
    reps, dt = 0, 0.0
    while dt < min_seconds or reps == 0:                                       # bounded sample: >= min_seconds of CPU work
        t0 = time.perf_counter()
        run(sample_steps)
        dt += time.perf_counter() - t0
        reps += 1
    ms_step = dt / (reps * sample_steps) * 1e3
    fps = sample_b * T / (ms_step * 1e-3 * D)
    desc = ("%s, batch %d x %d denoise steps x %d repeats timed (%.1f s), %s; "
            "frames/s = batch*frames / (ms_per_denoise_step * %d steps)"
            % (what, sample_b, sample_steps, reps, dt, "CFG" if use_cfg else "no CFG", D))
    return fps, ms_step, kind, desc


def cpu_sample_plan(B, D):
    """Bounded CPU sample: the bench batch itself (up to 64) x a few of the D steps."""
    if B <= 1:
        return 1, D
    return min(B, 64), 2


def run_reference(args):
    """--impl reference: the reference's own CPU implementation of the path on this box's host cores."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cfg, sd, B, T = build_case(args.workload, args.batch, args.frames)
    D = args.denoise_steps
    use_cfg = not args.no_cfg
    threads = os.cpu_count() or 1
    sample_b, sample_steps = cpu_sample_plan(B, D)
    vals, ms = [], []
    kind, desc = "port", ""
    for i in range(args.warmup + args.steps):
        fps, ms_step, kind, desc = cpu_rate(cfg, sd, T, D, use_cfg, sample_b, sample_steps, threads, min_seconds=6.0)
        if i >= args.warmup:
            vals.append(fps)
            ms.append(ms_step)
    v = sum(vals) / len(vals)
    line = {"metric": "motion frames/sec (CFG DDPM sampling)", "value": v, "unit": "frames/s", "impl": "reference",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": sum(ms) / len(ms) * D, "ms_per_denoise_step": sum(ms) / len(ms),
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": workload_name(args.workload, cfg, B, T, D, use_cfg), "denoise_steps": D,
                       "parallelism": "cpu"},
            "cpu_baseline": {"value": v, "unit": "frames/s", "cores": threads, "kind": kind, "sample": desc},
            "e2e": {"value": v, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))


# ------------------------------------------------------------------------------------------ B200 arm
class Case:
    """One workload on this rank: model + engine (weights broadcast from rank 0), diffusion schedule, synthetic
    conditioning as pinned host tensors (e2e) and device tensors (value), and the timed-call helpers."""

    def __init__(self, workload, B, T, D, precision, use_cfg, rank, world, dev, sample_base, seed, cfg_sd=None):
        import torch.distributed as dist
        from gesturediffusion_b200 import ClassifierFreeSampleModel, create_gaussian_diffusion
        from gesturediffusion_b200 import dist as gdist
        from gesturediffusion_b200.synthetic import build_model, make_inputs
        self.dist, self.gdist = dist, gdist
        self.workload, self.B, self.T, self.D = workload, B, T, D
        self.rank, self.world, self.dev, self.use_cfg = rank, world, dev, use_cfg
        self.cfg, self.sd = cfg_sd if cfg_sd else build_case(workload)[:2]
        self.J = self.cfg["njoints"]
        self.denoiser = build_model(self.cfg, self.sd, dev, precision)
        self.eng = self.denoiser.engine()
        # start-up collectives, timed once: rank 0's packed weight arenas (one broadcast per dtype) + the noise seed
        self._sync()
        t0 = time.perf_counter()
        gdist.broadcast_engine(self.eng, src=0)
        self.seed = gdist.broadcast_seed(seed, dev)
        self._sync()
        self.broadcast_ms = (time.perf_counter() - t0) * 1e3
        self.model = ClassifierFreeSampleModel(self.denoiser) if use_cfg else self.denoiser
        self.make_dif = lambda resp: self._dif(create_gaussian_diffusion(timestep_respacing=resp), sample_base)
        self.dif = self.make_dif([D])
        self.shape = (B, self.J, 1, T)
        self.global_batch = None                      # set by the caller (weak: B * world, strong: fixed)
        _, y_host = make_inputs(self.cfg, B, T, seed=1234 + sample_base, guidance=GUIDANCE)
        self.y_pin = {k: v.pin_memory() for k, v in y_host.items()}
        self.y_dev = {k: v.to(dev) for k, v in y_host.items()}
        self.out_pin = torch.empty(B, self.J, 1, T).pin_memory()         # this rank's shard of the motions
        self.h2d = sum(v.numel() * v.element_size() for v in self.y_pin.values())
        self.d2h = self.out_pin.numel() * 4

    def _dif(self, dif, sample_base):
        dif.rng, dif.seed, dif.sample_base = "philox", self.seed, sample_base
        return dif

    def _sync(self):
        torch.cuda.synchronize()
        if self.world > 1:
            self.dist.barrier()

    def one_call(self, e2e, diffusion=None, skip=0):
        """value path: conditioning resident in HBM, motions all-gathered on the device (every rank holds the batch).
        e2e path: conditioning H2D from pinned memory, and every rank copies ITS OWN shard of the motions to its pinned
        host buffer (one PCIe link per GPU in parallel; no rank moves another rank's shard to the host)."""
        dif = diffusion or self.dif
        y = {k: v.to(self.dev, non_blocking=True) for k, v in self.y_pin.items()} if e2e else self.y_dev
        out = dif.p_sample_loop(self.model, self.shape, clip_denoised=False, model_kwargs={"y": y}, skip_timesteps=skip)
        if e2e:
            self.out_pin.copy_(out, non_blocking=True)
            torch.cuda.current_stream().synchronize()
        elif self.world > 1:
            out = self.gdist.gather_motions(out, self.global_batch)
        return out

    def timed(self, n, e2e, diffusion=None):
        """CUDA events around n calls, barrier + synchronize on both sides, max over ranks (ms)."""
        self._sync()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(n):
            self.one_call(e2e, diffusion)
        e1.record()
        torch.cuda.synchronize()
        ms = torch.tensor([e0.elapsed_time(e1)], device=self.dev)
        if self.world > 1:
            self.dist.barrier()
            self.dist.all_reduce(ms, op=self.dist.ReduceOp.MAX)
        return float(ms.item())

    def timed_gather(self, reps=5):
        """The final all-gather alone (ms per call, max over ranks)."""
        if self.world == 1:
            return 0.0
        x = torch.zeros(self.shape, device=self.dev)
        self.gdist.gather_motions(x, self.global_batch)
        self._sync()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            self.gdist.gather_motions(x, self.global_batch)
        e1.record()
        torch.cuda.synchronize()
        ms = torch.tensor([e0.elapsed_time(e1) / reps], device=self.dev)
        self.dist.all_reduce(ms, op=self.dist.ReduceOp.MAX)
        return float(ms.item())

    def free(self):
        self.eng._ws.clear()
        self.denoiser = self.model = self.eng = self.dif = None
        torch.cuda.empty_cache()


def extra_case(workload, B, T, D, args, rank, world, dev, sample_base, global_batch, steps, peaks, respacing=None):
    """One more BASELINE config measured like the headline (value + e2e + step roofline); returns its JSON object."""
    cfg, sd, _, _ = build_case(workload)
    c = Case(workload, B, T, D, args.precision, True, rank, world, dev, sample_base, 20261104, cfg_sd=(cfg, sd))
    c.global_batch = global_batch
    dif = c.make_dif(respacing) if respacing is not None else c.dif
    n_den = dif.num_timesteps
    if n_den > 100:
        c.one_call(False, dif, skip=n_den - 10)               # builds the step graph of this schedule (10 steps)
    else:
        for _ in range(2):
            c.one_call(False, dif)
    c.one_call(True, dif, skip=max(n_den - 10, 0))
    ms = c.timed(steps, False, dif) / steps
    ms_e2e = c.timed(steps, True, dif) / steps
    frames = global_batch * T
    flops = c.eng.flops_per_step(B, T, 2)
    tf = flops / (ms / n_den * 1e-3) / 1e12
    obj = {"workload": workload_name(workload, cfg, B, T, n_den, True), "global_batch": global_batch,
           "per_gpu_batch": B, "value": frames / (ms * 1e-3), "unit": "frames/s", "ms_per_step": ms,
           "ms_per_denoise_step": ms / n_den, "calls_timed": steps,
           "e2e": {"value": frames / (ms_e2e * 1e-3), "unit": "frames/s", "h2d_bytes_per_step": c.h2d * world,
                   "d2h_bytes_per_step": c.d2h * world, "ms_per_step": ms_e2e},
           "broadcast_ms": c.broadcast_ms, "gather_ms": c.timed_gather(),
           "step_roofline": {"bound": "tensor", "achieved": tf, "peak": peaks["tflops_sustained"], "unit": "TFLOP/s",
                             "frac": tf / peaks["tflops_sustained"],
                             "algorithmic_gflop_per_denoise_step_per_gpu": flops / 1e9}}
    c.free()
    return obj


def dominant_kernel_roofline(c, peaks, profile_out):
    """CUDA events around every launch of six eager denoise steps of the headline workload: per-kernel breakdown and the
    roofline of the kernel with the largest share of the step (rank 0)."""
    from gesturediffusion_b200 import kernels as K
    D = c.D
    dif_p = c.make_dif([D])
    dif_p.use_cuda_graph = False
    K.PROFILE = []
    # Keep the GPU busy (a ~30 ms spin kernel) while the host enqueues the six eager steps and their events: otherwise the
    # GPU outruns the host, every [start, end] event pair also spans the host's launch latency of the kernel between
    # them, and the per-launch figure depends on the box's CPU (69.7 vs 83.0 us for the same kernel on two boxes).
    torch.cuda.synchronize()
    torch.cuda._sleep(int(45e6))
    dif_p.p_sample_loop(c.model, c.shape, clip_denoised=False, model_kwargs={"y": c.y_dev}, skip_timesteps=D - 6)
    summ = K.profile_summary(K.PROFILE)
    K.PROFILE = None
    tot = sum(t for _, t in summ.values())
    breakdown = sorted(((n, dict(m), cnt, t) for (n, m), (cnt, t) in summ.items()), key=lambda r: -r[3])
    # launches of one kernel at one (M, N, K) are one entry, whatever their residual / output byte variant (rw)
    merged = {}
    for n, m, cnt, t in breakdown:
        if n not in ("gemm_h16", "gemm_ln_h16", "ffn_ln_h16"):
            continue
        # the fused feed-forward kernel is two chained GEMMs: N = d, K = F stands for both (flops 4 M d F)
        key = (n, m["M"], m["N"], m["F"] if n == "ffn_ln_h16" else m["K"])
        c0_, t0_, b0_ = merged.get(key, (0, 0.0, 0.0))
        merged[key] = (c0_ + cnt, t0_ + t, b0_ + cnt * float(m.get("rw", 10)))
    gemm = sorted(((n, {"M": M_, "N": N_, "K": K_, "rw": b / cnt}, cnt, t) for (n, M_, N_, K_), (cnt, t, b) in merged.items()),
                  key=lambda r: -r[3])
    roof = None
    if gemm:
        # algorithmic work per launch: 2MNK flops; bytes = A (16 bit) + outputs (+ the residual pair for
        # the GEMM+LayerNorm kernels); the weights stay L2-resident and are not counted
        def algo(n, m):
            if n == "ffn_ln_h16":
                # linear1 + linear2: 4 M d F flops; bytes = the residual pair in (it is also the A operand) and
                # the pair (or its hi half) out, m["rw"] per element; H never leaves the SM
                return 4.0 * m["M"] * m["N"] * m["K"], m["M"] * m["N"] * float(m["rw"])
            fl = 2.0 * m["M"] * m["N"] * m["K"]
            by = m["M"] * m["K"] * 2.0 + (m["M"] * m["N"] * float(m.get("rw", 10)) if n == "gemm_ln_h16"
                                          else m["M"] * m["N"] * 2.0)
            return fl, by
        fl_all = sum(algo(n, m)[0] * cnt for n, m, cnt, t in gemm)
        tt = sum(t for n, m, cnt, t in gemm)
        n0, m0, c0, t0 = gemm[0]                           # the kernel with the largest share of the step
        fl0, by0 = algo(n0, m0)
        sec = t0 / c0 * 1e-3
        hbm_bound = by0 / (peaks["hbm_gbs"] * 1e9) > fl0 / (peaks["tflops_sustained"] * 1e12)
        if n0 == "ffn_ln_h16":
            label = "ffn_ln_tcgen05_kernel M=%d d=%d F=%d" % (m0["M"], m0["N"], m0["K"])
        else:
            label = "%s M=%d N=%d K=%d" % ("gemm_ln_tcgen05_kernel" if n0 == "gemm_ln_h16"
                                           else "gemm_h16_tcgen05_kernel", m0["M"], m0["N"], m0["K"])
        # DRAM and L2->SM bytes per launch of this kernel from the committed ncu capture of the kernel INSIDE a full
        # step replay (tools/ncu_traffic.py regenerates profiles/traffic.json); None when no capture matches the label
        traffic, l2_to_sm = None, None
        tpath = os.path.join(ROOT, "profiles", "traffic.json")
        if os.path.exists(tpath):
            ent = json.load(open(tpath)).get(label, {})
            traffic, l2_to_sm = ent.get("dram_bytes_per_launch"), ent.get("l2_to_sm_bytes_per_launch")
        if hbm_bound:
            ach, peak, unit = by0 / sec / 1e9, peaks["hbm_gbs"], "GB/s"
        else:
            ach, peak, unit = fl0 / sec / 1e12, peaks["tflops_sustained"], "TFLOP/s"
        roof = {"bound": "hbm" if hbm_bound else "tensor", "kernel": label, "achieved": ach, "peak": peak,
                "unit": unit, "frac": ach / peak, "traffic": traffic, "traffic_l2_to_sm": l2_to_sm,
                "traffic_over_algorithmic": (traffic / by0) if traffic else None,
                "l2_to_sm_over_algorithmic": (l2_to_sm / by0) if l2_to_sm else None,
                "peak_source": peaks["source"] + ", sustained figure (kernel timed inside a long step)",
                "launch_ms": t0 / c0, "share_of_step": t0 / tot, "_launches_per_step": c0 / 6.0,
                "algorithmic_bytes_per_launch": by0,
                "algorithmic_flops_per_launch": fl0,
                "tensor_tflops_same_kernel": fl0 / sec / 1e12,
                "all_gemms": {"achieved": fl_all / (tt * 1e-3) / 1e12, "unit": "TFLOP/s",
                              "frac_of_tensor_peak": fl_all / (tt * 1e-3) / 1e12 / peaks["tflops_sustained"],
                              "share_of_step": tt / tot}}
    if profile_out:
        with open(profile_out, "w") as fh:
            json.dump({"workload": workload_name(c.workload, c.cfg, c.B, c.T, D, c.use_cfg), "eager_denoise_steps": 6,
                       "total_ms": tot, "kernels": [{"name": n, "meta": m, "launches": cnt, "total_ms": t,
                                                     "share": t / tot} for n, m, cnt, t in breakdown]}, fh, indent=1)
    return roof


def run_b200(args):
    from gesturediffusion_b200 import _cabi
    from gesturediffusion_b200 import dist as gdist
    from gesturediffusion_b200 import sampler as S

    rank, world = gdist.init_process_group()
    if world != args.gpus:
        raise SystemExit("--gpus %d but WORLD_SIZE=%d (launch N>1 under torch.distributed.run)" % (args.gpus, world))
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    _cabi.load()                                     # fail loudly if the CUDA library is missing

    cfg, sd, B, T = build_case(args.workload, args.batch, args.frames)
    D = args.denoise_steps
    use_cfg = not args.no_cfg
    peaks = measured_peaks()
    # model: rank 0's weights are the job's weights (broadcast of the packed arenas), seed broadcast
    c = Case(args.workload, B, T, D, args.precision, use_cfg, rank, world, dev, rank * B, 20261104, cfg_sd=(cfg, sd))
    c.global_batch = B * world
    eng, dif = c.eng, c.dif

    for _ in range(max(args.warmup, 3)):
        c.one_call(False)
    c.one_call(True)

    sampler = ClockSampler(local) if rank == 0 else None
    if sampler:
        sampler.start()
    l0 = _cabi.launch_count() + S.STATS["graph_kernels"]
    ms_total = c.timed(args.steps, False)
    launches = _cabi.launch_count() + S.STATS["graph_kernels"] - l0
    ms_e2e = c.timed(args.steps, True)
    clocks = sampler.stop() if sampler else None

    frames = B * world * T * args.steps
    value = frames / (ms_total * 1e-3)
    ms_step = ms_total / args.steps
    e2e_value = frames / (ms_e2e * 1e-3)
    gather_ms = c.timed_gather()

    extra = {}
    if not args.skip_extras:
        dif1000 = c.make_dif("")
        c.one_call(False, dif1000, skip=990)                   # builds this schedule's step graph (10 steps)
        ms1000 = c.timed(1, False, dif1000)
        extra["ddpm1000"] = {"value": B * world * T / (ms1000 * 1e-3), "unit": "frames/s",
                             "ms_per_denoise_step": ms1000 / 1000.0, "calls": 1}

        # SURVEY 8.6 rows on the same workload, rank-local (no gather): 50-step DDIM (eta 0) and, for the 6-interleaved
        # GENEA layout, a 4-chunk take through generate_take (seed hand-off + de-normalise/split on the device)
        def timed_local(fn):
            fn()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            fn()
            e1.record()
            torch.cuda.synchronize()
            return e0.elapsed_time(e1)
        dif_ddim = c.make_dif("ddim%d" % D)
        ms_ddim = timed_local(lambda: dif_ddim.ddim_sample_loop(c.model, c.shape, clip_denoised=False,
                                                                model_kwargs={"y": c.y_dev}, eta=0.0))
        extra["ddim%d" % D] = {"value": B * T / (ms_ddim * 1e-3), "unit": "frames/s per GPU",
                               "ms_per_denoise_step": ms_ddim / D}
        if c.J % 6 == 0 and use_cfg:
            from gesturediffusion_b200.generate import generate_take
            chunks = [c.y_dev] * 4
            mean, std = torch.zeros(c.J, device=dev), torch.ones(c.J, device=dev)
            ms_take = timed_local(lambda: generate_take(c.model, dif, chunks, B, T, seed_poses=cfg["seed_poses"],
                                                        guidance_param=GUIDANCE, mean=mean, std=std))
            extra["take4"] = {"value": 4 * B * T / (ms_take * 1e-3), "unit": "frames/s per GPU", "chunks": 4,
                              "ms_per_denoise_step": ms_take / (4 * D)}

    # ---- roofline of the dominant kernel: CUDA events around every launch of an eager denoise pass (rank 0)
    roof = dominant_kernel_roofline(c, peaks, args.profile_out) if rank == 0 else None
    flops_step = eng.flops_per_step(B, T, 2 if use_cfg else 1)
    weights_mb = sum(a.numel() * a.element_size() for a in eng.arenas.values()) // 2 ** 20
    broadcast_ms, h2d, d2h = c.broadcast_ms, c.h2d, c.d2h

    if not args.skip_extras and args.workload == "cfg2a":
        c.free()
        # configs[3]: KIT shape, GLOBAL batch 1024 sharded over the ranks (strong scaling; N = 1 holds all 1024)
        lo, hi = gdist.shard_range(CFG4_GLOBAL_BATCH, rank, world)
        extra["cfg4"] = extra_case("cfg4", hi - lo, 196, D, args, rank, world, dev, lo, CFG4_GLOBAL_BATCH, 2, peaks)
        extra["cfg4"]["scaling"] = "strong (global batch %d / %d ranks)" % (CFG4_GLOBAL_BATCH, world)
        # configs[2]: the reference's native gesture config, batch 128 per GPU (weak)
        extra["cfg3"] = extra_case("cfg3", 128, 120, D, args, rank, world, dev, rank * 128, 128 * world, 3, peaks)
        # configs[1] as quoted: V2 + text conditioning, T = 200, CFG, the full 1000-step schedule
        extra["cfg2b"] = extra_case("cfg2b", 64, 200, D, args, rank, world, dev, rank * 64, 64 * world, 1, peaks,
                                    respacing="")

    if rank != 0:
        return
    step_tflops = flops_step / (ms_step / D * 1e-3) / 1e12
    if roof is not None:
        # The eager pass that carries the per-launch events cannot overlap a kernel's prologue with its predecessor (no
        # programmatic dependent launch across an event record), so `launch_ms` is a few us above what the kernel costs
        # inside the step graph.  Extra, clearly labelled estimate: the kernel's share of the eager pass x the graph's
        # time per denoise step / its launches per step (`frac` itself stays the directly measured figure).
        per_step = roof.pop("_launches_per_step", None)
        if per_step:
            est_ms = roof["share_of_step"] * (ms_step / D) / per_step
            scale = roof["launch_ms"] / est_ms
            roof["in_graph_estimate"] = {"launch_ms": est_ms, "achieved": roof["achieved"] * scale,
                                         "frac": roof["frac"] * scale,
                                         "how": "share_of_step x graph ms per denoise step / launches per step"}
    line = {"metric": "motion frames/sec (CFG DDPM sampling)", "value": value, "unit": "frames/s",
            "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms_step,
            "ms_per_denoise_step": ms_step / D, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": {"fp16": "f16 operands / f32 accumulate", "bf16": "bf16 operands / f32 accumulate",
                      "fp32": "f32"}[args.precision],
            "precision_note": PRECISION_NOTE if args.precision == "fp16" else None,
            "data": "synthetic (random-init weights of the reference architecture, N(0,1) seed poses / MFCC)",
            "config": {"workload": workload_name(args.workload, cfg, B, T, D, use_cfg), "global_batch": B * world,
                       "frames": T, "denoise_steps": D, "guidance": GUIDANCE if use_cfg else None,
                       "parallelism": "dp%d (batch-sharded, no per-step collective)" % world,
                       "noise": "in-kernel Philox4x32-10 keyed by (seed, global sample, step)",
                       "l2": "per-denoise-step working set ~%d MB of activations > 126 MB L2; weights (%d MB) stay "
                             "L2-resident as in production" % (2 * B * (T + 1) * 12 * cfg["latent_dim"] * 2 // 2 ** 20,
                                                               weights_mb),
                       "e2e_path": "H2D of y from pinned memory, p_sample_loop, every rank copies its own shard of the "
                                   "motions D2H to pinned memory (no device gather on the e2e path); value path: "
                                   "device-resident y, all-gather of the motions"},
            "e2e": {"value": e2e_value, "unit": "frames/s", "h2d_bytes_per_step": h2d * world,
                    "d2h_bytes_per_step": d2h * world, "ms_per_step": ms_e2e / args.steps},
            "gpu_launches": int(launches), "broadcast_ms": broadcast_ms, "gather_ms": gather_ms,
            "step_roofline": {"bound": "tensor", "achieved": step_tflops, "peak": peaks["tflops_sustained"],
                              "unit": "TFLOP/s", "frac": step_tflops / peaks["tflops_sustained"],
                              "algorithmic_gflop_per_denoise_step": flops_step / 1e9},
            "roofline": roof, "clocks": clocks}
    line.update(extra)
    if world == 1 and not args.skip_cpu_baseline:
        threads = os.cpu_count() or 1
        sb, ss = cpu_sample_plan(B, D)
        fps, ms_c, kind, desc = cpu_rate(cfg, sd, T, D, use_cfg, sb, ss, threads)
        line["cpu_baseline"] = {"value": fps, "unit": "frames/s", "cores": threads, "kind": kind, "sample": desc,
                                "ms_per_denoise_step_at_sample_batch": ms_c, "same_batch": sb == B}
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
