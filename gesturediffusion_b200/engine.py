"""DenoiserEngine: drives the sm_100a kernels for one MDM / MDM_Old denoiser evaluation.

Data layout in HBM (token-major, sequence-contiguous; the reference's [S,B,d] is never formed):
  a_in   act  [B*T, ld_in]      GEMM operand of InputProcess: x_t permuted (mdm.py:350-352) (+ mfcc, V1)
  X32    fp32 [nb*B*S, d]       residual stream (master copy), S = T + 1, token 0 = conditioning
  Xa     act  [nb*B*S, d]       bf16 copy feeding the tensor-core GEMMs (aliases X32 in fp32 mode)
  QKV    act  [nb*B*S, 3d]      in_proj output, q | k | v
  ATT    act  [nb*B*S, d]       attention output
  Y32    fp32 [nb*B*S, d]       sub-layer output + residual (LayerNorm input)
  Hh     act  [nb*B*S, F]       GELU(linear1)
  XL     fp32 [nb*B*T, d]       project_to_lat output (V2 only)
  OUT    fp32 [nb*B, J, T]      model output in the reference layout [.,J,1,T]
nb = 2 batches the conditional and unconditional branches of classifier-free guidance
(cfg_sampler.py:23-28) as 2B sequences; branch 0 = cond, branch 1 = uncond.
`act` is fp16 or bf16 ("fp16" / "bf16" modes: tcgen05 kind::f16 GEMMs, fp32 accumulate, fp32
residual / LayerNorm / softmax statistics; identical speed, fp16 is the default because its 3 extra
mantissa bits bring the per-step parity from ~4e-3 to ~5e-4 relative L2) or fp32 ("fp32" mode: SIMT
kernels, 1e-5 parity runs).
"""
import math

import os

import torch

from . import kernels as K
from .synthetic import MFCC_DIM

_ALIGN = 256


def _round_up(x, m):
    return (x + m - 1) // m * m


class _Arena:
    """Collects host tensors, then materialises them in ONE device buffer per dtype so that the
    multi-GPU start-up broadcast (dist.py) is a single collective per dtype."""

    def __init__(self):
        self.items = {}

    def add(self, name, t, dtype):
        self.items[name] = (t.detach().to(torch.float64 if t.dtype == torch.float64 else t.dtype), dtype)

    def materialise(self, device):
        views, bufs = {}, {}
        for dtype in (torch.float16, torch.bfloat16, torch.float32):
            names = [n for n, (_, dt) in self.items.items() if dt == dtype]
            esz = torch.empty((), dtype=dtype).element_size()
            offs, total = {}, 0
            for n in names:
                offs[n] = total
                total += _round_up(self.items[n][0].numel() * esz, _ALIGN) // esz
            host = torch.zeros(max(total, 1), dtype=dtype)
            for n in names:
                src = self.items[n][0]
                host[offs[n]: offs[n] + src.numel()] = src.reshape(-1).to(dtype)
            dev = host.to(device)
            bufs[dtype] = dev
            for n in names:
                src = self.items[n][0]
                views[n] = dev[offs[n]: offs[n] + src.numel()].view(src.shape)
        return views, bufs


def _pad_cols(w, ld):
    if w.shape[1] == ld:
        return w
    out = torch.zeros(w.shape[0], ld, dtype=w.dtype)
    out[:, : w.shape[1]] = w
    return out


class DenoiserEngine:
    """Packed weights + workspaces + kernel schedule for one denoiser on one GPU."""

    def __init__(self, state_dict, cfg, device, precision="fp16"):
        if precision not in ("fp16", "bf16", "fp32"):
            raise ValueError("precision must be 'fp16', 'bf16' or 'fp32'")
        device = torch.device(device)
        if device.type != "cuda":
            raise K._cabi.GdError("DenoiserEngine needs a CUDA device (no CPU path)")
        K._cabi.load()
        self.cfg = dict(cfg)
        self.device = device
        self.precision = precision
        self.act_dtype = {"fp16": torch.float16, "bf16": torch.bfloat16, "fp32": torch.float32}[precision]
        self.h16 = precision != "fp32"
        self.arch = cfg["arch"]
        self.d = cfg["latent_dim"]
        self.J = cfg["njoints"]
        self.F = cfg["ff_size"]
        self.L = cfg["num_layers"]
        self.H = cfg["num_heads"]
        self.n_seed = cfg["seed_poses"]
        self.use_text = bool(cfg.get("use_text", False)) and self.arch == "v2"
        self.cl_head = cfg.get("cl_head", 8)
        if self.d % 128 != 0:
            raise ValueError("latent_dim must be a multiple of 128 (LayerNorm/attention tiling)")
        if self.d // self.H not in (32, 64, 128):
            raise ValueError("latent_dim/num_heads must be 32, 64 or 128")
        self._ws = {}
        self.fuse_ln = True            # GEMM + residual + LayerNorm in one kernel where gd_gemm_h16_ln supports d
        # residual stream between the fused kernels as a 16-bit (hi, lo) pair instead of fp32 (GD_SPLIT_RES=0: fp32)
        self.split_residual = os.environ.get("GD_SPLIT_RES", "1") != "0"
        # linear1 + GELU + linear2 + residual + LayerNorm as ONE kernel (gd_ffn_h16_ln_split) where it supports the shape
        # (d = 512, F a multiple of 256): 69 us against 79.5 us for the two kernels it replaces at d = 512 / F = 1024, and
        # H (M x F x 2 bytes per layer) never goes to HBM.  GD_FUSE_FFN=0 selects the two-kernel path.
        self.fuse_ffn = os.environ.get("GD_FUSE_FFN", "1") == "1"
        # ... from this many token rows on: a cluster streams all of W1 and W2 per 128-row tile, so with few tiles (small
        # batches) the two GEMM kernels, which spread N over the SMs, are faster (batch 1 x 196 frames: 0.40 vs 0.50 ms per
        # denoise step; break-even near 25 row tiles, profiles/r01g_sweep.md)
        self.fuse_ffn_min_rows = int(os.environ.get("GD_FUSE_FFN_MIN_ROWS", "4096"))
        # GEMM + residual + LayerNorm in one kernel only from this many token rows on: a 2-CTA cluster owns a whole
        # 128-row tile, so at batch 1 (197 rows = 2 tiles) four CTAs stream the complete weight matrix (16.3 us per launch,
        # half of the batch-1 step) where the plain GEMM spreads N over 8-16 CTAs (7 us) and gd_layernorm adds ~3 us.
        # Measured on the batch-B no-guidance step (ms per denoise step, fused / unfused, profiles/r02r_fuse_ln_small.txt):
        # B = 1: 0.346 / 0.293, 8: 0.387 / 0.333, 16: 0.453 / 0.403, 24 (4728 rows): 0.497 / 0.445, 32 (6304 rows):
        # 0.541 / 0.594, 64: 0.862 / 0.960.
        self.fuse_ln_min_rows = int(os.environ.get("GD_FUSE_LN_MIN_ROWS", "5500"))
        # ... and up to this many rows the library's few-row form of the same call (one 8-CTA cluster per row tile): the
        # library says how many rows fit this GPU in one wave of clusters (gd_gemm_ln_small_max_rows; GD_LN_SMALL_MAX_ROWS
        # overrides it there)
        with torch.cuda.device(device):
            self.fuse_ln_small_max_rows = int(K._cabi.load().gd_gemm_ln_small_max_rows())
        self._pack(state_dict)

    # ------------------------------------------------------------------ weights
    def _pack(self, sd):
        sd = {k: v.detach().cpu() for k, v in sd.items()}
        d, J = self.d, self.J
        act = self.act_dtype
        ar = _Arena()
        f32 = torch.float32

        def lin(name, w, b, wdtype=None):
            ar.add(name + ".w", _pad_cols(w, _round_up(w.shape[1], 8)), wdtype or act)
            ar.add(name + ".b", b, f32)

        pe = sd["sequence_pos_encoder.pe"][:, 0, :].float()
        ar.add("pe", pe, f32)
        # one-off fp32 setup products keep fp32 weights
        ar.add("temb0.w", sd["embed_timestep.time_embed.0.weight"], f32)
        ar.add("temb0.b", sd["embed_timestep.time_embed.0.bias"], f32)
        ar.add("temb2.w", sd["embed_timestep.time_embed.2.weight"], f32)
        ar.add("temb2.b", sd["embed_timestep.time_embed.2.bias"], f32)
        ar.add("seed.w", sd["seed_pose_encoder.seed_embed.weight"], f32)
        ar.add("seed.b", sd["seed_pose_encoder.seed_embed.bias"], f32)
        if self.arch == "v1":
            self.k_in = J + MFCC_DIM
            lin("in", sd["input_process.poseEmbedding.weight"], sd["input_process.poseEmbedding.bias"])
        else:
            self.k_in = J
            # InputProcess and project_to_lat are both affine with nothing in between
            # (mdm.py:144,151-169): fold W_p . W_in once (float64) so one GEMM produces the pose term.
            Wp = sd["project_to_lat.weight"].double()
            Win = sd["input_process.poseEmbedding.weight"].double()
            W_pose, W_aud, W_coa = Wp[:, :d], Wp[:, d: d + MFCC_DIM], Wp[:, d + MFCC_DIM:]
            b_fold = W_pose @ sd["input_process.poseEmbedding.bias"].double() + sd["project_to_lat.bias"].double()
            ar.add("in.w", _pad_cols((W_pose @ Win).float(), _round_up(J, 8)), act)
            ar.add("aud.w", W_aud.float().contiguous(), f32)
            ar.add("aud.b", b_fold.float(), f32)
            ar.add("coa.w", W_coa.float().contiguous(), f32)
            ar.add("inv_freq", sd["rel_pos.inv_freq"].float(), f32)
            if self.use_text:
                ar.add("text.w", sd["embed_text.weight"], f32)
                ar.add("text.b", sd["embed_text.bias"], f32)
        for l in range(self.L):
            p = "seqTransEncoder.layers.%d." % l
            lin("l%d.qkv" % l, sd[p + "self_attn.in_proj_weight"], sd[p + "self_attn.in_proj_bias"])
            lin("l%d.o" % l, sd[p + "self_attn.out_proj.weight"], sd[p + "self_attn.out_proj.bias"])
            lin("l%d.f1" % l, sd[p + "linear1.weight"], sd[p + "linear1.bias"])
            lin("l%d.f2" % l, sd[p + "linear2.weight"], sd[p + "linear2.bias"])
            for nm in ("norm1", "norm2"):
                ar.add("l%d.%s.g" % (l, nm), sd[p + nm + ".weight"], f32)
                ar.add("l%d.%s.b" % (l, nm), sd[p + nm + ".bias"], f32)
        lin("out", sd["output_process.poseFinal.weight"], sd["output_process.poseFinal.bias"])
        self.w, self.arenas = ar.materialise(self.device)
        self._derive_tables()

    def _derive_tables(self):
        """Step-invariant products computed once on the device with the fp32 SIMT GEMM:
        the timestep-embedding table over every pe row (TimestepEmbedder, mdm.py:309-310) and,
        for V2, its image under the coarse block of project_to_lat."""
        w, d = self.w, self.d
        n_t = w["pe"].shape[0]
        hid = torch.empty(n_t, d, device=self.device)
        self.temb_table = torch.empty(n_t, d, device=self.device)
        K.gemm(w["pe"], w["temb0.w"], K.epilogue(bias=w["temb0.b"], act=K.GD_ACT_SILU, out_f32=hid))
        K.gemm(hid, w["temb2.w"], K.epilogue(bias=w["temb2.b"], out_f32=self.temb_table))
        self.pe0 = w["pe"][0].contiguous()
        if self.arch == "v2":
            self.coarse_tab = torch.empty(n_t, d, device=self.device)
            K.gemm(self.temb_table, w["coa.w"], K.epilogue(out_f32=self.coarse_tab))
            self._rope = {}
        torch.cuda.current_stream().synchronize()

    def rope_tables(self, T):
        """cos/sin of cat(pos x inv_freq, pos x inv_freq) for positions 0..T
        (local_attention.py:43-53), fp32 like the reference."""
        if T not in self._rope:
            pos = torch.arange(T + 1, device=self.device).float()
            fr = torch.outer(pos, self.w["inv_freq"])
            fr = torch.cat((fr, fr), dim=-1)
            self._rope[T] = (fr.cos().contiguous(), fr.sin().contiguous())
        return self._rope[T]

    # ------------------------------------------------------------------ workspaces
    def workspace(self, B, T, nb):
        """Persistent buffers for (batch, frames, branches); pointers stay fixed so a captured
        step graph can be replayed."""
        key = (B, T, nb)
        if key in self._ws:
            return self._ws[key]
        if self.arch == "v2" and T % 10 != 0:
            raise ValueError("sequence length must be divisible by window size for local attention "
                             "(T=%d, window=10; model/local_attention.py:100-112)" % T)
        if T + 1 > self.w["pe"].shape[0]:
            raise ValueError("sequence too long for the positional table")
        dev, d, F, J, act = self.device, self.d, self.F, self.J, self.act_dtype
        S = T + 1
        R = nb * B * S

        def buf(*shape, dtype=torch.float32):
            return torch.zeros(*shape, device=dev, dtype=dtype)

        ws = dict(B=B, T=T, S=S, nb=nb, R=R)
        ws["a_in"] = buf(B * T, _round_up(self.k_in, 8), dtype=act)
        ws["X32"] = buf(R, d)
        if self.h16:
            # the residual stream as a 16-bit (hi, lo) pair in one allocation
            ws["Xpair"] = buf(2, R, d, dtype=act)
            ws["Xa"], ws["Xlo"] = ws["Xpair"][0], ws["Xpair"][1]
        else:
            ws["Xa"] = ws["X32"]
        ws["QKV"] = buf(R, 3 * d, dtype=act)
        ws["ATT"] = buf(R, d, dtype=act)
        ws["Y32"] = buf(R, d)
        ws["Hh"] = buf(R, F, dtype=act)
        ws["OUT"] = buf(nb * B, J, T)
        ws["t"] = torch.zeros(B, device=dev, dtype=torch.int64)
        ws["seedemb"] = buf(nb * B, d)
        if self.arch == "v2":
            ws["XL"] = buf(nb * B * T, d)
            ws["AM"] = buf(B * T, d)
            ws["mfcc_rows"] = buf(B * T, 32)
            ws["coarse_seed"] = buf(nb * B, d)
            ws["coarse_out"] = buf(nb * B, d)
        self._ws[key] = ws
        return ws

    def _outs(self, t):
        """Epilogue output kwargs for an `act` buffer."""
        return dict(out_h16=t) if t.dtype != torch.float32 else dict(out_f32=t)

    # ------------------------------------------------------------------ per-call conditioning
    def prepare_cond(self, ws, y, branches):
        """Everything that does not change across the denoising steps of one p_sample_loop call
        (SURVEY 2b K3/K4): seed-pose embedding per branch (zero seed -> bias when unconditional,
        mdm.py:125-127,242-250), the mfcc operand columns (V1) or W_a.mfcc + b term (V2), and
        W_c.seed_emb (V2).  `branches` is a list of bools: uncond flag per batched branch."""
        B, T, nb, d, J = ws["B"], ws["T"], ws["nb"], self.d, self.J
        w = self.w
        assert len(branches) == nb
        seed = y["seed"].to(self.device, torch.float32)
        if tuple(seed.shape) != (B, J, 1, self.n_seed):
            raise ValueError("y['seed'] must be [B, J, 1, seed_poses] = %s, got %s"
                             % ((B, J, 1, self.n_seed), tuple(seed.shape)))
        seed_flat = seed.reshape(B, J * self.n_seed).contiguous()
        mfcc = y["mfcc"].to(self.device, torch.float32)
        if tuple(mfcc.shape) != (B, MFCC_DIM, 1, T):
            raise ValueError("y['mfcc'] must be [B, 26, 1, T]")
        mfcc = mfcc.reshape(B, MFCC_DIM, T).contiguous()
        n_seed_out = w["seed.w"].shape[0]
        col0 = d - n_seed_out                       # text_dim when use_text else 0
        for c, uncond in enumerate(branches):
            dst = ws["seedemb"][c * B:(c + 1) * B]
            if uncond:
                dst[:, col0:] = w["seed.b"]
                if self.use_text:
                    dst[:, :col0] = w["text.b"]
            else:
                K.gemm(seed_flat, w["seed.w"], K.epilogue(bias=w["seed.b"], out_f32=dst[:, col0:]))
                if self.use_text:
                    txt = y["text"].to(self.device, torch.float32).contiguous()
                    K.gemm(txt, w["text.w"], K.epilogue(bias=w["text.b"], out_f32=dst[:, :col0]))
        if self.arch == "v1":
            K.pack_rows(mfcc, ws["a_in"], B, MFCC_DIM, T, col_off=J)
        else:
            K.pack_rows(mfcc, ws["mfcc_rows"], B, MFCC_DIM, T, col_off=0)
            K.gemm(ws["mfcc_rows"], w["aud.w"], K.epilogue(bias=w["aud.b"], out_f32=ws["AM"]), K=MFCC_DIM)
            K.gemm(ws["seedemb"], w["coa.w"], K.epilogue(out_f32=ws["coarse_seed"]))

    def load_x(self, ws, x):
        """x [B,J,1,T] fp32 -> operand rows (InputProcess permute, mdm.py:350-352)."""
        K.pack_rows(x.reshape(ws["B"], self.J, ws["T"]), ws["a_in"], ws["B"], self.J, ws["T"], col_off=0)

    # ------------------------------------------------------------------ one denoiser evaluation
    def forward_step(self, ws):
        """Reads ws['a_in'], ws['t'] (+ prepared conditioning); writes ws['OUT'] [nb*B, J, T].
        No host synchronisation, no allocation: capturable in a CUDA graph."""
        B, T, S, nb, R = ws["B"], ws["T"], ws["S"], ws["nb"], ws["R"]
        d, J, w = self.d, self.J, self.w
        X32, Xa = ws["X32"], ws["Xa"]
        xa_b = Xa if self.h16 else None
        if self.arch == "v1":
            # token 0 = time_embed(pe[t]) + seed_embed + pe[0]   (mdm_old.py:94-99,109-112)
            K.cond_tokens(ws["t"], self.temb_table, ws["seedemb"], self.pe0, X32, xa_b, B, nb, S, d)
            # frames: poseEmbedding(cat(x, mfcc)) + pe[1+t], written to rows 1..T of both branches
            K.gemm(ws["a_in"], w["in.w"],
                   K.epilogue(bias=w["in.b"], rowtab=w["pe"], tab_period=T, tab_offset=1, rowmap=(T, S, 1),
                              dup=nb, dup_row_stride=B * S, out_f32=X32, out_h16=xa_b), K=self.k_in)
        else:
            K.cond_tokens(ws["t"], self.temb_table, ws["seedemb"], None, X32, xa_b, B, nb, S, d,
                          coarse_tab=self.coarse_tab, coarse_seed=ws["coarse_seed"], coarse_out=ws["coarse_out"])
            # project_to_lat(cat(poseEmbedding(x), mfcc, coa)) = (W_p W_in) x + [W_a mfcc + b] + W_c coa
            K.gemm(ws["a_in"], w["in.w"],
                   K.epilogue(rowtab=ws["AM"], tab_period=B * T, tab_offset=0, seqbias=ws["coarse_out"],
                              seq_rows=T, seq_count=B, dup=nb, dup_row_stride=B * T, out_f32=ws["XL"]), K=self.k_in)
            rc, rs = self.rope_tables(T)
            K.local_attention(ws["XL"], rc, rs, X32, xa_b, nb * B, T, d, self.cl_head, 10)
        n_seq = nb * B
        self._encoder_stack(ws, 0, n_seq, S)
        # OutputProcess + output[1:] + permute to [.,J,1,T] (mdm.py:219-222,372-380)
        K.gemm(Xa, w["out.w"], K.epilogue(bias=w["out.b"], out_t=ws["OUT"], t_s=S, t_t=T, t_skip=1), K=d)

    def _encoder_stack(self, ws, seq0, seq1, S):
        """The L post-norm encoder layers (mdm.py:90-96,216) on sequences [seq0, seq1) of the workspace."""
        d, w = self.d, self.w
        r0, r1 = seq0 * S, seq1 * S
        rows = lambda name: ws[name][r0:r1]
        X32, Xa = rows("X32"), rows("Xa")
        xa_b = Xa if self.h16 else None
        QKV, ATT, Hh = rows("QKV"), rows("ATT"), rows("Hh")
        dh = d // self.H
        fused_ln = self.fuse_ln and self.h16 and d in (256, 512) and (r1 - r0) >= self.fuse_ln_min_rows
        # few rows (batch 1-8): gd_gemm_h16_ln runs its 8-CTA-cluster kernel (gemm_ln_small.cu), fp32 residual stream
        small_ln = self.fuse_ln and self.h16 and d == 512 and (r1 - r0) <= self.fuse_ln_small_max_rows
        for l in range(self.L):
            p = "l%d." % l
            K.gemm(Xa, w[p + "qkv.w"], K.epilogue(bias=w[p + "qkv.b"], **self._outs(QKV)), K=d)
            K.attention(QKV, ATT, seq1 - seq0, S, self.H, dh)
            if fused_ln and self.split_residual:
                # same two fused kernels with the fp32 residual stream carried through HBM as a 16-bit pair
                # (Xa = round16(x), Xlo = round16(x - Xa)): 4 instead of 6 bytes per element on the way out, and the
                # hi half is the operand the next GEMM reads anyway.  Layer 0 takes the fp32 X32 of the input stage.
                Xlo, last = rows("Xlo"), l == self.L - 1
                res = dict(res_f32=X32) if l == 0 else dict(res_hi=Xa, res_lo=Xlo)
                K.gemm_ln_split(ATT, w[p + "o.w"], w[p + "o.b"], w[p + "norm1.g"], w[p + "norm1.b"],
                                out_hi=Xa, out_lo=Xlo, K=d, **res)
                if self.fuse_ffn and r1 - r0 >= self.fuse_ffn_min_rows and K.ffn_ln_supported(Xa, d, self.F):
                    K.ffn_ln_split(Xa, Xlo, w[p + "f1.w"], w[p + "f1.b"], w[p + "f2.w"], w[p + "f2.b"],
                                   w[p + "norm2.g"], w[p + "norm2.b"], out_hi=Xa, out_lo=None if last else Xlo)
                    continue
                K.gemm(Xa, w[p + "f1.w"], K.epilogue(bias=w[p + "f1.b"], act=K.GD_ACT_GELU_ERF, **self._outs(Hh)), K=d)
                K.gemm_ln_split(Hh, w[p + "f2.w"], w[p + "f2.b"], w[p + "norm2.g"], w[p + "norm2.b"],
                                out_hi=Xa, out_lo=None if last else Xlo, res_hi=Xa, res_lo=Xlo, K=self.F)
                continue
            if fused_ln or small_ln:
                # out_proj + residual + norm1 and linear2 + residual + norm2, each in ONE kernel (2-CTA cluster
                # LayerNorm epilogue); the residual stream X32 is updated in place
                K.gemm_ln(ATT, w[p + "o.w"], w[p + "o.b"], X32, w[p + "norm1.g"], w[p + "norm1.b"],
                          out_f32=X32, out_h16=xa_b, K=d)
                K.gemm(Xa, w[p + "f1.w"], K.epilogue(bias=w[p + "f1.b"], act=K.GD_ACT_GELU_ERF, **self._outs(Hh)), K=d)
                K.gemm_ln(Hh, w[p + "f2.w"], w[p + "f2.b"], X32, w[p + "norm2.g"], w[p + "norm2.b"],
                          out_f32=X32, out_h16=xa_b, K=self.F)
                continue
            K.gemm(ATT, w[p + "o.w"], K.epilogue(bias=w[p + "o.b"], residual=X32, out_f32=rows("Y32")), K=d)
            K.layernorm(rows("Y32"), w[p + "norm1.g"], w[p + "norm1.b"], out_f32=X32, out_h16=xa_b)
            K.gemm(Xa, w[p + "f1.w"], K.epilogue(bias=w[p + "f1.b"], act=K.GD_ACT_GELU_ERF, **self._outs(Hh)), K=d)
            K.gemm(Hh, w[p + "f2.w"], K.epilogue(bias=w[p + "f2.b"], residual=X32, out_f32=rows("Y32")), K=self.F)
            K.layernorm(rows("Y32"), w[p + "norm2.g"], w[p + "norm2.b"], out_f32=X32, out_h16=xa_b)

    def kernels_per_step(self, rows):
        """Kernel launches of one denoiser evaluation over `rows` token rows (nb * B * S)."""
        per_layer = 5 if (self.fuse_ln and self.h16 and self.d in (256, 512) and rows >= self.fuse_ln_min_rows) else 7
        if self.fuse_ln and self.h16 and self.d == 512 and rows <= self.fuse_ln_small_max_rows:
            per_layer = 5
        if (per_layer == 5 and self.split_residual and self.fuse_ffn and self.d == 512 and self.F >= 512
                and self.F % 256 == 0 and rows >= self.fuse_ffn_min_rows):
            per_layer = 4
        return 2 + (1 if self.arch == "v2" else 0) + per_layer * self.L + 1

    def flops_per_step(self, B, T, nb):
        """Algorithmic flops of one denoiser evaluation (BASELINE.md section 5 formula)."""
        d, F, L, J = self.d, self.F, self.L, self.J
        S = T + 1
        E = S * L * (8 * d * d + 4 * d * F + 4 * S * d)
        if self.arch == "v1":
            return B * (2 * T * self.k_in * d + nb * (E + 2 * T * d * J))
        return B * (2 * T * J * d + 2 * T * d * d + nb * (2 * d * d + 62 * d * T + E + 2 * T * d * J))
