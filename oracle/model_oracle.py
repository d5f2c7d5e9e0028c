"""Oracle (test infrastructure) for the denoiser side of the hot path.

Functional, batch-first restatement -- driven directly by a reference-format
``state_dict`` -- of
  model/mdm_old.py:84-122          MDM_Old.forward  (V1: plain trans_enc)
  model/mdm.py:105-224             MDM.forward      (V2: local attention + RoPE + trans_enc)
  model/mdm.py:277-310             sinusoid table + TimestepEmbedder
  model/mdm.py:340-392             Input/Output projections, SeedPoseEncoder
  model/local_attention.py:36-62   look_around, SinusoidalEmbeddings, rotate_half, apply_rotary_pos_emb
  model/local_attention.py:92-172  LocalAttention.forward (window 10, causal, look_backward 1)
  torch.nn.TransformerEncoderLayer (post-norm, erf GELU, eps 1e-5; call sites mdm.py:90-96,216)
The reference computes in [S,B,d]; this restatement computes in [B,S,d] with explicit
matmul/softmax/layer-norm so that it shares no code path with nn.TransformerEncoder.
Never imported by the product package.
"""
import math

import torch
import torch.nn.functional as F


def _lin(x, sd, prefix):
    return x @ sd[prefix + ".weight"].T + sd[prefix + ".bias"]


def sinusoid_table(d_model, max_len, dtype=torch.float32):
    """mdm.py:282-289 (computed in float32 like the reference buffer)."""
    pos = torch.arange(0, max_len, dtype=torch.float).unsqueeze(1)
    div = torch.exp(torch.arange(0, d_model, 2).float() * (-math.log(10000.0) / d_model))
    pe = torch.zeros(max_len, d_model)
    pe[:, 0::2] = torch.sin(pos * div)
    pe[:, 1::2] = torch.cos(pos * div)
    return pe.to(dtype)


def timestep_embedding(sd, timesteps, d, dtype):
    """mdm.py:309-310: Linear -> SiLU -> Linear over pe[t]."""
    key = "embed_timestep.sequence_pos_encoder.pe"
    pe = sd[key][:, 0, :] if key in sd else sinusoid_table(d, 5000, dtype)
    h = _lin(pe[timesteps].to(dtype), sd, "embed_timestep.time_embed.0")
    h = h * torch.sigmoid(h)
    return _lin(h, sd, "embed_timestep.time_embed.2")


def encoder_layer(x, sd, prefix, n_heads):
    """Post-norm nn.TransformerEncoderLayer in eval mode, batch-first [B,S,d]."""
    B, S, d = x.shape
    dh = d // n_heads
    qkv = x @ sd[prefix + ".self_attn.in_proj_weight"].T + sd[prefix + ".self_attn.in_proj_bias"]
    q, k, v = qkv.split(d, dim=-1)
    q = q.view(B, S, n_heads, dh).transpose(1, 2)
    k = k.view(B, S, n_heads, dh).transpose(1, 2)
    v = v.view(B, S, n_heads, dh).transpose(1, 2)
    att = torch.softmax((q @ k.transpose(-1, -2)) * (1.0 / math.sqrt(dh)), dim=-1) @ v
    att = att.transpose(1, 2).reshape(B, S, d)
    x = F.layer_norm(x + _lin(att, sd, prefix + ".self_attn.out_proj"), (d,),
                     sd[prefix + ".norm1.weight"], sd[prefix + ".norm1.bias"], 1e-5)
    h = _lin(x, sd, prefix + ".linear1")
    h = 0.5 * h * (1.0 + torch.erf(h * (1.0 / math.sqrt(2.0))))
    x = F.layer_norm(x + _lin(h, sd, prefix + ".linear2"), (d,),
                     sd[prefix + ".norm2.weight"], sd[prefix + ".norm2.bias"], 1e-5)
    return x


def encoder_stack(x, sd, n_layers, n_heads, taps=None):
    for l in range(n_layers):
        x = encoder_layer(x, sd, "seqTransEncoder.layers.%d" % l, n_heads)
        if taps is not None:
            taps["layer%d" % l] = x
    return x


def _seed_embedding(sd, y, B, uncond, dtype):
    flat = y["seed"].to(dtype).squeeze(2).reshape(B, -1)
    if uncond:
        flat = torch.zeros_like(flat)
    return _lin(flat, sd, "seed_pose_encoder.seed_embed")


def mdm_v1_forward(sd, cfg, x, timesteps, y, uncond=False, taps=None):
    """mdm_old.py:84-122.  cfg: dict(latent_dim, num_layers, num_heads)."""
    dtype = x.dtype
    B, J, _, T = x.shape
    d = cfg["latent_dim"]
    tok0 = timestep_embedding(sd, timesteps, d, dtype) + _seed_embedding(sd, y, B, uncond, dtype)
    feats = torch.cat((x, y["mfcc"].to(dtype)), dim=1)[:, :, 0, :].transpose(1, 2)   # [B,T,J+26]
    h = _lin(feats, sd, "input_process.poseEmbedding")
    key = "sequence_pos_encoder.pe"
    pe = sd[key][:, 0, :] if key in sd else sinusoid_table(d, 5000, dtype)
    seq = torch.cat((tok0.unsqueeze(1), h), dim=1) + pe[: T + 1].to(dtype)
    if taps is not None:
        taps["enc_in"] = seq
    seq = encoder_stack(seq, sd, cfg["num_layers"], cfg["num_heads"], taps)
    out = _lin(seq[:, 1:], sd, "output_process.poseFinal")                             # [B,T,J]
    return out.transpose(1, 2).unsqueeze(2)


def rope_angles(n, dim_head, dtype):
    """local_attention.py:43-53: freqs = cat(t x inv_freq, t x inv_freq)."""
    inv = 1.0 / (10000 ** (torch.arange(0, dim_head, 2).float() / dim_head))
    ang = torch.outer(torch.arange(n).float(), inv)
    return torch.cat((ang, ang), dim=-1).to(dtype)


def rope_apply(x, ang):
    """local_attention.py:55-62: x*cos + rotate_half(x)*sin, half-split convention."""
    half = x.shape[-1] // 2
    rot = torch.cat((-x[..., half:], x[..., :half]), dim=-1)
    return x * ang.cos() + rot * ang.sin()


def local_window_attention(q, window=10):
    """local_attention.py:92-172 specialised to how mdm.py:186-189 calls it:
    q = k = v [N, T, e], causal, look_backward 1, look_forward 0, mask all-True.
    Query at in-window offset i of window w sees window w-1 (all `window` keys, absent
    for w == 0) and keys 0..i of window w.  Score scale e**-0.5; masked scores are
    -finfo.max before the softmax over 2*window slots (so they contribute exactly 0)."""
    N, T, e = q.shape
    if T % window != 0:
        raise ValueError("sequence length must be divisible by window size for local attention")
    W = T // window
    qw = q.view(N, W, window, e)
    pad = torch.full((N, 1, window, e), -1.0, dtype=q.dtype)            # pad_value = -1
    prev = torch.cat((pad, qw[:, :-1]), dim=1)
    kv = torch.cat((prev, qw), dim=2)                                    # [N,W,2w,e]
    sim = torch.einsum("nwie,nwje->nwij", qw, kv) * (e ** -0.5)
    neg = -torch.finfo(sim.dtype).max
    pos_q = torch.arange(T).view(W, window)
    pos_k = torch.cat((torch.cat((torch.full((1, window), -1), pos_q[:-1]), dim=0), pos_q), dim=1)
    causal = pos_q.unsqueeze(-1) < pos_k.unsqueeze(-2)                   # [W,w,2w]
    sim = sim.masked_fill(causal.unsqueeze(0), neg)
    padded = torch.zeros(W, 2 * window, dtype=torch.bool)
    padded[0, :window] = True                                            # look-back of window 0
    sim = sim.masked_fill(padded.view(1, W, 1, 2 * window), neg)
    att = sim.softmax(dim=-1)
    return torch.einsum("nwij,nwje->nwie", att, kv).reshape(N, T, e)


def mdm_v2_forward(sd, cfg, x, timesteps, y, uncond=False, taps=None):
    """mdm.py:105-224.  cfg: dict(latent_dim, num_layers, num_heads, use_text, cl_head=8)."""
    dtype = x.dtype
    B, J, _, T = x.shape
    d = cfg["latent_dim"]
    H = cfg.get("cl_head", 8)
    e = d // H
    seed_emb = _seed_embedding(sd, y, B, uncond, dtype)
    if cfg.get("use_text", False):
        txt = y["text"].to(dtype)                       # synthetic CLIP-like embedding [B, clip_dim]
        if uncond:
            txt = torch.zeros_like(txt)
        seed_emb = torch.cat((_lin(txt, sd, "embed_text"), seed_emb), dim=1)
    coa = seed_emb + timestep_embedding(sd, timesteps, d, dtype)                       # [B,d]
    pose = _lin(x[:, :, 0, :].transpose(1, 2), sd, "input_process.poseEmbedding")      # [B,T,d]
    audio = y["mfcc"].to(dtype)[:, :, 0, :].transpose(1, 2)                            # [B,T,26]
    feats = torch.cat((pose, audio, coa.unsqueeze(1).expand(B, T, d)), dim=-1)
    z = _lin(feats, sd, "project_to_lat")                                              # [B,T,d]
    if taps is not None:
        taps["proj"] = z
    zh = z.view(B, T, H, e).transpose(1, 2).reshape(B * H, T, e)
    zh = rope_apply(zh, rope_angles(T, e, dtype))
    zh = local_window_attention(zh, 10)
    z = zh.view(B, H, T, e).transpose(1, 2).reshape(B, T, d)
    if taps is not None:
        taps["local"] = z
    seq = torch.cat((coa.unsqueeze(1), z), dim=1)                                      # [B,T+1,d]
    sh = seq.view(B, T + 1, H, e).transpose(1, 2)
    sh = rope_apply(sh, rope_angles(T + 1, e, dtype))
    seq = sh.transpose(1, 2).reshape(B, T + 1, d)
    if taps is not None:
        taps["enc_in"] = seq
    seq = encoder_stack(seq, sd, cfg["num_layers"], cfg["num_heads"], taps)
    out = _lin(seq[:, 1:], sd, "output_process.poseFinal")
    return out.transpose(1, 2).unsqueeze(2)


def make_model_fn(sd, cfg, dtype=torch.float32):
    """Returns model_fn(x, t, y, uncond=False).  cfg['arch'] in {'v1','v2'}."""
    sdc = {k: (v.to(dtype) if v.is_floating_point() else v) for k, v in sd.items()}
    fwd = mdm_v1_forward if cfg["arch"] == "v1" else mdm_v2_forward

    def model_fn(x, t, y, uncond=False):
        if isinstance(y, dict) and y.get("uncond", False):
            uncond = True
        with torch.no_grad():
            return fwd(sdc, cfg, x.to(dtype), t, y, uncond=uncond)
    return model_fn
