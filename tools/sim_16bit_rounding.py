"""CPU simulation of where 16-bit operand rounding enters the denoiser (test/analysis tool, not product):
rounds GEMM operands / attention probabilities to bf16 or fp16 at the points the engine does and reports the
relative L2 against the reference golden output.  Evidence for DESIGN.md section "16-bit format"."""
import math, sys, torch
sys.path.insert(0, "/root/repo/tests"); sys.path.insert(0, "/root/repo")
import torch.nn.functional as F
from helpers import load_model_case, rel_l2
import oracle.model_oracle as mo

def run(case, dt, round_w=True, round_a=True, round_attn=True):
    cfg, sd, x, y, gold = load_model_case(case)
    r = (lambda t: t.to(dt).float()) 
    ra = r if round_a else (lambda t: t)
    rw = r if round_w else (lambda t: t)
    orig_lin = mo._lin
    def lin(xx, sd_, prefix):
        big = prefix.startswith("seqTransEncoder") or prefix.startswith("input_process") or prefix.startswith("output_process") or prefix.startswith("project_to_lat")
        if big:
            return ra(xx) @ rw(sd_[prefix + ".weight"]).T + sd_[prefix + ".bias"]
        return orig_lin(xx, sd_, prefix)
    def enc(xx, sd_, prefix, n_heads):
        B, S, d = xx.shape; dh = d // n_heads
        qkv = ra(xx) @ rw(sd_[prefix + ".self_attn.in_proj_weight"]).T + sd_[prefix + ".self_attn.in_proj_bias"]
        if round_attn: qkv = r(qkv)
        q, k, v = qkv.split(d, dim=-1)
        q = q.view(B, S, n_heads, dh).transpose(1, 2); k = k.view(B, S, n_heads, dh).transpose(1, 2); v = v.view(B, S, n_heads, dh).transpose(1, 2)
        p = torch.softmax((q @ k.transpose(-1, -2)) / math.sqrt(dh), dim=-1)
        if round_attn: p = r(p)
        att = (p @ v).transpose(1, 2).reshape(B, S, d)
        xx = F.layer_norm(xx + lin(att, sd_, prefix + ".self_attn.out_proj"), (d,), sd_[prefix + ".norm1.weight"], sd_[prefix + ".norm1.bias"], 1e-5)
        h = lin(xx, sd_, prefix + ".linear1"); h = 0.5 * h * (1.0 + torch.erf(h / math.sqrt(2.0)))
        xx = F.layer_norm(xx + lin(h, sd_, prefix + ".linear2"), (d,), sd_[prefix + ".norm2.weight"], sd_[prefix + ".norm2.bias"], 1e-5)
        return xx
    mo._lin, old_enc = lin, mo.encoder_layer
    mo.encoder_layer = enc
    try:
        fn = mo.make_model_fn(sd, cfg)
        c = fn(x, gold["t"], y); u = fn(x, gold["t"], y, uncond=True)
        g = u + y["scale"].view(-1,1,1,1) * (c - u)
    finally:
        mo._lin, mo.encoder_layer = orig_lin, old_enc
    return rel_l2(c, gold["cond"]), rel_l2(g, gold["guided"])

for case in ["v1_humanml", "v2_gesture"]:
    for name, dt in [("bf16", torch.bfloat16), ("fp16", torch.float16)]:
        print(case, name, "all: cond %.2e guided %.2e" % run(case, dt))
    print(case, "bf16 weights exact:  cond %.2e guided %.2e" % run(case, torch.bfloat16, round_w=False))
    print(case, "bf16 acts exact:     cond %.2e guided %.2e" % run(case, torch.bfloat16, round_a=False, round_attn=False))
    print(case, "bf16 no attn round:  cond %.2e guided %.2e" % run(case, torch.bfloat16, round_attn=False))
