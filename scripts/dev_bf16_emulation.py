"""CPU experiment: restatement with bf16 rounding injected at the engine's storage points, toggled per class of
point, to attribute the gradient error (profiles/parity_r1_status.md).  Not product code, not a test."""
import math, sys, itertools, statistics, torch
import torch.nn.functional as F
sys.path.insert(0, ".")
from oracle import restatement as R

FLAGS = dict(w=True, act_branch=True, act_stream=True, grad_branch=True, grad_stream=True)
DT = dict(fwd=torch.bfloat16, grad=torch.bfloat16)     # storage dtype of forward tensors / gradient tensors

class _Rnd(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, f, b):
        ctx.b = b
        return x.to(DT["fwd"]).float() if f else x
    @staticmethod
    def backward(ctx, g):
        return (g.to(DT["grad"]).float() if ctx.b else g), None, None

def rb(x):  # branch tensor
    return _Rnd.apply(x, FLAGS["act_branch"], FLAGS["grad_branch"])
def rs(x):  # residual-stream tensor
    return _Rnd.apply(x, FLAGS["act_stream"], FLAGS["grad_stream"])
def W(sd, k):
    w = sd[k]
    return _Rnd.apply(w, FLAGS["w"], False) if w.dim() >= 3 else w

def res_block(sd, pre, x, emb):
    h1 = rb(F.silu(R.group_norm32(x, sd[pre + "in_layers.0.weight"], sd[pre + "in_layers.0.bias"])))
    emb_out = F.linear(F.silu(emb), sd[pre + "emb_layers.1.weight"], sd[pre + "emb_layers.1.bias"])
    h2 = rb(F.conv2d(h1, W(sd, pre + "in_layers.2.weight"), sd[pre + "in_layers.2.bias"], padding=1) + emb_out[:, :, None, None])
    h3 = rb(F.silu(R.group_norm32(h2, sd[pre + "out_layers.0.weight"], sd[pre + "out_layers.0.bias"])))
    h = F.conv2d(h3, W(sd, pre + "out_layers.3.weight"), sd[pre + "out_layers.3.bias"], padding=1)
    if pre + "skip_connection.weight" in sd:
        x = rb(F.conv2d(x, W(sd, pre + "skip_connection.weight"), sd[pre + "skip_connection.bias"]))
    return rs(x + h)

def attn_block(sd, pre, x):
    b, c, *sp = x.shape
    xr = x.reshape(b, c, -1)
    hn = rb(R.group_norm32(xr, sd[pre + "norm.weight"], sd[pre + "norm.bias"]))
    qkv = rb(F.conv1d(hn, W(sd, pre + "qkv.weight"), sd[pre + "qkv.bias"]))
    a = rb(R.qkv_attention_legacy(qkv, 1))
    h = F.conv1d(a, W(sd, pre + "proj_out.weight"), sd[pre + "proj_out.bias"])
    return rs(xr + h).reshape(b, c, *sp)

def forward(sd, cfg, image, cond, featmean, t):
    a = F.relu(F.linear(featmean, sd["audio_transformer.transform.0.weight"], sd["audio_transformer.transform.0.bias"]))
    S = cfg["image_size"]
    a = a.view(-1, a.shape[1], 1, 1).expand(-1, -1, S, S)
    im = F.conv2d(cond, sd["cond_conv_in.weight"])
    x = rb(torch.cat([image, im, a], 1))
    mc, nrb = cfg["model_channels"], cfg["num_res_blocks"]
    emb = F.linear(R.timestep_embedding(t, mc), sd["time_embed.0.weight"], sd["time_embed.0.bias"])
    emb = F.linear(F.silu(emb), sd["time_embed.2.weight"], sd["time_embed.2.bias"])
    hs = []
    h = rs(F.conv2d(x, W(sd, "input_blocks.0.0.weight"), sd["input_blocks.0.0.bias"], padding=1)); hs.append(h)
    idx, nl = 1, len(cfg["channel_mult"])
    for level in range(nl):
        for _ in range(nrb):
            h = res_block(sd, f"input_blocks.{idx}.0.", h, emb)
            h = attn_block(sd, f"input_blocks.{idx}.1.", h)
            hs.append(h); idx += 1
        if level != nl - 1:
            h = rs(F.conv2d(h, W(sd, f"input_blocks.{idx}.0.op.weight"), sd[f"input_blocks.{idx}.0.op.bias"], stride=2, padding=1))
            hs.append(h); idx += 1
    h = res_block(sd, "middle_block.0.", h, emb); h = attn_block(sd, "middle_block.1.", h); h = res_block(sd, "middle_block.2.", h, emb)
    idx = 0
    for level in reversed(range(nl)):
        for i in range(nrb + 1):
            h = torch.cat([h, hs.pop()], 1)
            h = res_block(sd, f"output_blocks.{idx}.0.", h, emb)
            h = attn_block(sd, f"output_blocks.{idx}.1.", h)
            if level and i == nrb:
                h = rb(F.interpolate(h, scale_factor=2, mode="nearest"))
                h = rs(F.conv2d(h, W(sd, f"output_blocks.{idx}.2.conv.weight"), sd[f"output_blocks.{idx}.2.conv.bias"], padding=1))
            idx += 1
    h = rb(F.silu(R.group_norm32(h, sd["out.0.weight"], sd["out.0.bias"])))
    return F.conv2d(h, W(sd, "out.2.weight"), sd["out.2.bias"], padding=1)

S, N = (int([a for a in sys.argv[1:] if a.isdigit()][0]) if [a for a in sys.argv[1:] if a.isdigit()] else 32), 2
cfg = dict(image_size=S, model_channels=64, num_res_blocks=2, attention_resolutions=(1, 2, 4), channel_mult=(1, 2, 4), num_heads=1)
import lipreading_video_generation_b200 as P
class Stub(torch.nn.Module):
    def forward(self, a): return a["feats"]
import os; os.environ["B2D_WAV2VEC2_RANDOM_INIT"] = "1"
m = P.UNetAudio(image_size=S, in_channels=3, model_channels=64, out_channels=3, num_res_blocks=2, attention_resolutions=(1, 2, 4),
                dropout=0.0, audio_feature_dim=768, projected_audio_dim=128, audio_encoder=Stub())
sd0 = (R.protocol_params if 'protocol' in sys.argv else R.fill_params)(m.state_dict(), seed=3)
S_arg = [a for a in sys.argv[1:] if a.isdigit()]
g = torch.Generator().manual_seed(5)
x0 = torch.rand(N, 3, S, S, generator=g) * 2 - 1; cond = torch.rand(N, 3, S, S, generator=g) * 2 - 1
feats = torch.randn(N, 12, 768, generator=g) * 0.5; eps = torch.randn(N, 3, S, S, generator=g); t = torch.randint(0, 1000, (N,), generator=g)
xt = R.add_noise(R.linear_tables(1000, 0.00085, 0.012), x0, eps, t)

def run(fn):
    sd = {k: v.clone().requires_grad_(True) for k, v in sd0.items()}
    p = fn(sd); l = R.mse(p, eps); l.backward()
    return p.detach(), {k: v.grad for k, v in sd.items()}

ref_p, ref_g = run(lambda sd: R.unet_audio_forward(sd, cfg, xt, cond, feats.mean(1), t))
rel = lambda a, b: ((a - b).norm() / (b.norm() + 1e-20)).item()
variants = {
    "all bf16 (engine as built)": {},
    "fp32 fwd stream": dict(act_stream=False),
    "fp32 grad stream": dict(grad_stream=False),
    "fp32 fwd+grad stream": dict(act_stream=False, grad_stream=False),
    "fp32 weights only": dict(w=False),
    "fp32 branch grads": dict(grad_branch=False),
    "fp32 all grads": dict(grad_branch=False, grad_stream=False),
    "fp32 all acts": dict(act_branch=False, act_stream=False),
    "only weights bf16": dict(act_branch=False, act_stream=False, grad_branch=False, grad_stream=False),
    "fp16 fwd tensors, bf16 grads": dict(_fwd=torch.float16),
    "fp16 fwd tensors, fp16 grads x2^14": dict(_fwd=torch.float16, _grad="f16s"),
}
print(f"{'variant':32s} {'pred':>9s} {'g median':>9s} {'g worst':>9s} {'n>2e-2':>7s}")
for name, ov in variants.items():
    for k in FLAGS: FLAGS[k] = True
    DT["fwd"] = ov.pop("_fwd", torch.bfloat16); DT["grad"] = ov.pop("_grad", torch.bfloat16)
    if DT["grad"] == "f16s": continue
    FLAGS.update(ov)
    p, gr = run(lambda sd: forward(sd, cfg, xt, cond, feats.mean(1), t))
    errs = [rel(gr[k], ref_g[k]) for k in ref_g if ref_g[k] is not None and gr[k] is not None]
    print(f"{name:32s} {rel(p, ref_p):9.2e} {statistics.median(errs):9.2e} {max(errs):9.2e} {sum(e > 2e-2 for e in errs):7d}")
