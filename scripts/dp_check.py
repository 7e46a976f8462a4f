"""Data-parallel gradient identity on real NCCL ranks (SURVEY section 4 test 5; run under torchrun on >= 2 GPUs):
rank-averaged gradients of the sharded batch (GradientAllReducer: torso buckets from the backward's stage hooks, the
wav2vec2 encoder through the persistent flat buffer and per-parameter hooks) == gradients of the concatenated batch
computed on one rank with no reducer.  The full UNetAudio incl. the HF wav2vec2 encoder (kept in eval mode here:
LayerDrop / dropout / SpecAugment draw per-rank random numbers that a 1-rank run cannot replay) is used, so the 94 M
encoder gradients go through the hook path.  A second pass drops one encoder layer's gradients on rank 1 only and
checks that every rank still holds identical averaged gradients (fixed bucket layout, missing = zeros).

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 scripts/dp_check.py
Writes one JSON line to stdout and gpurun_out/dp_check.json (rank 0)."""
import json
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ.setdefault("B2D_WAV2VEC2_RANDOM_INIT", "1")
import torch
import torch.distributed as dist
import lipreading_video_generation_b200 as P
from lipreading_video_generation_b200.parallel import GradientAllReducer, broadcast_parameters, shard_range

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
S, per = 32, 2
N = per * world
torch.manual_seed(0)
model = P.UNetAudio(image_size=S, in_channels=3, model_channels=64, out_channels=3, num_res_blocks=1,
                    attention_resolutions=(1, 2), channel_mult=(1, 2), dropout=0.0, audio_feature_dim=768,
                    projected_audio_dim=128)
g = torch.Generator().manual_seed(1)
with torch.no_grad():
    for name, p in model.named_parameters():
        if (".out_layers.3." in name) or (".proj_out." in name) or name.startswith("out.2."):
            p.copy_(torch.randn(p.shape, generator=g) * 0.02)
model = model.to(dev).train()
model.audio_encoder.eval()
broadcast_parameters(model)
xt = torch.randn(N, 3, S, S, generator=g).to(dev)
cond = (torch.rand(N, 3, S, S, generator=g) * 2 - 1).to(dev)
audio = torch.randn(N, 4000, generator=g).to(dev)
eps = torch.randn(N, 3, S, S, generator=g).to(dev)
t = torch.randint(0, 1000, (N,), generator=g).to(dev)
crit = P.MSELoss()
names = [k for k, _ in model.named_parameters()]


def grads(lo, hi, reducer=None, drop=None):
    model.zero_grad(set_to_none=True)
    loss = crit(model(xt[lo:hi], cond[lo:hi], {"input_values": audio[lo:hi]}, t[lo:hi]), eps[lo:hi])
    loss.backward()
    if drop:
        for k, p in model.named_parameters():
            if k.startswith(drop):
                p.grad = None
    if reducer is not None:
        reducer.finish()
    torch.cuda.synchronize()
    return {k: (p.grad.detach().float().clone() if p.grad is not None else None) for k, p in model.named_parameters()}


def rel(a, b, floor=0.0):
    return ((a - b).norm() / max(b.norm().item(), floor)).item()


full = grads(0, N)                                   # 1-rank gradients of the concatenated batch (every rank computes them)
red = GradientAllReducer(model, bucket_mb=4.0)
lo, hi = shard_range(N, rank, world)
got = grads(lo, hi, red)
# Some encoder gradients are zero in exact arithmetic (a key bias cannot change a softmax, a bias in front of a
# normalisation cancels): their computed values are rounding noise of size ~1e-6 of the typical gradient norm, so the
# relative error is taken against max(||ref||, 1e-3 * median gradient norm of the group).
enc_norms = sorted(full[k].norm().item() for k in names if full[k] is not None and k.startswith("audio_encoder."))
floor_enc = 1e-3 * enc_norms[len(enc_norms) // 2]
errs = sorted((rel(got[k], full[k], floor_enc if k.startswith("audio_encoder.") else 0.0), k)
              for k in names if full[k] is not None)
n_torso = len(model.engine().torso_params())
worst_torso = max((e, k) for e, k in errs if not k.startswith("audio_encoder."))
worst_enc = max((e, k) for e, k in errs if k.startswith("audio_encoder."))
if rank == 0:
    print("worst torso", worst_torso, "worst encoder", worst_enc, "encoder floor", floor_enc, flush=True)
worst_torso, worst_enc = worst_torso[0], worst_enc[0]
# every rank must hold bit-identical averaged gradients
ident = True
for k in names:
    if got[k] is None:
        continue
    ref = got[k].clone()
    dist.broadcast(ref, 0)
    ident = ident and torch.equal(ref, got[k])
# LayerDrop-style mismatch: rank 1 skips encoder layer 3 this step (HF draws the skip per rank) -> its parameters get
# no gradient there; the bucket layout must not change and every rank must end with identical averaged gradients
red.detach()
red2 = GradientAllReducer(model, bucket_mb=4.0)
enc = model.audio_encoder.wav2vec2.encoder
all_layers = enc.layers
if rank == 1:
    enc.layers = torch.nn.ModuleList([l for i, l in enumerate(all_layers) if i != 3])
model.zero_grad(set_to_none=True)
loss = crit(model(xt[lo:hi], cond[lo:hi], {"input_values": audio[lo:hi]}, t[lo:hi]), eps[lo:hi])
loss.backward()
red2.finish()
enc.layers = all_layers
torch.cuda.synchronize()
ident2 = True
for k, p in model.named_parameters():
    gk = p.grad.detach().float().clone()
    ref = gk.clone()
    dist.broadcast(ref, 0)
    ident2 = ident2 and torch.equal(ref, gk)
flags = torch.tensor([worst_torso, worst_enc, 0.0 if ident else 1.0, 0.0 if ident2 else 1.0], device=dev)
dist.all_reduce(flags, op=dist.ReduceOp.MAX)
ok = flags[0].item() <= 5e-3 and flags[1].item() <= 2e-2 and flags[2].item() == 0.0 and flags[3].item() == 0.0
if rank == 0:
    line = {"check": "dp_gradient_identity", "world": world, "images_per_rank": per, "n_param_tensors": len(names),
            "n_torso_tensors": n_torso, "worst_rel_torso": flags[0].item(), "worst_rel_wav2vec2": flags[1].item(),
            "ranks_bit_identical": flags[2].item() == 0.0, "ranks_bit_identical_with_missing_layer": flags[3].item() == 0.0,
            "bytes_reduced_per_step": red.bytes_reduced, "ok": ok}
    print(json.dumps(line))
    os.makedirs("gpurun_out", exist_ok=True)
    with open("gpurun_out/dp_check.json", "w") as f:
        f.write(json.dumps(line) + "\n")
dist.barrier()
dist.destroy_process_group()
sys.exit(0 if ok else 1)
