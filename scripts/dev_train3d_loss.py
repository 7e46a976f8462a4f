"""Loss trajectory of the train3d_small bench workload on one GPU (diagnostic: where does it stop being finite?)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import lipreading_video_generation_b200 as P
lr = float(sys.argv[1]) if len(sys.argv) > 1 else 1e-2
which = sys.argv[2] if len(sys.argv) > 2 else "ours"
world = int(os.environ.get("WORLD_SIZE", "1")); rank = int(os.environ.get("RANK", "0"))
dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", "0")))
torch.cuda.set_device(dev)
if world > 1:
    torch.distributed.init_process_group("nccl", device_id=dev)
S, frames = 64, 16
torch.manual_seed(0)
model = P.UNetModel(image_size=S, in_channels=195, model_channels=64, out_channels=3, num_res_blocks=2,
                    attention_resolutions=(1, 2, 4), dropout=0.1, channel_mult=(1, 2, 4), dims=3)
g = torch.Generator().manual_seed(1)
with torch.no_grad():
    for name, p in model.named_parameters():
        if (".out_layers.3." in name) or (".proj_out." in name) or name.startswith("out.2."):
            p.copy_(torch.randn(p.shape, generator=g) * 0.02)
model = model.to(dev).train()
opt = P.Adam(model.parameters(), lr=lr) if which == "ours" else torch.optim.Adam(model.parameters(), lr=lr, foreach=True)
crit = P.MSELoss()
reducer = None
if world > 1:
    from lipreading_video_generation_b200.parallel import GradientAllReducer, SequenceParallelAttention, broadcast_parameters
    broadcast_parameters(model); sp = SequenceParallelAttention(model); reducer = GradientAllReducer(model)
gi = torch.Generator().manual_seed(1234)
x = torch.randn(1, 195, frames, S, S, generator=gi).to(dev)
eps = torch.randn(1, 3, frames, S, S, generator=gi).to(dev)
t = torch.randint(0, 1000, (1,), generator=gi).to(dev)
for i in range(8):
    opt.zero_grad(set_to_none=True)
    out = model(x, t)
    loss = crit(out, eps)
    loss.backward()
    if reducer is not None:
        reducer.finish()
    gn = torch.stack([p.grad.float().norm() for p in model.parameters()]).norm().item()
    opt.step()
    wmax = max(p.detach().abs().max().item() for p in model.parameters())
    if rank == 0:
        print(f"world {world} {which} lr {lr} step {i}: loss {loss.item():.5f} out absmax {out.abs().max().item():.3f} grad norm {gn:.4e} weight absmax {wmax:.3f}")
if world > 1:
    torch.distributed.destroy_process_group()
