"""Row N1 parity on N GPUs (run under torchrun): one UNetModel(dims=3) clip, train step with the attention sharded by
query rows across the ranks vs the same step un-sharded on every rank.  Prints per-rank relative errors of output,
loss and every gradient, and checks that the ranks hold identical results after the collectives.
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 scripts/seqpar_check.py"""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
import lipreading_video_generation_b200 as P
from lipreading_video_generation_b200.parallel import SequenceParallelAttention, broadcast_parameters

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
T, S = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (8, 64)
torch.manual_seed(0)
model = P.UNetModel(image_size=S, in_channels=195, model_channels=64, out_channels=3, num_res_blocks=1,
                    attention_resolutions=(1, 2), dropout=0.0, channel_mult=(1, 2), dims=3)
g = torch.Generator().manual_seed(1)
with torch.no_grad():
    for name, p in model.named_parameters():
        if (".out_layers.3." in name) or (".proj_out." in name) or name.startswith("out.2."):
            p.copy_(torch.randn(p.shape, generator=g) * 0.02)
model = model.to(dev).train()
broadcast_parameters(model)
x = torch.randn(1, 195, T, S, S, generator=g).to(dev)
eps = torch.randn(1, 3, T, S, S, generator=g).to(dev)
t = torch.tensor([417], device=dev)
crit = P.MSELoss()


def run():
    for p in model.parameters():
        p.grad = None
    out = model(x, t)
    loss = crit(out, eps)
    loss.backward()
    return out.detach().clone(), loss.detach().clone(), [p.grad.detach().clone() for p in model.parameters()]


ref = run()                                           # un-sharded on every rank
sp = SequenceParallelAttention(model, min_rows=256)
assert model.engine().seq_parallel is sp
got = run()
rel = lambda a, b: ((a.float() - b.float()).norm() / (b.float().norm() + 1e-20)).item()
e_out, e_loss = rel(got[0], ref[0]), abs(got[1].item() - ref[1].item()) / abs(ref[1].item())
e_g = [rel(a, b) for a, b in zip(got[2], ref[2])]
# after the collectives every rank must hold the same numbers
mx = torch.tensor([e_out, e_loss, max(e_g)], device=dev)
same = got[0].clone()
dist.broadcast(same, 0)
ident = torch.equal(same, got[0])
dist.all_reduce(mx, op=dist.ReduceOp.MAX)
if rank == 0:
    print(f"world {world} L_full {T * S * S}: output rel {mx[0].item():.3e}, loss rel {mx[1].item():.3e}, "
          f"worst gradient rel {mx[2].item():.3e} over {len(e_g)} tensors; bytes moved {sp.bytes_moved}; "
          f"ranks identical: {ident}")
ok = mx[0].item() <= 2e-3 and mx[1].item() <= 1e-4 and mx[2].item() <= 5e-3
dist.barrier()
dist.destroy_process_group()
sys.exit(0 if ok else 1)
