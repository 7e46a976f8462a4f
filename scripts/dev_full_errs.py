"""Full train.py-config UNetAudio (model_channels 64, (1,2,4), 2 res blocks, attention at every level) at S=64:
product on GPU vs CPU oracle restatement.  Prints output / loss / gradient errors."""
import os, sys, time, torch
sys.path.insert(0, ".")
os.environ["B2D_WAV2VEC2_RANDOM_INIT"] = "1"
import lipreading_video_generation_b200 as P
from oracle import restatement as R

class Stub(torch.nn.Module):
    def forward(self, audio): return audio["feats"]

S = int(sys.argv[1]) if len(sys.argv) > 1 else 64
N = int(sys.argv[2]) if len(sys.argv) > 2 else 2
INIT = sys.argv[3] if len(sys.argv) > 3 else 'fill'
cfg = dict(image_size=S, in_channels=3, model_channels=64, out_channels=3, num_res_blocks=2,
           attention_resolutions=(1, 2, 4), dropout=0.0, audio_feature_dim=768, projected_audio_dim=128)
model = P.UNetAudio(**cfg, audio_encoder=Stub())
sd = (R.protocol_params if INIT == 'protocol' else R.fill_params)(model.state_dict(), seed=3)
print('config: S', S, 'N', N, 'init', INIT)
model.load_state_dict(sd)
model = model.cuda().train()
g = torch.Generator().manual_seed(5)
x0 = torch.rand(N, 3, S, S, generator=g) * 2 - 1; cond = torch.rand(N, 3, S, S, generator=g) * 2 - 1
feats = torch.randn(N, 12, 768, generator=g) * 0.5; eps = torch.randn(N, 3, S, S, generator=g)
t = torch.randint(0, 1000, (N,), generator=g)
tab = R.linear_tables(1000, 0.00085, 0.012)
xt = R.add_noise(tab, x0, eps, t)
pred = model(xt.cuda(), cond.cuda(), {"feats": feats.cuda()}, t.cuda())
loss = P.MSELoss()(pred, eps.cuda()); loss.backward(); torch.cuda.synchronize()
t0 = time.time()
sdr = {k: v.clone().requires_grad_(True) for k, v in sd.items()}
rcfg = dict(image_size=S, model_channels=64, num_res_blocks=2, attention_resolutions=(1, 2, 4), channel_mult=(1, 2, 4), num_heads=1)
rp = R.unet_audio_forward(sdr, rcfg, xt, cond, feats.mean(1), t); rl = R.mse(rp, eps); rl.backward()
print(f"oracle cpu time {time.time()-t0:.1f}s")
rel = lambda a, b: ((a.detach().float().cpu() - b.detach().float()).norm() / (b.detach().float().norm() + 1e-20)).item()
print("pred rel", rel(pred, rp), "loss", loss.item(), rl.item())
errs = sorted((rel(p.grad, sdr[k].grad), k) for k, p in model.named_parameters())
import statistics
print("grad rel err: median", statistics.median(e for e, _ in errs), "worst", errs[-8:])
print("n > 2e-2:", sum(e > 2e-2 for e, _ in errs), "of", len(errs))
