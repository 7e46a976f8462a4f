"""List host<->device synchronisation points inside one train step (torch.cuda.set_sync_debug_mode('warn'))."""
import os, sys, warnings, traceback
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ.setdefault("B2D_WAV2VEC2_RANDOM_INIT", "1")
import torch
import bench

dev = torch.device("cuda:0")
model = bench.build_model(64, dev)
step = bench.TrainStep(model, 64, 1, dev)
n = 16
x0, cond, audio = bench.make_batch(n, 64, 1234, dev)
t = torch.randint(0, 1000, (n,), device=dev)
for _ in range(3):
    step(x0, cond, audio, t)
torch.cuda.synchronize()
seen = {}
def showwarning(message, category, filename, lineno, file=None, line=None):
    if "synchroniz" not in str(message):
        return
    st = [f for f in traceback.extract_stack()[:-1] if "warnings" not in f.filename]
    key = tuple((f.filename.split("site-packages/")[-1], f.lineno) for f in st[-4:])
    seen[key] = seen.get(key, 0) + 1
warnings.showwarning = showwarning
warnings.simplefilter("always")
torch.cuda.set_sync_debug_mode("warn")
step(x0, cond, audio, t)
torch.cuda.set_sync_debug_mode("default")
for k, v in seen.items():
    print(v, "x", " <- ".join(f"{a}:{b}" for a, b in reversed(k)))
print("sync points:", sum(seen.values()))
