"""Kernel launches of ONE eager training step by kernel name (torch.profiler): how many of them are not ours (at::).
  python scripts/count_launches.py [fct|unet|ae32k_image] [batch]"""
import collections
import os
import re
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import hybridnet
from hybridnet import ops, parallel

name = sys.argv[1] if len(sys.argv) > 1 else "fct"
B = int(sys.argv[2]) if len(sys.argv) > 2 else 4
dev = torch.device("cuda")
torch.manual_seed(0)
model = hybridnet.MODELS[name]().to(dev).train()
dead = parallel.used_parameter_names(name)
used = {n for n, p in model.named_parameters() if p.requires_grad and n not in dead}
arena = parallel.GradArena(model, used=used, buckets=4).direct()
opt = hybridnet.optim.AdamW(model.parameters(), arena, lr=1e-4)
recon = name in ("autoencoder", "ae32k_image")
crit = hybridnet.MSELoss() if recon else hybridnet.DiceLoss()
x = torch.rand(B, 3, 256, 256, device=dev)
y = (torch.rand(B, 1, 256, 256, device=dev) > 0.5).float()
cap = sum(p.numel() for p in model.parameters()) + (1 << 18)


def step():
    opt.zero_grad()
    ops.begin_step(dev, cap)
    out = model(x)
    out = out[-1] if isinstance(out, tuple) else out
    loss = crit(out, x if recon else y)
    loss.backward()
    arena.finish()
    opt.step()


for _ in range(3):
    step()
torch.cuda.synchronize()
with torch.profiler.profile(activities=[torch.profiler.ProfilerActivity.CUDA]) as prof:
    step()
    torch.cuda.synchronize()
cnt = collections.Counter()
for e in prof.events():
    if e.device_type == torch.autograd.DeviceType.CUDA:
        cnt[re.sub(r"^void |<unnamed>::|\(anonymous namespace\)::", "", e.name).split("(")[0][:70] or e.name[:70]] += 1
tot = sum(cnt.values())
foreign = {k: v for k, v in cnt.items() if k.startswith("at::") or "Memset" in k or "Memcpy" in k or "nccl" in k.lower()}
print(f"{name} B={B}: {tot} device activities in one step; not hybridnet kernels: {sum(foreign.values())}")
for k, v in sorted(foreign.items(), key=lambda kv: -kv[1]):
    print(f"  {v:4d}  {k}")
print("top of ours:")
for k, v in [kv for kv in cnt.most_common(40) if kv[0] not in foreign][:12]:
    print(f"  {v:4d}  {k}")
