"""Debug aid: UNet gradients with the weight gradients serial / on the side stream / deferred, pairwise distances."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import hybridnet
from hybridnet import ops, parallel
from tests.util import rel_l2

torch.manual_seed(0)
feat = int(sys.argv[1]) if len(sys.argv) > 1 else 32
m = hybridnet.UNet(3, 1, init_features=feat).cuda()
m = m.train() if len(sys.argv) > 2 and sys.argv[2] == "train" else m.eval()
arena = parallel.GradArena(m).direct(True)
x, y = torch.rand(4, 3, 128, 128).cuda(), (torch.rand(4, 1, 128, 128) > 0.5).float().cuda()
crit = hybridnet.DiceLoss()


def grads(overlap, defer):
    ops.WGRAD_OVERLAP, ops.WGRAD_DEFER = overlap, defer
    arena.zero_grad()
    ops.begin_step(x.device, sum(p.numel() for p in m.parameters()) + (1 << 16))
    crit(m(x), y).backward()
    torch.cuda.synchronize()
    return arena.arena.clone()


res = {}
for name, cfg in (("serial", (False, False)), ("serial2", (False, False)), ("side", (True, False)), ("side2", (True, False)),
                  ("defer", (True, True)), ("defer2", (True, True))):
    res[name] = grads(*cfg)
base = res["serial"]
for k, v in res.items():
    print(f"{k:8s} vs serial: {rel_l2(v, base):.3e}")
# per-parameter worst offenders of the deferred run
worst = []
for n, (s, e) in zip(arena.names, arena.slices):
    a, b = res["defer"][s:e], base[s:e]
    if float(b.norm()) > 0:
        worst.append((rel_l2(a, b), n))
for r, n in sorted(worst, reverse=True)[:8]:
    print(f"  {r:.3e} {n}")
