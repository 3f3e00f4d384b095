"""Per-tensor gradient error of a product model vs the fp32 oracle (and the oracle's ideal-bf16 emulation vs fp32) on
the trained fixture at 256x256: where does the whole-gradient error come from?
Usage: python scripts/grad_diag.py fct [B]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from oracle import nets  # noqa: E402
from tests import test_parity_full_gpu as T  # noqa: E402
from tests.golden import trained  # noqa: E402
from tests.util import cosine, rel_l2  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "fct"
B = int(sys.argv[2]) if len(sys.argv) > 2 else 2
sd = T._trained(name)
x, y = trained.blobs(B, 256, 4242)
outs, loss, grads = T._product(name, sd, x, y)
r_outs, r_loss, r_grads = T._oracle_gpu(name, sd, x, y, 2 if name == "fct" else None)
with nets.bf16_emulation():
    e_outs, e_loss, e_grads = T._oracle_gpu(name, sd, x, y, 2 if name == "fct" else None)
tot = sum(float((grads[k].float() - r).norm() ** 2) for k, r in r_grads.items())
tot_e = sum(float((e_grads[k].float() - r).norm() ** 2) for k, r in r_grads.items())
gn = sum(float(r.norm() ** 2) for r in r_grads.values())
print(f"{name} B={B}: product rel {(tot / gn) ** 0.5:.4f}   emulation rel {(tot_e / gn) ** 0.5:.4f}")
print(f"outputs: product rel {rel_l2(outs[-1], r_outs[-1]):.5f}  emulation rel {rel_l2(e_outs[-1], r_outs[-1]):.5f}   "
      f"product max-abs {float((outs[-1] - r_outs[-1]).abs().max()):.4f}  emulation max-abs {float((e_outs[-1] - r_outs[-1]).abs().max()):.4f}")
ga = torch.cat([grads[k].flatten().float() for k in r_grads]); gb = torch.cat([r_grads[k].flatten().float() for k in r_grads])
ge = torch.cat([e_grads[k].flatten().float() for k in r_grads])
print(f"whole gradient: product cos {cosine(ga, gb):.5f} rel {rel_l2(ga, gb):.4f}   emulation cos {cosine(ge, gb):.5f} rel {rel_l2(ge, gb):.4f}")
rows = []
for k, r in r_grads.items():
    d = float((grads[k].float() - r).norm() ** 2)
    de = float((e_grads[k].float() - r).norm() ** 2)
    rows.append((d / tot, k, cosine(grads[k], r), rel_l2(grads[k], r), cosine(e_grads[k], r), rel_l2(e_grads[k], r), float(r.norm()), de / max(tot_e, 1e-30)))
rows.sort(reverse=True)
print(f"{'share':>6s} {'emu sh':>6s} {'cos':>8s} {'rel':>7s} {'emu cos':>8s} {'emu rel':>7s} {'|g|':>9s}  tensor")
for sh, k, c, rl, ce, re, n, she in rows[:40]:
    print(f"{sh:6.3f} {she:6.3f} {c:8.5f} {rl:7.4f} {ce:8.5f} {re:7.4f} {n:9.3g}  {k}")
