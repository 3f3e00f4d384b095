"""Per-parameter gradient cosine of the product vs the golden reference for i2i2m, next to the
bf16-emulating oracle's and to an fp32 oracle run on a bf16-rounded input (sensitivity probe)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from oracle import nets
from tests.util import cosine
from tests.test_models_gpu import _product, _oracle, G

name = sys.argv[1] if len(sys.argv) > 1 else "i2i2m"
fx = torch.load(os.path.join(G, name + ".pt"))
sd = nets.init_state_dict(name, fx["seed"], fx["jitter"])
runs = []
for i in range(2):
    outs, loss, grads, m = _product(name, sd, fx["x"], fx["y"])
    runs.append(grads)
_, _, e_grads, _ = _oracle(name, sd, fx["x"], fx["y"], emulate=True)
xr = fx["x"].bfloat16().float()
_, _, p_grads, _ = _oracle(name, sd, xr, fx["y"])
for k, r in fx["grads"].items():
    if float(r.norm()) > 1e-4:
        c = [cosine(g[k], r) for g in runs]
        ce, cp = cosine(e_grads[k], r), cosine(p_grads[k], r)
        flag = " <<<" if min(c) < min(0.99, 1 - 2 * (1 - ce)) else ""
        if min(c) < 0.995 or flag:
            print(f"{k:55s} n={r.numel():7d} |r|={float(r.norm()):.2e} ours={c[0]:.4f},{c[1]:.4f} emu={ce:.4f} x-bf16-perturbed-fp32={cp:.4f}{flag}")
