"""Is the 5 % gradient-scale gap of FCT on the trained fixture in the loss kernels or in the model?"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import hybridnet  # noqa: E402
from oracle import nets  # noqa: E402
from tests import test_parity_full_gpu as T  # noqa: E402
from tests.golden import trained  # noqa: E402
from tests.util import cosine, rel_l2  # noqa: E402

name = "fct"
sd = T._trained(name)
x, y = trained.blobs(2, 256, 4242)


def product(kernel_loss):
    m = hybridnet.MODELS[name]()
    m.load_state_dict(sd, strict=True)
    m = m.cuda().train()
    for mod in m.modules():
        if isinstance(mod, torch.nn.Dropout):
            mod.p = 0.0
    out = m(x.cuda())
    out.retain_grad()
    loss = hybridnet.DiceLoss()(out, y.cuda()) if kernel_loss else trained.loss_of(name, out, x.cuda(), y.cuda())
    loss.backward()
    g = torch.cat([p.grad.flatten().float().cpu() for p in m.parameters() if p.grad is not None])
    return out.detach().cpu(), out.grad.detach().cpu(), g, float(loss)


ok, gk, Gk, lk = product(True)
ot, gt, Gt, lt = product(False)
r_outs, r_loss, r_grads = T._oracle_gpu(name, sd, x, y)
names = [k for k in r_grads]
import hybridnet as H  # noqa
m = H.MODELS[name]()
order = [n for n, p in m.named_parameters() if n in r_grads]
Gr = torch.cat([r_grads[k].flatten().float() for k in order])
print("loss kernel/torch/oracle", lk, lt, r_loss)
print("dL/dout kernel vs torch-op: rel", rel_l2(gk, gt), "cos", cosine(gk, gt), "norm ratio", float(gk.norm() / gt.norm()))
print("grads kernel-loss vs oracle: rel", rel_l2(Gk, Gr), "cos", cosine(Gk, Gr), "norm ratio", float(Gk.norm() / Gr.norm()))
print("grads torch-loss  vs oracle: rel", rel_l2(Gt, Gr), "cos", cosine(Gt, Gr), "norm ratio", float(Gt.norm() / Gr.norm()))
# head: sigma' from the bf16-stored mask vs from fp32
p = r_outs[-1]
print("mask stats: frac>0.99", float((p > 0.99).float().mean()), "frac<0.01", float((p < 0.01).float().mean()), "mean", float(p.mean()))
sg_true = p * (1 - p)
pb = ok
sg_b = pb * (1 - pb)
w = gt.abs()
print("sum |dL/dout| * sigma': fp32", float((w * sg_true).sum()), " bf16-stored", float((w * sg_b).sum()))
