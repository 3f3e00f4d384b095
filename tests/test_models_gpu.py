"""End-to-end parity of the drop-in modules (CUDA path through the C ABI) against
 (1) the committed golden vectors produced by the real reference classes, and
 (2) the CPU oracle (oracle/nets.py, fp32) on the same seeded inputs and weights.
Tolerances (fp32 reference vs bf16-operand / fp32-accumulate kernels):
 outputs max-abs <= 5e-2 and rel-L2 <= 2e-2; mask IoU@0.5 >= 0.999 where masks straddle 0.5;
 parameter-gradient cosine >= 0.999 and rel-L2 <= 3e-2 over the whole gradient vector.
"""
import os

import pytest
import torch

from oracle import nets
from tests.util import cosine, iou, max_abs, rel_l2

pytestmark = pytest.mark.gpu
G = os.path.join(os.path.dirname(__file__), "golden")
OUT_ABS, OUT_REL, GRAD_COS, GRAD_REL = 5e-2, 2e-2, 0.999, 3e-2


def _loss(name, outs, x, y):
    if name in ("autoencoder", "ae32k_image"):
        return torch.nn.functional.mse_loss(outs[-1], x)
    return nets.dice_loss(outs[-1], y)


def _oracle(name, sd, x, y, training=True):
    params = {k: (v.clone().requires_grad_(True) if v.is_floating_point() and "running" not in k else v.clone())
              for k, v in sd.items()}
    sink = {}
    out = nets.FORWARDS[name](params, x, nets.Ctx(training=training, dropout="off", bn_sink=sink))
    outs = out if isinstance(out, tuple) else (out,)
    loss = _loss(name, outs, x, y)
    loss.backward()
    return [o.detach() for o in outs], float(loss), {k: p.grad for k, p in params.items() if torch.is_tensor(p) and p.requires_grad}, sink


def _product(name, sd, x, y, training=True):
    import hybridnet
    m = hybridnet.MODELS[name]()
    m.load_state_dict(sd, strict=True)
    m = m.cuda()
    m.train(training)
    for mod in m.modules():                       # dropout off: RNG streams cannot match (H5)
        if isinstance(mod, (torch.nn.Dropout, torch.nn.Dropout2d)):
            mod.p = 0.0
    xd, yd = x.cuda(), y.cuda()
    out = m(xd)
    outs = out if isinstance(out, tuple) else (out,)
    loss = _loss(name, outs, xd, yd)
    loss.backward()
    grads = {k: (p.grad.detach().cpu() if p.grad is not None else None) for k, p in m.named_parameters()}
    return [o.detach().cpu() for o in outs], float(loss), grads, m


def _compare(outs, ref_outs, grads, ref_grads, rel=OUT_REL):
    for o, r in zip(outs, ref_outs):
        assert o.shape == r.shape
        assert max_abs(o, r) <= OUT_ABS, max_abs(o, r)
        assert rel_l2(o, r) <= rel, rel_l2(o, r)
    ga, gb = [], []
    for k, r in ref_grads.items():
        g = grads[k]
        if r is None:
            assert g is None, f"{k}: never-used parameter must keep grad=None (SURVEY note U)"
            continue
        assert g is not None, k
        ga.append(g.flatten())
        gb.append(r.flatten())
        if float(r.norm()) > 1e-4:
            assert cosine(g, r) > 0.99, (k, cosine(g, r))
    ga, gb = torch.cat(ga), torch.cat(gb)
    assert cosine(ga, gb) >= GRAD_COS, cosine(ga, gb)
    assert rel_l2(ga, gb) <= GRAD_REL, rel_l2(ga, gb)


@pytest.mark.parametrize("name", ["unet", "autoencoder", "i2i2m", "fct"])
def test_against_reference_golden(name):
    """Golden vectors from the unmodified reference classes (tests/golden/make_golden.py)."""
    fx = torch.load(os.path.join(G, name + ".pt"))
    sd = nets.init_state_dict(name, fx["seed"], fx["jitter"])
    outs, loss, grads, m = _product(name, sd, fx["x"], fx["y"])
    for o, r in zip(outs, fx["outputs"]):
        assert max_abs(o, r) <= OUT_ABS and rel_l2(o, r) <= OUT_REL, (max_abs(o, r), rel_l2(o, r))
    assert abs(loss - fx["loss"]) <= 5e-3
    for k, n in fx["grad_norms"].items():
        if n is None:
            assert grads[k] is None, k
        else:
            assert grads[k] is not None, k
    for k, r in fx["grads"].items():
        if float(r.norm()) > 1e-4:
            assert cosine(grads[k], r) > 0.99, (k, cosine(grads[k], r))
    msd = m.state_dict()
    for k, r in fx["bn_after"].items():
        assert torch.allclose(msd[k].cpu().float(), r.float(), atol=5e-3, rtol=2e-2), k


@pytest.mark.parametrize("name,B,S", [("unet", 2, 64), ("autoencoder", 2, 64), ("fct", 2, 64), ("i2i2m", 1, 64),
                                      ("ae32k_image", 2, 32), ("ae32k_mask", 2, 32), ("latent2mask", 2, 32)])
def test_against_oracle_small(name, B, S):
    sd = nets.init_state_dict(name, seed=5, jitter=0.2)
    g = torch.Generator().manual_seed(1)
    x = torch.rand(B, 3, S, S, generator=g)
    y = (torch.rand(B, 1, S, S, generator=g) > 0.5).float()
    if name == "latent2mask":
        sd_o = {k: v for k, v in sd.items()}
    ref_outs, ref_loss, ref_grads, sink = _oracle(name, sd, x, y)
    outs, loss, grads, m = _product(name, sd, x, y)
    if name == "latent2mask":      # frozen encoder: only decoder ("1.*") gradients exist
        assert all(grads[k] is None for k in grads if k.startswith("0."))
        ref_grads = {k: v for k, v in ref_grads.items() if k.startswith("1.")}
    _compare(outs, ref_outs, grads, ref_grads)
    assert abs(loss - ref_loss) <= 5e-3
    msd = m.state_dict()
    for k, r in sink.items():
        assert torch.allclose(msd[k].cpu().float(), r.float(), atol=5e-3, rtol=2e-2), k


@pytest.mark.parametrize("name", ["unet", "fct"])
def test_full_resolution_256(name):
    """BASELINE.json frame size (256x256), B=1 so the CPU oracle finishes in seconds."""
    sd = nets.init_state_dict(name, seed=7, jitter=0.1)
    g = torch.Generator().manual_seed(2)
    x = torch.rand(1, 3, 256, 256, generator=g)
    y = (torch.rand(1, 1, 256, 256, generator=g) > 0.5).float()
    ref_outs, ref_loss, ref_grads, _ = _oracle(name, sd, x, y)
    outs, loss, grads, _ = _product(name, sd, x, y)
    _compare(outs, ref_outs, grads, ref_grads)
    # mask IoU on a threshold the masks straddle (sigmoid outputs sit near 0.5 at init, H6)
    thr = float(ref_outs[-1].median())
    margin = (ref_outs[-1] - thr).abs() > 2e-2          # pixels not within bf16 noise of the threshold
    assert iou(outs[-1][margin], ref_outs[-1][margin], thr) >= 0.999


def test_eval_mode_inference_matches_oracle():
    """Inference path (config 4): eval-mode BN uses running statistics, no grad graph."""
    name = "latent2mask"
    sd = nets.init_state_dict(name, seed=9, jitter=0.2)
    for k in sd:
        if k.endswith("running_var"):
            sd[k] = sd[k] + 0.3
    x = torch.rand(2, 3, 64, 64, generator=torch.Generator().manual_seed(3))
    ref = nets.FORWARDS[name](sd, x, nets.Ctx(training=False))
    import hybridnet
    m = hybridnet.LatentToMask()
    m.load_state_dict(sd, strict=True)
    m = m.cuda().eval()
    with torch.no_grad():
        out = m(x.cuda()).cpu()
    assert max_abs(out, ref) <= OUT_ABS and rel_l2(out, ref) <= OUT_REL


def test_linearity_of_weight_gradient_at_full_batch():
    """Size-independent property at BASELINE batch (B=32, 256^2 would need the CPU oracle for minutes):
    the weight gradient of a conv layer is additive over the batch: dW(B=8) == sum of dW on the two halves."""
    from hybridnet import ops
    torch.manual_seed(0)
    x = torch.rand(8, 16, 64, 64, device="cuda")
    w = (torch.rand(16, 16, 3, 3, device="cuda") - 0.5).requires_grad_(True)
    gy = torch.rand(8, 16, 64, 64, device="cuda") - 0.5

    def dW(xs, gs):
        w.grad = None
        out = ops.to_nchw(ops.conv2d(ops.to_nhwc(xs), w, None, act="none"), 16)
        out.backward(gs)
        return w.grad.clone()
    full = dW(x, gy)
    halves = dW(x[:4], gy[:4]) + dW(x[4:], gy[4:])
    assert rel_l2(full, halves) < 1e-4
