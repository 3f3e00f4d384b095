"""Generate golden vectors by running the UNMODIFIED reference classes.

Run ONCE in the build container (where /root/reference exists):

    python tests/golden/make_golden.py

It imports UNet.py / AutoEncoder.py / AE_Unet_Combined.py from the reference
checkout and exec's FCT.py lines 1-255 with stub modules (plain ``import FCT``
needs torchsummary/tensorboardX and would start training, FCT.py:410-412),
loads the deterministic weights of ``oracle.nets.init_state_dict`` with
``strict=True`` (this is also the state_dict key/shape check), runs
forward + backward in fp32 on CPU and stores

  * tests/golden/<model>.pt       small-resolution input, output(s), loss,
                                  per-parameter grad L2 norms, a few full grads
  * tests/golden/state_dict_keys.json   reference key -> shape for every model

The committed fixtures are what ``tests/test_oracle_golden.py`` pins the oracle
against; nothing at test/bench time reads /root/reference.
"""
import contextlib
import io
import json
import os
import sys
import types

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = os.environ.get("HN_REFERENCE", "/root/reference")
sys.path.insert(0, ROOT)
sys.path.insert(0, REF)

from oracle import nets  # noqa: E402


def load_reference():
    for name in ("torchsummary", "tensorboardX", "dataset", "metric", "pytorch_msssim"):
        m = types.ModuleType(name)
        m.summary = m.SummaryWriter = m.DataLoader = m.DiceLoss = m.JaccardScore = object
        m.ms_ssim = None
        sys.modules.setdefault(name, m)
    with contextlib.redirect_stdout(io.StringIO()):
        import UNet as ref_unet
        import AutoEncoder as ref_ae
        import AE_Unet_Combined as ref_comb
        import Metrics as ref_metrics
    src = "\n".join(open(os.path.join(REF, "FCT.py")).read().split("\n")[:255])
    ns = {}
    exec(compile(src, "FCT.py", "exec"), ns)
    return {
        "unet": ref_unet.UNet, "autoencoder": ref_ae.AutoEncoder,
        "i2i2m": ref_comb.Image2Image2Mask, "fct": ns["FCT"],
    }, ref_metrics.DiceLoss


def main():
    torch.set_num_threads(8)
    classes, DiceLoss = load_reference()
    keys = {}
    cases = {"unet": (2, 64), "autoencoder": (2, 64), "i2i2m": (1, 64), "fct": (1, 64)}
    for name, cls in classes.items():
        torch.manual_seed(0)
        model = cls()
        ref_sd = model.state_dict()
        keys[name] = {k: list(v.shape) for k, v in ref_sd.items()}
        sd = nets.init_state_dict(name, seed=11, jitter=0.25)
        assert list(sd.keys()) == list(ref_sd.keys()), name
        model.load_state_dict(sd, strict=True)
        model.train()
        for m in model.modules():          # dropout off: RNG streams cannot match
            if isinstance(m, (torch.nn.Dropout, torch.nn.Dropout2d)):
                m.p = 0.0
        B, S = cases[name]
        g = torch.Generator().manual_seed(1)
        x = torch.rand(B, 3, S, S, generator=g)
        y = (torch.rand(B, 1, S, S, generator=g) > 0.5).float()
        out = model(x)
        outs = out if isinstance(out, tuple) else (out,)
        if name == "autoencoder":
            loss = torch.nn.functional.mse_loss(outs[1], x)
        else:
            loss = DiceLoss()(outs[-1], y)
        loss.backward()
        gn = {k: (float(p.grad.norm()) if p.grad is not None else None)
              for k, p in model.named_parameters()}
        named = dict(model.named_parameters())
        small = sorted((k for k, p in named.items() if p.grad is not None),
                       key=lambda k: named[k].numel())
        full = {k: named[k].grad.clone() for k in small[:12] + small[-2:]
                if named[k].numel() <= 40000}
        bn_after = {k: v.clone() for k, v in model.state_dict().items()
                    if "running_" in k or "num_batches" in k}
        torch.save({"model": name, "seed": 11, "jitter": 0.25, "x": x, "y": y,
                    "outputs": [o.detach() for o in outs], "loss": float(loss),
                    "grad_norms": gn, "grads": full,
                    "bn_after": {k: v for k, v in list(bn_after.items())[:12]}},
                   os.path.join(HERE, name + ".pt"))
        print(name, "loss", float(loss), "outs", [tuple(o.shape) for o in outs])
    # Autoencoder32K: no reference source; record our restated key list so a
    # change is visible (counts pinned by SURVEY.md App. A)
    for name in ("ae32k_image", "latent2mask"):
        sd = nets.init_state_dict(name, 0)
        keys[name] = {k: list(v.shape) for k, v in sd.items()}
    with open(os.path.join(HERE, "state_dict_keys.json"), "w") as f:
        json.dump(keys, f, indent=0)
    make_trained(classes)


def make_trained(classes):
    """Pin of the trained-weights recipe (tests/golden/trained.py): the recipe run through the REAL reference
    classes; for the bytecode-only Autoencoder32K family through the oracle (unpinned)."""
    from tests.golden import trained
    for name, cls in classes.items():
        model = cls()
        model.load_state_dict(nets.init_state_dict(name, trained.INIT_SEED.get(name, 0)), strict=True)
        losses, _ = trained.train_module(model, name)
        sd = {k: v.detach().clone() for k, v in model.state_dict().items()}
        x, _ = trained.blobs(2, trained.S, trained.EVAL_SEED)
        with torch.no_grad():
            out = model(x)
        out = out[-1] if isinstance(out, tuple) else out
        fx = trained.summary(sd, losses, out)
        fx["source"] = "reference classes"
        torch.save(fx, os.path.join(HERE, f"trained_{name}.pt"))
        print("trained", name, [round(l, 4) for l in losses[::5]], "frac>0.5", float((out > 0.5).float().mean()))
    for name in ("ae32k_image", "latent2mask"):
        sd, losses = trained.train_oracle(name)
        x, _ = trained.blobs(2, trained.S, trained.EVAL_SEED)
        with torch.no_grad():
            out = nets.FORWARDS[name](sd, x, nets.Ctx(training=True, dropout="off"))
        fx = trained.summary(sd, losses, out)
        fx["source"] = "oracle/nets.py (no reference source for this family: parity unpinned)"
        torch.save(fx, os.path.join(HERE, f"trained_{name}.pt"))
        print("trained", name, [round(l, 4) for l in losses[::5]])


if __name__ == "__main__":
    main()
