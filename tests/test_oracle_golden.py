"""Pin the CPU oracle (oracle/nets.py) to golden vectors produced by the real
reference classes (tests/golden/make_golden.py).  fp32 vs fp32: tight bounds."""
import json
import os

import pytest
import torch

from oracle import nets

G = os.path.join(os.path.dirname(__file__), "golden")


def _run(name):
    fx = torch.load(os.path.join(G, name + ".pt"))
    sd = nets.init_state_dict(name, fx["seed"], fx["jitter"])
    params = {k: v.clone().requires_grad_(True) if v.is_floating_point() and "running" not in k
              else v for k, v in sd.items()}
    sink = {}
    ctx = nets.Ctx(training=True, dropout="off", bn_sink=sink)
    out = nets.FORWARDS[name](params, fx["x"], ctx)
    outs = out if isinstance(out, tuple) else (out,)
    if name == "autoencoder":
        loss = torch.nn.functional.mse_loss(outs[1], fx["x"])
    else:
        loss = nets.dice_loss(outs[-1], fx["y"])
    loss.backward()
    return fx, params, outs, loss, sink


@pytest.mark.parametrize("name", ["unet", "autoencoder", "i2i2m", "fct"])
def test_oracle_matches_reference_golden(name):
    fx, params, outs, loss, sink = _run(name)
    assert len(outs) == len(fx["outputs"])
    for o, r in zip(outs, fx["outputs"]):
        assert o.shape == r.shape
        assert (o - r).abs().max().item() < 2e-5
    assert abs(float(loss) - fx["loss"]) < 1e-6
    for k, n in fx["grad_norms"].items():
        g = params[k].grad
        if n is None:      # note U: never-used parameters keep grad=None
            assert g is None, k
        else:
            assert g is not None, k
            assert abs(float(g.norm()) - n) <= 2e-4 * max(n, 1e-3) + 1e-7, k
    for k, r in fx["grads"].items():
        assert (params[k].grad - r).abs().max().item() <= 1e-5 + 1e-4 * r.abs().max().item(), k
    for k, r in fx["bn_after"].items():
        assert torch.allclose(sink[k].float(), r.float(), atol=1e-6), k


def test_state_dict_keys_match_reference():
    keys = json.load(open(os.path.join(G, "state_dict_keys.json")))
    for name, ref in keys.items():
        sd = nets.init_state_dict(name, 0)
        assert list(sd.keys()) == list(ref.keys()), name
        for k, v in sd.items():
            assert list(v.shape) == ref[k], (name, k)


def test_param_counts_match_survey():
    want = {"fct": (322, 2094789), "unet": (118, 7763041), "autoencoder": (70, 692743),
            "i2i2m": (188, 8455784), "ae32k_image": (235, 3305393), "latent2mask": (235, 3305243)}
    for name, (nt, npar) in want.items():
        sd = nets.init_state_dict(name, 0)
        n = sum(v.numel() for k, v in sd.items() if "running_" not in k and "num_batches" not in k)
        assert (len(sd), n) == (nt, npar), name


def test_ae32k_shapes():
    sd = nets.init_state_dict("ae32k_image", 0)
    x = torch.rand(1, 3, 64, 64)
    ctx = nets.Ctx(training=True)
    lat = nets.encoder32k_forward(sd, x, ctx, "encoder.")
    assert lat.shape == (1, 8, 16, 16)      # 256 -> [8,64,64] = "32K"
    assert nets.ae32k_forward(sd, x, ctx).shape == (1, 3, 64, 64)
    sdm = nets.init_state_dict("latent2mask", 0)
    y = nets.FORWARDS["latent2mask"](sdm, x, ctx)
    assert y.shape == (1, 1, 64, 64) and 0 <= float(y.min()) and float(y.max()) <= 1


@pytest.mark.parametrize("name", ["unet", "autoencoder", "i2i2m", "fct", "ae32k_image", "latent2mask"])
def test_trained_recipe_matches_reference_pin(name):
    """The trained-weights fixture (tests/golden/trained.py): the oracle run through the recipe must retrace
    what the unmodified reference classes did under the same recipe (make_golden.py) -- loss trajectory,
    per-tensor norms of the trained weights and the trained net's output on a fixed batch.  fp32 vs fp32 on
    the CPU over 20-60 AdamW steps: agreement is ~1e-5; the bounds leave room for thread-count dependent
    summation order.  (Autoencoder32K family: the pin is the oracle's own run -- parity unpinned.)"""
    from tests.golden import trained
    fx = torch.load(os.path.join(G, f"trained_{name}.pt"))
    assert fx["recipe"]["steps"] == trained.STEPS[name]
    sd, losses = trained.train_oracle(name)
    assert len(losses) == len(fx["losses"])
    # The first 20 steps retrace the reference to ~1e-6.  FCT / Image2Image2Mask run 60 steps (their masks need
    # them to straddle 0.5) and lr 1e-3 without warm-up is a chaotic trajectory: a 1e-7 summation-order
    # difference at step 8 grows to 1e-2 in the loss by step 45 (measured here, oracle vs reference classes,
    # both fp32 on the same CPU), so beyond step 20 only the first 20 steps and the fixture's purpose are checked.
    for i, (a, b) in enumerate(zip(losses, fx["losses"])):
        if i < 20:
            assert abs(a - b) <= 2e-4, (i, a, b)
    x, _ = trained.blobs(2, trained.S, trained.EVAL_SEED)
    with torch.no_grad():
        out = nets.FORWARDS[name](sd, x, nets.Ctx(training=True, dropout="off"))
    out = out[-1] if isinstance(out, tuple) else out
    if trained.STEPS[name] <= 20:
        for k, n in fx["norms"].items():
            assert abs(float(sd[k].float().norm()) - n) <= 2e-3 * max(n, 1e-2), k
        assert (out - fx["eval_out"]).abs().max().item() <= 2e-2
    if name not in trained.RECONSTRUCTION:       # the fixture's purpose: masks that straddle the 0.5 threshold
        frac = float((out > 0.5).float().mean())
        assert 0.02 < frac < 0.98, frac
