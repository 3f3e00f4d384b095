"""Trained-weights fixture recipe ("trained-for-20-steps weight set").

Random-init weights make two things degenerate: sigmoid masks sit at 0.485..0.497, so mask IoU at
threshold 0.5 says nothing, and a ~37-layer train-mode BatchNorm stack (Autoencoder32K) is an expander
under which ANY bf16 arithmetic lands 30-45 % from fp32.  A few optimizer steps fix both: after
``STEPS[model]`` (20-60) AdamW steps on the synthetic blob dataset below the masks straddle 0.5 and an ideal-bf16
emulation of the oracle stays within 0.7 % of fp32 on Autoencoder32K's output.

The weights are NOT committed (UNet alone is 31 MB): they are re-derived from this deterministic recipe
-- ``init_state_dict(name, INIT_SEED)`` -> ``STEPS[model]`` x (blobs(B, S, SEED0 + i) -> forward -> loss -> AdamW) with
dropout off -- by the oracle (``oracle/nets.py``), on the CPU (``-m "not gpu"`` tests) or on the GPU in
fp32 with TF32 off (``-m gpu`` tests, as the checker only).  What IS committed is the pin:
``tests/golden/trained_<model>.pt`` holds the loss trajectory, the per-tensor L2 norms of the trained
weights and the trained net's output on a fixed batch, produced by running the SAME recipe through the
unmodified reference classes (``make_golden.py``; FCT / UNet / AutoEncoder / Image2Image2Mask) -- for the
bytecode-only Autoencoder32K family the pin comes from the oracle itself (parity unpinned, as everywhere).
"""
import torch

S, B, LR, SEED0, EVAL_SEED = 64, 4, 1e-3, 100, 999
# FCT and Image2Image2Mask learn slowly: at 20 steps 97-100 % of their mask pixels are still above 0.5
STEPS = {"fct": 60, "i2i2m": 60, "unet": 20, "autoencoder": 20, "ae32k_image": 20, "ae32k_mask": 20, "latent2mask": 20}
# init_state_dict seed: with seed 0 the AutoEncoder's ReLU head (AutoEncoder.py:67) is dead on all three output
# channels (every pre-activation negative -> zero output, zero gradient, nothing ever trains); seed 10 is the first
# seed with all three channels alive.  Image2Image2Mask contains that AutoEncoder.
INIT_SEED = {"autoencoder": 10, "i2i2m": 10}
RECONSTRUCTION = ("autoencoder", "ae32k_image")
_trained_cache = {}


def blobs(B, S, seed, device="cpu"):
    """Frames with two bright discs on a dim noisy background; mask = the discs.  Generated on the CPU
    from one torch.Generator so that every device sees identical data."""
    g = torch.Generator().manual_seed(seed)
    yy, xx = torch.meshgrid(torch.arange(S).float(), torch.arange(S).float(), indexing="ij")
    x = torch.rand(B, 3, S, S, generator=g) * 0.35
    y = torch.zeros(B, 1, S, S)
    for b in range(B):
        for _ in range(2):
            cx, cy = (torch.rand(2, generator=g) * S).tolist()
            r = float(torch.rand(1, generator=g)) * S * 0.22 + S * 0.08
            m = ((xx - cx) ** 2 + (yy - cy) ** 2) < r * r
            y[b, 0][m] = 1.0
            col = torch.rand(3, generator=g) * 0.5 + 0.5
            for c in range(3):
                x[b, c][m] = x[b, c][m] * 0.3 + col[c]
    return x.clamp(0, 1).to(device), y.to(device)


def loss_of(name, out, x, y):
    """MSE for the reconstruction nets (Model.py:57 without the MS-SSIM term), DiceLoss (Metrics.py:14-22)
    on the last output for the mask nets."""
    out = out[-1] if isinstance(out, tuple) else out
    if name in RECONSTRUCTION:
        return torch.nn.functional.mse_loss(out, x)
    p, t = out[:, 0].reshape(-1), y[:, 0].reshape(-1)
    return 1.0 - (2.0 * (p * t).sum() + 1.0) / (p.sum() + t.sum() + 1.0)


def train_oracle(name, device="cpu", steps=None):
    """The recipe through oracle/nets.py.  Returns (trained state_dict on `device`, loss trajectory)."""
    from oracle import nets
    sd = nets.init_state_dict(name, INIT_SEED.get(name, 0))
    if name == "latent2mask":
        # pyc:AE_32K:L232-245: the frozen encoder IS the trained image autoencoder's encoder
        # (saved_model/EncoderDecoder_road_image2image32K.tar), not a random one
        key = ("ae32k_image", str(device))
        if key not in _trained_cache:
            _trained_cache[key] = train_oracle("ae32k_image", device)[0]
        for k, v in _trained_cache[key].items():
            if k.startswith("encoder."):
                sd["0." + k[len("encoder."):]] = v.detach().cpu().clone()
    frozen = (lambda k: k.startswith("0.")) if name == "latent2mask" else (lambda k: False)
    params = {k: (v.to(device).clone().requires_grad_(True)
                  if v.is_floating_point() and "running" not in k and not frozen(k) else v.to(device).clone())
              for k, v in sd.items()}
    opt = torch.optim.AdamW([p for p in params.values() if p.requires_grad], lr=LR)
    losses = []
    for i in range(steps or STEPS[name]):
        x, y = blobs(B, S, SEED0 + i, device)
        sink = {}
        out = nets.FORWARDS[name](params, x, nets.Ctx(training=True, dropout="off", bn_sink=sink))
        loss = loss_of(name, out, x, y)
        opt.zero_grad()
        loss.backward()
        opt.step()
        params.update(sink)                      # BatchNorm running statistics / num_batches_tracked
        losses.append(float(loss.detach()))
    return {k: v.detach() for k, v in params.items()}, losses


def train_module(model, name, steps=None):
    """The same recipe through an nn.Module (the reference classes in make_golden.py)."""
    model.train()
    for m in model.modules():
        if isinstance(m, (torch.nn.Dropout, torch.nn.Dropout2d)):
            m.p = 0.0
    opt = torch.optim.AdamW(model.parameters(), lr=LR)
    losses = []
    for i in range(steps or STEPS[name]):
        x, y = blobs(B, S, SEED0 + i)
        loss = loss_of(name, model(x), x, y)
        opt.zero_grad()
        loss.backward()
        opt.step()
        losses.append(float(loss.detach()))
    return losses, opt


def summary(sd, losses, eval_out):
    return {"losses": [float(l) for l in losses],
            "norms": {k: float(v.float().norm()) for k, v in sd.items() if v.is_floating_point()},
            "eval_out": eval_out.detach().cpu().float(),
            "recipe": {"steps": len(losses), "S": S, "B": B, "lr": LR, "seed0": SEED0, "eval_seed": EVAL_SEED}}
