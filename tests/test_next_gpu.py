"""GPU parity of the "next" rows either side of the hot path:
N1 fused AdamW (vs torch.optim.AdamW, state_dict both ways), N3 loss kernels (vs the reference's formulas,
Metrics.py:14-39 / F.mse_loss), N4 device input pipeline (vs the reference's torch ops, Model.py:42-53), plus the
regression tests of the round-1 advisor findings (fold cache, gradient sink, pack cache)."""
import random

import pytest
import torch

from tests.util import max_abs, rel_l2

pytestmark = pytest.mark.gpu


# ---- N3 ---------------------------------------------------------------------------------------------------
def _ref_dice(p, t, smooth=1.0):          # Metrics.py:14-22
    p, t = p[:, 0].contiguous().view(-1), t[:, 0].contiguous().view(-1)
    return 1.0 - (2.0 * (p * t).sum() + smooth) / (p.sum() + t.sum() + smooth)


@pytest.mark.parametrize("B,C,S", [(2, 1, 64), (32, 1, 256), (3, 2, 32)])
def test_dice_loss_matches_reference_formula(B, C, S):
    import hybridnet
    g = torch.Generator().manual_seed(0)
    p = torch.rand(B, C, S, S, generator=g).cuda().requires_grad_(True)
    t = (torch.rand(B, C, S, S, generator=g) > 0.5).float().cuda()
    loss = hybridnet.DiceLoss()(p, t)
    assert loss.dim() == 0 and isinstance(loss.item(), float)       # `loss.item()` as in FCT.py:334
    (loss * 3.0).backward()
    pr = p.detach().double().requires_grad_(True)
    ref = _ref_dice(pr, t.double())
    (ref * 3.0).backward()
    assert abs(float(loss) - float(ref)) < 2e-6
    assert rel_l2(p.grad, pr.grad) < 1e-5
    # a second call reuses the self-cleaning workspace
    loss2 = hybridnet.DiceLoss()(p.detach(), t)
    assert abs(float(loss2) - float(ref)) < 2e-6


def test_mse_loss_and_psnr_match_torch():
    import hybridnet
    g = torch.Generator().manual_seed(1)
    a = torch.rand(4, 3, 64, 64, generator=g).cuda().requires_grad_(True)
    b = torch.rand(4, 3, 64, 64, generator=g).cuda()
    crit = hybridnet.MSELoss()
    loss = crit(a, b)
    loss.backward()
    ar = a.detach().double().requires_grad_(True)
    ref = torch.nn.functional.mse_loss(ar, b.double())
    ref.backward()
    assert abs(float(loss) - float(ref)) < 1e-6
    assert rel_l2(a.grad, ar.grad) < 1e-5
    psnr = 20 * torch.log10(1.0 / torch.sqrt(ref))                   # Model.py:30
    assert abs(float(crit.last_psnr) - float(psnr)) < 1e-3


def test_jaccard_score_has_reference_semantics():
    import hybridnet
    g = torch.Generator().manual_seed(2)
    p = torch.rand(2, 1, 64, 64, generator=g)
    p[p < 0.3] = 0.0                       # the reference counts every NON-ZERO value as True (Metrics.py:35-36)
    t = (torch.rand(2, 1, 64, 64, generator=g) > 0.5).float()
    out = hybridnet.JaccardScore()(p.cuda(), t.cuda())
    yp, yt = p[:, 0].contiguous().view(-1), t[:, 0].contiguous().view(-1)
    ref = torch.sum(torch.logical_and(yt, yp)) / torch.sum(torch.logical_or(yt, yp))
    assert torch.is_tensor(out) and abs(out.item() - ref.item()) < 1e-6


# ---- N1 ---------------------------------------------------------------------------------------------------
def _toy():
    import hybridnet
    torch.manual_seed(0)
    return hybridnet.UNet(3, 1, init_features=8).cuda().train()


def test_fused_adamw_matches_torch_and_exchanges_state_dicts():
    import hybridnet
    from hybridnet import parallel
    m = _toy()
    names = [n for n, _ in m.named_parameters()]
    ref_params = [p.detach().clone().requires_grad_(True) for p in m.parameters()]
    arena = parallel.GradArena(m)
    opt = hybridnet.optim.AdamW(m.parameters(), arena, lr=1e-3)
    ref = torch.optim.AdamW(ref_params, lr=1e-3)
    assert all(p.data_ptr() >= opt.flat.data_ptr() for p in arena.params)       # parameters live in the flat buffer
    g = torch.Generator().manual_seed(3)
    for step in range(4):
        opt.zero_grad()
        for p, r in zip(m.parameters(), ref_params):
            gr = torch.randn(p.shape, generator=g).cuda() * 0.1
            p.grad.copy_(gr)                                   # views of the arena
            r.grad = gr.clone()
        opt.step()
        ref.step()
        for n, p, r in zip(names, m.parameters(), ref_params):
            assert max_abs(p, r) < 2e-6, (step, n)
    assert opt.steps_done() == 4
    # our state_dict -> torch.optim.AdamW
    sd = opt.state_dict()
    tsd = ref.state_dict()
    assert set(sd["param_groups"][0]) == set(tsd["param_groups"][0])
    assert sd["param_groups"][0]["params"] == tsd["param_groups"][0]["params"]
    assert set(sd["state"]) == set(tsd["state"])
    for i in sd["state"]:
        assert float(sd["state"][i]["step"]) == float(tsd["state"][i]["step"]) == 4.0
        assert max_abs(sd["state"][i]["exp_avg"], tsd["state"][i]["exp_avg"]) < 1e-6
        assert max_abs(sd["state"][i]["exp_avg_sq"], tsd["state"][i]["exp_avg_sq"]) < 1e-6
    ref2_params = [p.detach().clone().requires_grad_(True) for p in m.parameters()]
    ref2 = torch.optim.AdamW(ref2_params, lr=1e-3)
    ref2.load_state_dict(sd)
    # torch's state_dict -> ours (fresh model/optimizer), then both continue identically
    m2 = _toy()
    m2.load_state_dict(m.state_dict())
    arena2 = parallel.GradArena(m2)
    opt2 = hybridnet.optim.AdamW(m2.parameters(), arena2, lr=1e-3)
    opt2.load_state_dict(tsd)
    assert opt2.steps_done() == 4
    opt2.zero_grad()
    for p, r in zip(m2.parameters(), ref2_params):
        gr = torch.randn(p.shape, generator=g).cuda() * 0.1
        p.grad.copy_(gr)
        r.grad = gr.clone()
    opt2.step()
    ref2.step()
    for p, r in zip(m2.parameters(), ref2_params):
        assert max_abs(p, r) < 2e-6


def test_fused_adamw_leaves_never_used_parameters_bit_identical_and_trains_under_graph():
    """SURVEY note U: FCT's 14 never-used tensors get no weight decay and no update; the optimizer step inside a
    captured step replays with the device-side step count."""
    import hybridnet
    from hybridnet import ops, parallel
    torch.manual_seed(0)
    dev = torch.device("cuda")
    m = hybridnet.FCT().to(dev).train()
    dead = parallel.used_parameter_names("fct")
    before = {n: p.detach().clone() for n, p in m.named_parameters()}
    arena = parallel.GradArena(m, used={n for n, _ in m.named_parameters() if n not in dead}).direct()
    try:
        opt = hybridnet.optim.AdamW(m.parameters(), arena, lr=1e-3)
        crit = hybridnet.DiceLoss()
        x = torch.rand(2, 3, 64, 64, device=dev)
        y = (torch.rand(2, 1, 64, 64, device=dev) > 0.5).float()
        cap_floats = sum(p.numel() for p in m.parameters()) + (1 << 16)

        def step(xd, yd):
            opt.zero_grad()
            ops.begin_step(dev, cap_floats)
            loss = crit(m(xd), yd)
            loss.backward()
            arena.finish()
            opt.step()
            return loss

        l0 = float(step(x, y))
        cap = parallel.CapturedStep(step, x, y, warmup=1)
        n0 = opt.steps_done()
        losses = [float(cap()) for _ in range(6)]
        assert opt.steps_done() == n0 + 6                       # the graph bumps the device-side step itself
        assert losses[-1] < l0                                  # it trains
        for n, p in m.named_parameters():
            if n in dead:
                assert torch.equal(p.detach(), before[n]), n
            else:
                assert not torch.equal(p.detach(), before[n]), n
    finally:
        arena.direct(False)
        ops.end_zero_pool()


# ---- N4 ---------------------------------------------------------------------------------------------------
def test_augment_pack_matches_reference_ops():
    from hybridnet import data
    from tests.util import nchw_f32
    g = torch.Generator().manual_seed(4)
    u8 = torch.randint(0, 256, (3, 64, 48, 3), generator=g, dtype=torch.uint8)
    image = (u8.float() / 255.0).permute(0, 3, 1, 2).contiguous()           # ToTensor(), Dataloader.py:19-23
    # (a) plain pack of uint8 frames
    packed, nchw = data.augment_pack(u8.cuda(), want_nchw=True)
    assert torch.equal(nchw.cpu(), image)
    assert max_abs(nchw_f32(packed.nhwc.cpu(), 3), image.to(torch.bfloat16).float()) == 0.0
    assert float(packed.nhwc[..., 3:].abs().max()) == 0.0
    # (b) hflip (Model.py:44-45)
    _, nchw = data.augment_pack(u8.cuda(), flip=True, want_nchw=True)
    assert torch.equal(nchw.cpu(), torch.flip(image, dims=[3]))
    # (c) blackout patches without noise (the patch arithmetic of Model.py:50-53), fp32 input
    patches = [(5, 7), (40, 30)]
    _, nchw = data.augment_pack(image.cuda(), patches=patches, want_nchw=True)
    ref = image.clone()
    for x, y in patches:
        ref[:, :, x:x + 16, y:y + 16] = 0.0
    assert torch.equal(nchw.cpu(), ref)
    # (d) the noise branch starts from the UNFLIPPED image (Model.py:46-47) and adds N(0, 0.05^2)
    big = torch.rand(4, 3, 128, 128, generator=g)
    _, nchw = data.augment_pack(big.cuda(), flip=True, noise_std=0.05, seed=7, want_nchw=True)
    noise = nchw.cpu() - big
    assert abs(float(noise.mean())) < 2e-3 and abs(float(noise.std()) - 0.05) < 2e-3
    k = float(((noise / 0.05) ** 4).mean())                                  # Gaussian kurtosis = 3
    assert 2.7 < k < 3.3, k
    _, again = data.augment_pack(big.cuda(), flip=True, noise_std=0.05, seed=8, want_nchw=True)
    assert float((again.cpu() - nchw.cpu()).abs().max()) > 1e-3               # another seed, another draw


def test_augmentation_decisions_follow_the_reference_order():
    """draw_augmentation consumes python `random` exactly like Model.py:44-53, so a seeded run branches the same."""
    from hybridnet import data
    for seed in range(40):
        random.seed(seed)
        flip, std, patches = data.draw_augmentation(256, 256)
        random.seed(seed)
        r_flip = random.random() > 0.5
        r_noise, r_patches = False, []
        if random.random() > 0.8:
            r_noise = True
            if random.random() > 0.5:
                for _ in range(random.randint(0, 3)):
                    x = random.randint(0, 256 - 16)
                    y = random.randint(0, 256 - 16)
                    r_patches.append((x, y))
        assert (flip, std > 0, patches) == (r_flip, r_noise, r_patches)


def test_models_accept_packed_frames():
    import hybridnet
    from hybridnet import data
    torch.manual_seed(0)
    m = hybridnet.UNet(3, 1, init_features=8).cuda().eval()
    x = torch.rand(2, 3, 64, 64)
    with torch.no_grad():
        a = m(x.cuda())
        b = m(data.augment_pack(x.cuda()))
    assert torch.equal(a, b)


def test_avgpool_pyramid_is_differentiable():
    """nn.AvgPool2d propagates gradient to the frames (FCT.py:238-240); round 1 detached them."""
    from hybridnet import ops
    x = torch.rand(2, 8, 16, 16).cuda().requires_grad_(True)
    y = ops.to_nchw(ops.avgpool2(ops.to_nhwc(x)), 8)
    gy = torch.rand(2, 8, 8, 8).cuda()
    y.backward(gy)
    xr = x.detach().to(torch.bfloat16).float().requires_grad_(True)
    yr = torch.nn.functional.avg_pool2d(xr, 2)
    yr.backward(gy.to(torch.bfloat16).float())
    assert rel_l2(y, yr) < 4e-3 and rel_l2(x.grad, xr.grad) < 4e-3


# ---- advisor findings (round 1) -----------------------------------------------------------------------------
def test_eval_bn_fold_follows_training_mode_updates_without_optimizer():
    """LatentToMask's frozen encoder still updates its running statistics in train mode; an eval forward after it
    must not reuse BatchNorm folds made from the old statistics."""
    import hybridnet
    from oracle import nets
    sd = nets.init_state_dict("latent2mask", seed=9, jitter=0.2)
    x = torch.rand(2, 3, 64, 64, generator=torch.Generator().manual_seed(3))
    m = hybridnet.LatentToMask()
    m.load_state_dict(sd, strict=True)
    m = m.cuda()

    def eval_out():
        m.eval()
        with torch.no_grad():
            return m(x.cuda()).cpu()

    first = eval_out()                          # builds the folds
    m.train()
    for _ in range(3):
        m(x.cuda())                             # no optimizer step: only the running statistics move
    second = eval_out()
    ref = nets.FORWARDS["latent2mask"]({k: v.cpu() for k, v in m.state_dict().items()}, x, nets.Ctx(training=False))
    assert max_abs(second, first) > 1e-3        # the statistics did change the output
    assert max_abs(second, ref) <= 5e-2 and rel_l2(second, ref) <= 2e-2


def test_gradient_sink_two_backwards_per_zero_grad_accumulate():
    import hybridnet
    from hybridnet import parallel
    torch.manual_seed(0)
    m = hybridnet.UNet(3, 1, init_features=8).cuda().eval()         # eval BN: deterministic
    arena = parallel.GradArena(m)
    xs = [torch.rand(2, 3, 32, 32).cuda() for _ in range(2)]

    def two_backwards():
        arena.zero_grad()
        for x in xs:
            m(x).mean().backward()
        return arena.arena.clone()

    plain = two_backwards()
    try:
        arena.direct(True)
        direct = two_backwards()
    finally:
        arena.direct(False)
    assert rel_l2(direct, plain) < 1e-3


def test_weight_gradients_on_the_side_stream_match_the_serial_order():
    """Sunk weight gradients run on a side stream and rejoin once, in an end-of-backward engine callback
    (ops._conv_backward): the arena must hold the same gradients as with everything on one stream, the side stream must
    be joined when backward() returns, and a backward that raises half-way must not leave it dangling."""
    import hybridnet
    from hybridnet import ops, parallel
    torch.manual_seed(0)
    # eval-mode BatchNorm: with batch statistics two runs of the SAME schedule already differ by percents at random
    # initialisation (atomic summation order -> statistics -> bf16 rounding flips through 18 BatchNorm layers)
    m = hybridnet.UNet(3, 1, init_features=32).cuda().eval()
    arena = parallel.GradArena(m).direct(True)
    x, y = torch.rand(4, 3, 128, 128).cuda(), (torch.rand(4, 1, 128, 128) > 0.5).float().cuda()
    crit = hybridnet.DiceLoss()

    def grads():
        arena.zero_grad()
        ops.begin_step(x.device, sum(p.numel() for p in m.parameters()) + (1 << 16))
        crit(m(x), y).backward()
        assert ops.deferred_side_stream() is None          # joined by the callback
        torch.cuda.synchronize()
        return arena.arena.clone()

    try:
        assert ops.WGRAD_OVERLAP and ops.WGRAD_DEFER
        side = grads()
        again = grads()
        ops.WGRAD_OVERLAP = False
        serial = grads()
        ops.WGRAD_OVERLAP = True
        assert rel_l2(side, serial) < 1e-3 and rel_l2(again, serial) < 1e-3

        class Boom(torch.autograd.Function):
            @staticmethod
            def forward(ctx, t):
                return t.clone()

            @staticmethod
            def backward(ctx, g):
                raise RuntimeError("boom")
        arena.zero_grad()
        with pytest.raises(RuntimeError, match="boom"):
            # the head's weight gradient is deferred first, then the failing node stops the pass
            crit(m.conv(Boom.apply(torch.rand(4, 32, 128, 128, device="cuda", requires_grad=True))).sigmoid(), y).backward()
        arena.zero_grad()                                    # start of the next step: rejoins whatever was left
        assert ops.deferred_side_stream() is None
        assert rel_l2(grads(), serial) < 1e-3
    finally:
        ops.WGRAD_OVERLAP = True
        arena.direct(False)
        ops.end_zero_pool()


def test_pack_cache_keeps_buffers_across_weight_updates():
    from hybridnet import ops
    w = (torch.rand(64, 64, 3, 3).cuda() - 0.5).requires_grad_(True)
    buf, _ = ops._packed(w, "fwd", 64)
    addr = buf.data_ptr()
    before = buf.clone()
    with torch.no_grad():
        w.mul_(2.0)                              # bumps _version
    buf2, _ = ops._packed(w, "fwd", 64)
    assert buf2.data_ptr() == addr               # refilled in place: a captured graph may hold this address
    assert rel_l2(buf2.float(), 2.0 * before.float()) < 1e-2
    ops.invalidate_packed()
    buf3, _ = ops._packed(w, "fwd", 64)
    assert buf3.data_ptr() == addr
