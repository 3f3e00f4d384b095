"""Per-operator parity: every CUDA kernel (through the C ABI, via hybridnet.ops) against the
same op in plain PyTorch fp32.  Operands are pre-rounded to bf16 so the comparison measures
the kernel's fp32 accumulation + one bf16 output rounding.
Tolerances: forward rel-L2 <= 4e-3 (SURVEY 8(c) per-conv bound), gradients rel-L2 <= 1e-2.
"""
import math

import pytest
import torch
import torch.nn.functional as F

from tests.util import bf16r, nchw_f32, nhwc_bf16, rel_l2

pytestmark = pytest.mark.gpu
FWD_TOL, BWD_TOL = 4e-3, 1e-2


@pytest.fixture(scope="module")
def ops():
    import hybridnet
    from hybridnet import _lib
    _lib.call("hn_check_device")
    return hybridnet.ops


def _dev():
    return torch.device("cuda:0")


def _rand(*shape, seed=0, scale=1.0):
    g = torch.Generator().manual_seed(seed)
    return bf16r((torch.rand(*shape, generator=g) * 2 - 1) * scale).to(_dev())


def test_layout_roundtrip(ops):
    x = _rand(2, 3, 16, 24, seed=1)
    y = ops.to_nhwc(x)
    assert y.shape == (2, 16, 24, 8) and y.dtype == torch.bfloat16
    assert torch.equal(y[..., 3:], torch.zeros_like(y[..., 3:]))
    assert torch.equal(ops.to_nchw(y, 3), x)
    assert torch.equal(y, nhwc_bf16(x))


CONV_CASES = [
    # Cin, Cout, k, stride, pad, dil, H, act, bias
    (3, 8, 3, 1, 1, 1, 32, "relu", True),        # FCT block_1.conv1_a (RGB padded to 8)
    (8, 8, 3, 1, 1, 1, 32, "relu", True),
    (16, 16, 3, 1, 2, 2, 16, "none", True),      # Wide_Focus conv2 (dilation 2)
    (32, 32, 3, 1, 3, 3, 16, "none", True),      # Wide_Focus conv3 (dilation 3)
    (8, 1, 3, 1, 1, 1, 32, "sigmoid", True),     # DS_out conv3
    (8, 3, 3, 1, 1, 1, 16, "relu", True),        # AutoEncoder head
    (64, 128, 3, 1, 1, 1, 16, "none", False),    # UNet level 3
    (128, 64, 3, 1, 1, 1, 8, "relu", True),
    (3, 64, 7, 2, 3, 1, 32, "none", False),      # Encoder_32K conv1
    (64, 64, 3, 2, 1, 1, 16, "none", False),     # Bottleneck conv2 stride 2
    (64, 256, 1, 1, 0, 1, 16, "none", False),    # Bottleneck 1x1
    (256, 512, 1, 2, 0, 1, 16, "none", False),   # downsample 1x1 stride 2
    (192, 64, 3, 1, 1, 1, 16, "none", True),     # Decoder_32K conv5
    (32, 1, 1, 1, 0, 1, 16, "sigmoid", True),    # UNet head 1x1
    (64, 64, 3, 1, 1, 1, 64, "none", False),     # UNet level 2 at W >= 64: the row-halo form of the tcgen05 weight gradient
    (128, 96, 3, 1, 1, 1, 64, "none", True),     # two 64-channel boxes per side, Cout not a multiple of 64
    (64, 128, 3, 1, 2, 2, 128, "none", False),   # halo with dilation 2, two 64-pixel steps per image row
]


@pytest.mark.parametrize("case", CONV_CASES)
def test_conv2d_fwd_bwd(ops, case):
    Cin, Cout, k, stride, pad, dil, H, act, has_bias = case
    B = 2
    x = _rand(B, Cin, H, H, seed=2).requires_grad_(True)
    w = _rand(Cout, Cin, k, k, seed=3, scale=1.0 / math.sqrt(Cin * k * k)).requires_grad_(True)
    b = _rand(Cout, seed=4, scale=0.1).requires_grad_(True) if has_bias else None
    ref = F.conv2d(x, w, b, stride, pad, dil)
    ref = {"relu": F.relu, "sigmoid": torch.sigmoid, "none": lambda t: t}[act](ref)
    gy = _rand(*ref.shape, seed=5)
    ref.backward(gy)
    xr, wr, br = x.grad.clone(), w.grad.clone(), (b.grad.clone() if has_bias else None)
    x.grad = w.grad = None
    if has_bias:
        b.grad = None

    xn = ops.to_nhwc(x)
    y = ops.conv2d(xn, w, b, stride=stride, pad=pad, dil=dil, act=act)
    out = ops.to_nchw(y, Cout)
    assert out.shape == ref.shape
    assert rel_l2(out, ref) < FWD_TOL
    out.backward(gy)
    assert rel_l2(x.grad, xr) < BWD_TOL
    assert rel_l2(w.grad, wr) < BWD_TOL
    if has_bias:
        assert rel_l2(b.grad, br) < BWD_TOL


@pytest.mark.parametrize("cin,cout,k,H,tc", [(128, 128, 3, 32, True), (256, 512, 1, 32, True), (64, 64, 3, 16, True),
                                             (64, 96, 3, 16, False)])      # K chunk 32 with N = 256: generic kernel
def test_stride2_input_gradient_runs_on_tcgen05(ops, cin, cout, k, H, tc):
    """Bottleneck conv2 / downsample (stride 2): dx comes from ONE hn_conv_tc_fwd launch (2x2 window over dz,
    four output parities on N, pixel-shuffle store) and equals torch's conv input gradient."""
    import contextlib
    from hybridnet import _lib
    B = 2
    x = _rand(B, cin, H, H, seed=12).requires_grad_(True)
    w = _rand(cout, cin, k, k, seed=13, scale=1.0 / math.sqrt(cin * k * k)).requires_grad_(True)
    ref = F.conv2d(x, w, None, 2, k // 2)
    gy = _rand(*ref.shape, seed=15)
    ref.backward(gy)
    xr = x.grad.clone()
    x.grad = None
    y = ops.to_nchw(ops.conv2d(ops.to_nhwc(x), w, None, stride=2, pad=k // 2), cout)
    names = []

    @contextlib.contextmanager
    def rec(name, args):
        names.append((name, args[0].pshuf if name.startswith("hn_conv") else 0))
        yield

    _lib.set_profiler(rec)
    try:
        y.backward(gy)
    finally:
        _lib.set_profiler(None)
    if tc:
        assert ("hn_conv_tc_fwd", cin) in names, names
        assert not any(n == "hn_conv2d_fwd" for n, _ in names), names
    else:
        assert any(n == "hn_conv2d_fwd" for n, _ in names), names
    assert rel_l2(x.grad, xr) < BWD_TOL


def test_repack_all_refreshes_every_packed_weight(ops):
    """ops.repack_all(): one hn_pack_weight_multi launch re-derives all bf16 weight copies (forward, flipped-tap
    input-gradient, stride-2 input-gradient and transposed-conv layouts) after an in-place weight update."""
    import contextlib
    from hybridnet import _lib
    ops.clear_packed()
    x = _rand(2, 64, 16, 16, seed=21).requires_grad_(True)
    w3 = _rand(64, 64, 3, 3, seed=22, scale=0.05).requires_grad_(True)       # stride 1: fwd + dgradf
    w2 = _rand(128, 64, 3, 3, seed=23, scale=0.05).requires_grad_(True)      # stride 2: fwd + s2dgrad
    wt = _rand(128, 32, 2, 2, seed=24, scale=0.05).requires_grad_(True)      # ConvTranspose k2s2: tfwd + tdgrad

    def run():
        for t in (x, w3, w2, wt):
            t.grad = None
        h = ops.conv2d(ops.to_nhwc(x), w3, None)
        h = ops.conv2d(h, w2, None, stride=2, pad=1)
        y = ops.to_nchw(ops.conv2d(h, wt, None, stride=2, pad=0, transposed=True), 32)
        y.backward(torch.ones_like(y))
        return y.detach().clone(), x.grad.clone()

    y1, g1 = run()
    with torch.no_grad():
        for t in (w3, w2, wt):
            t.mul_(2.0)
    names = []

    @contextlib.contextmanager
    def rec(name, args):
        names.append(name)
        yield

    _lib.set_profiler(rec)
    try:
        n = ops.repack_all()
        y2, g2 = run()
    finally:
        _lib.set_profiler(None)
    assert n == 6, n
    assert names.count("hn_pack_weight_multi") == 1 and not any(k.startswith("hn_pack_weight") and k != "hn_pack_weight_multi" for k in names), names
    assert rel_l2(y2, 8.0 * y1) < 1e-6 and rel_l2(g2, 8.0 * g1) < 1e-6      # powers of two: exact in bf16
    with torch.no_grad():
        w3.mul_(0.5)
    y3, _ = run()                      # without repack_all the version check makes the layer repack itself
    assert rel_l2(y3, 4.0 * y1) < 1e-6
    ops.invalidate_packed()


def test_conv_virtual_concat_upsample_residual(ops):
    """cat([a,b]) folded into the K loop; nearest-x2 upsample folded into the gather; residual epilogue."""
    B, H = 2, 16
    a = _rand(B, 16, H, H, seed=1).requires_grad_(True)
    b2 = _rand(B, 16, H, H, seed=2).requires_grad_(True)
    w = _rand(16, 32, 3, 3, seed=3, scale=0.06).requires_grad_(True)
    bias = _rand(16, seed=4, scale=0.1).requires_grad_(True)
    res = _rand(B, 16, H, H, seed=6).requires_grad_(True)
    ref = F.conv2d(torch.cat([a, b2], 1), w, bias, 1, 1) + res
    gy = _rand(*ref.shape, seed=5)
    ref.backward(gy)
    want = [t.grad.clone() for t in (a, b2, w, bias, res)]
    for t in (a, b2, w, bias, res):
        t.grad = None
    y = ops.conv2d(ops.to_nhwc(a), w, bias, x2=ops.to_nhwc(b2), res=ops.to_nhwc(res))
    out = ops.to_nchw(y, 16)
    assert rel_l2(out, ref) < FWD_TOL
    out.backward(gy)
    for t, r in zip((a, b2, w, bias, res), want):
        assert rel_l2(t.grad, r) < BWD_TOL

    s = _rand(B, 8, H // 2, H // 2, seed=7).requires_grad_(True)
    w2 = _rand(8, 8, 3, 3, seed=8, scale=0.1).requires_grad_(True)
    ref = F.relu(F.conv2d(F.interpolate(s, scale_factor=2.0, mode="nearest"), w2, None, 1, 1))
    gy = _rand(*ref.shape, seed=9)
    ref.backward(gy)
    sr, wr = s.grad.clone(), w2.grad.clone()
    s.grad = w2.grad = None
    out = ops.to_nchw(ops.conv2d(ops.to_nhwc(s), w2, None, ups=1, act="relu"), 8)
    assert rel_l2(out, ref) < FWD_TOL
    out.backward(gy)
    assert rel_l2(s.grad, sr) < BWD_TOL and rel_l2(w2.grad, wr) < BWD_TOL


@pytest.mark.parametrize("C,k", [(8, 3), (16, 3), (32, 1), (32, 3)])
def test_small_channel_conv_residual_and_upsample(ops, C, k):
    """The small-channel kernels (hn_conv_sc_fwd_res, hn_conv_wgrad_sc) with the residual epilogue (Transformer:
    conv1(x1) + x, FCT.py:93-95) and with the nearest-x2 upsample folded into the halo gather (Block_decoder,
    FCT.py:167-172): maps with H % 16 == 0 and W % 32 == 0 take that path."""
    from hybridnet import _lib
    B, H, W = 2, 32, 64
    x = _rand(B, C, H, W, seed=1).requires_grad_(True)
    w = _rand(C, C, k, k, seed=3, scale=0.5 / math.sqrt(C * k * k)).requires_grad_(True)
    bias = _rand(C, seed=4, scale=0.1).requires_grad_(True)
    res = _rand(B, C, H, W, seed=6).requires_grad_(True)
    ref = F.conv2d(x, w, bias, 1, k // 2) + res
    gy = _rand(*ref.shape, seed=5)
    ref.backward(gy)
    want = [t.grad.clone() for t in (x, w, bias, res)]
    for t in (x, w, bias, res):
        t.grad = None
    import contextlib
    names = []

    @contextlib.contextmanager
    def record(name, args):
        names.append(name)
        yield
    _lib.set_profiler(record)
    try:
        out = ops.to_nchw(ops.conv2d(ops.to_nhwc(x), w, bias, pad=k // 2, res=ops.to_nhwc(res)), C)
        assert rel_l2(out, ref) < FWD_TOL
        out.backward(gy)
    finally:
        _lib.set_profiler(None)
    assert "hn_conv_sc_fwd_res" in names and "hn_conv2d_fwd" not in names, names
    for t, r in zip((x, w, bias, res), want):
        assert rel_l2(t.grad, r) < BWD_TOL

    s = _rand(B, C, H // 2, W // 2, seed=7).requires_grad_(True)
    w2 = _rand(C, C, 3, 3, seed=8, scale=0.5 / math.sqrt(C * 9)).requires_grad_(True)
    ref = F.relu(F.conv2d(F.interpolate(s, scale_factor=2.0, mode="nearest"), w2, None, 1, 1))
    gy = _rand(*ref.shape, seed=9)
    ref.backward(gy)
    sr, wr = s.grad.clone(), w2.grad.clone()
    s.grad = w2.grad = None
    names.clear()
    _lib.set_profiler(record)
    try:
        out = ops.to_nchw(ops.conv2d(ops.to_nhwc(s), w2, None, ups=1, act="relu"), C)
        assert rel_l2(out, ref) < FWD_TOL
        out.backward(gy)
    finally:
        _lib.set_profiler(None)
    assert "hn_conv_sc_fwd" in names and "hn_conv_wgrad_sc" in names and "hn_conv2d_wgrad" not in names, names
    assert rel_l2(s.grad, sr) < BWD_TOL and rel_l2(w2.grad, wr) < BWD_TOL


@pytest.mark.parametrize("B,H,W,Cin,Cout,dil,ups", [
    (3, 64, 96, 32, 32, 1, 0),      # 36 tiles, 9216 dW words: clusters of 8 CTAs (32 CTAs, four of them take two tiles)
    (2, 128, 128, 16, 16, 2, 0),    # 64 tiles, double-buffered, clusters
    (4, 128, 128, 8, 8, 3, 0),      # 128 tiles: plain grid (small flush), two tiles per CTA through both tile buffers
    (2, 128, 128, 16, 32, 1, 1),    # upsample folded into the halo gather (general loader path), clusters
    (6, 64, 64, 32, 64, 1, 0),      # Cout = 64: four 16-row blocks, 5 chunks per warp
])
def test_small_channel_weight_gradient_many_tiles(ops, B, H, W, Cin, Cout, dil, ups):
    """hn_conv_wgrad_sc beyond one tile per CTA: the two tile buffers, the parked accumulator slices and the cluster
    reduction through distributed shared memory (launches whose grid x dW size reaches 1e5 atomics) against torch's
    conv weight gradient; bias gradient and input gradient ride along."""
    import contextlib
    from hybridnet import _lib
    h, w_ = (H // 2, W // 2) if ups else (H, W)
    x = _rand(B, Cin, h, w_, seed=21).requires_grad_(True)
    w = _rand(Cout, Cin, 3, 3, seed=22, scale=1.0 / math.sqrt(Cin * 9)).requires_grad_(True)
    b = _rand(Cout, seed=23, scale=0.1).requires_grad_(True)
    xin = F.interpolate(x, scale_factor=2.0, mode="nearest") if ups else x
    ref = F.conv2d(xin, w, b, 1, dil, dil)
    gy = _rand(*ref.shape, seed=24)
    ref.backward(gy)
    want = [t.grad.clone() for t in (x, w, b)]
    for t in (x, w, b):
        t.grad = None
    names = []

    @contextlib.contextmanager
    def record(name, args):
        names.append(name)
        yield
    _lib.set_profiler(record)
    try:
        out = ops.to_nchw(ops.conv2d(ops.to_nhwc(x), w, b, pad=dil, dil=dil, ups=ups), Cout)
        assert rel_l2(out, ref) < FWD_TOL
        out.backward(gy)
    finally:
        _lib.set_profiler(None)
    assert "hn_conv_wgrad_sc" in names and "hn_conv2d_wgrad" not in names, names
    for t, r in zip((x, w, b), want):
        assert rel_l2(t.grad, r) < BWD_TOL
    # twice the same launch: the second accumulates onto the first (fp32 atomics: order differs, values agree)
    g1 = w.grad.clone()
    w.grad = None
    x.grad = None
    b.grad = None
    ops.to_nchw(ops.conv2d(ops.to_nhwc(x), w, b, pad=dil, dil=dil, ups=ups), Cout).backward(gy)
    assert rel_l2(w.grad, g1) < 1e-5


@pytest.mark.parametrize("cin,cout", [(64, 32), (512, 256), (256, 384)])
def test_conv_transpose_k2s2(ops, cin, cout):
    B, H = 2, 8
    x = _rand(B, cin, H, H, seed=1).requires_grad_(True)
    w = _rand(cin, cout, 2, 2, seed=2, scale=1.0 / math.sqrt(cin)).requires_grad_(True)
    b = _rand(cout, seed=3, scale=0.1).requires_grad_(True)
    ref = F.conv_transpose2d(x, w, b, stride=2)
    gy = _rand(*ref.shape, seed=4)
    ref.backward(gy)
    want = [t.grad.clone() for t in (x, w, b)]
    for t in (x, w, b):
        t.grad = None
    out = ops.to_nchw(ops.conv2d(ops.to_nhwc(x), w, b, stride=2, pad=0, transposed=True), cout)
    assert out.shape == ref.shape
    assert rel_l2(out, ref) < FWD_TOL
    out.backward(gy)
    for t, r in zip((x, w, b), want):
        assert rel_l2(t.grad, r) < BWD_TOL


@pytest.mark.parametrize("C,training,act,with_res", [(32, True, "relu", False), (256, True, "relu", True),
                                                      (192, True, "relu", False), (64, False, "relu", False),
                                                      (1, True, "sigmoid", False), (3, True, "relu", False)])
def test_conv_bn_act(ops, C, training, act, with_res):
    import torch.nn as nn
    B, H, Cin = 4, 8, 16
    x = _rand(B, Cin, H, H, seed=1).requires_grad_(True)
    w = _rand(C, Cin, 3, 3, seed=2, scale=0.1).requires_grad_(True)
    res = _rand(B, C, H, H, seed=8).requires_grad_(True) if with_res else None
    bn_ref, bn = nn.BatchNorm2d(C).to(_dev()), nn.BatchNorm2d(C).to(_dev())
    with torch.no_grad():
        for m in (bn_ref, bn):
            m.weight.copy_(_rand(C, seed=3) * 0.5 + 1.0)
            m.bias.copy_(_rand(C, seed=4) * 0.3)
            m.running_mean.copy_(_rand(C, seed=5) * 0.1)
            m.running_var.copy_(_rand(C, seed=6).abs() + 0.5)
    bn_ref.train(training)
    bn.train(training)
    z = bn_ref(F.conv2d(x, w, None, 1, 1))
    if with_res:
        z = z + res
    ref = torch.sigmoid(z) if act == "sigmoid" else F.relu(z)
    gy = _rand(*ref.shape, seed=7)
    ref.backward(gy)
    want = [x.grad.clone(), w.grad.clone(), bn_ref.weight.grad.clone(), bn_ref.bias.grad.clone()]
    res_want = res.grad.clone() if with_res else None
    x.grad = w.grad = None
    if with_res:
        res.grad = None
    y = ops.conv_bn_act(ops.to_nhwc(x), w, None, bn, res=ops.to_nhwc(res) if with_res else None, act=act)
    out = ops.to_nchw(y, C)
    assert rel_l2(out, ref) < 6e-3
    out.backward(gy)
    got = [x.grad, w.grad, bn.weight.grad, bn.bias.grad]
    # the bf16-stored pre-BN tensor can flip the sign of a normalised value near zero, i.e. a ReLU-mask
    # bit, relative to the fp32 reference: ~0.2% of elements -> a few % rel-L2 on gradients
    for g, r in zip(got, want):
        assert rel_l2(g, r) < 5e-2
    if with_res:
        assert rel_l2(res.grad, res_want) < 5e-2
    if training:
        assert torch.allclose(bn.running_mean, bn_ref.running_mean, atol=2e-3)
        assert torch.allclose(bn.running_var, bn_ref.running_var, rtol=1e-2, atol=1e-3)
        assert int(bn.num_batches_tracked) == 1


def test_maxpool_dropout_add_gelu3_layernorm(ops):
    B, C, H = 2, 16, 16
    x = _rand(B, C, H, H, seed=1).requires_grad_(True)
    ref = F.max_pool2d(x, 2)
    gy = _rand(*ref.shape, seed=2)
    ref.backward(gy)
    xr = x.grad.clone()
    x.grad = None
    out = ops.to_nchw(ops.maxpool2(ops.to_nhwc(x)), C)
    assert torch.equal(out, ref)
    out.backward(gy)
    assert torch.equal(x.grad, xr)

    # ties (post-ReLU zeros): the first maximum in scan order takes the gradient, as in PyTorch
    xt = torch.zeros(1, 8, 4, 4, device=_dev()).requires_grad_(True)
    F.max_pool2d(xt, 2).sum().backward()
    want = xt.grad.clone()
    xt.grad = None
    ops.to_nchw(ops.maxpool2(ops.to_nhwc(xt)), 8).sum().backward()
    assert torch.equal(xt.grad, want)

    # dropout: keep-rate, scaling, and backward uses the same mask
    xn = ops.to_nhwc(torch.ones(4, 8, 64, 64, device=_dev())).requires_grad_(True)
    ops.manual_seed(123)
    y = ops.dropout(xn, 0.3, True)
    keep = (y > 0).float().mean().item()
    assert abs(keep - 0.7) < 0.01
    assert torch.allclose(y[y > 0].float(), torch.tensor(1 / 0.7, device=_dev()), rtol=1e-2)
    y.sum().backward()
    assert torch.equal((xn.grad > 0), (y > 0))
    y2 = ops.dropout(xn.detach(), 0.3, True, channelwise=True)      # Dropout2d: whole (b,c) planes
    plane = (y2 > 0).float().mean(dim=(1, 2))
    assert torch.all((plane == 0) | (plane == 1))
    assert ops.dropout(xn, 0.3, False) is xn

    a = _rand(B, C, H, H, seed=3).requires_grad_(True)
    b = _rand(B, C, H, H, seed=4).requires_grad_(True)
    out = ops.to_nchw(ops.add(ops.to_nhwc(a), ops.to_nhwc(b)), C)
    assert rel_l2(out, a + b) < FWD_TOL

    zs = [_rand(B, C, H, H, seed=s).requires_grad_(True) for s in (5, 6, 7)]
    ref = sum(F.gelu(z) for z in zs)
    gy = _rand(*ref.shape, seed=8)
    ref.backward(gy)
    want = [z.grad.clone() for z in zs]
    for z in zs:
        z.grad = None
    out = ops.to_nchw(ops.gelu3_sum(*[ops.to_nhwc(z) for z in zs]), C)
    assert rel_l2(out, ref) < 6e-3
    out.backward(gy)
    for z, r in zip(zs, want):
        assert rel_l2(z.grad, r) < BWD_TOL

    for Cc in (8, 16, 32, 64, 128):
        x = _rand(B, Cc, 8, 8, seed=9).requires_grad_(True)
        g = (_rand(Cc, seed=10) * 0.5 + 1).requires_grad_(True)
        be = (_rand(Cc, seed=11) * 0.2).requires_grad_(True)
        ref = F.layer_norm(x.permute(0, 2, 3, 1), (Cc,), g, be, 1e-5).permute(0, 3, 1, 2)
        gy = _rand(*ref.shape, seed=12)
        ref.backward(gy)
        want = [t.grad.clone() for t in (x, g, be)]
        for t in (x, g, be):
            t.grad = None
        out = ops.to_nchw(ops.layernorm_c(ops.to_nhwc(x), g, be), Cc)
        assert rel_l2(out, ref) < FWD_TOL
        out.backward(gy)
        for t, r in zip((x, g, be), want):
            assert rel_l2(t.grad, r) < BWD_TOL


@pytest.mark.parametrize("C,B,H,W", [(8, 2, 12, 12), (16, 2, 12, 12), (32, 2, 12, 12), (64, 2, 12, 12), (128, 2, 12, 12),
                                     (8, 4, 40, 72), (32, 3, 20, 36)])      # several grid-stride rounds, H != W, no power of two
def test_dwconv_relu_ln_qkv(ops, C, B, H, W):
    x = _rand(B, C, H, W, seed=1).requires_grad_(True)
    # the reference's twelve separate parameter tensors: conv_{q,k,v}.weight [C,1,3,3] / .bias, layernorm_{q,k,v}.weight / .bias
    w = [_rand(C, 1, 3, 3, seed=20 + t, scale=0.4).requires_grad_(True) for t in range(3)]
    b = [_rand(C, seed=30 + t, scale=0.2).requires_grad_(True) for t in range(3)]
    g = [(_rand(C, seed=40 + t) * 0.5 + 1).requires_grad_(True) for t in range(3)]
    be = [(_rand(C, seed=50 + t) * 0.2).requires_grad_(True) for t in range(3)]
    leaves = [x] + w + b + g + be
    refs = []
    for t in range(3):
        y = F.relu(F.conv2d(x, w[t], b[t], 1, 1, 1, C))
        refs.append(F.layer_norm(y.permute(0, 2, 3, 1), (C,), g[t], be[t], 1e-5).permute(0, 3, 1, 2))
    gys = [_rand(*refs[0].shape, seed=6 + t) for t in range(3)]
    torch.autograd.backward(refs, gys)
    want = [t.grad.clone() for t in leaves]
    for t in leaves:
        t.grad = None
    outs = ops.dwconv_relu_ln_qkv(ops.to_nhwc(x), tuple(w), tuple(b), tuple(g), tuple(be))
    outs = [ops.to_nchw(o, C) for o in outs]
    for o, r in zip(outs, refs):
        assert rel_l2(o, r) < FWD_TOL
    torch.autograd.backward(outs, gys)
    for t, r in zip(leaves, want):
        assert rel_l2(t.grad, r) < 1.5e-2


@pytest.mark.parametrize("D,N,tc", [(4, 256, 0), (4, 100, 0), (8, 256, 0), (16, 192, 0), (32, 64, 0), (64, 64, 0), (64, 4, 0),
                                    (8, 1024, 0), (4, 128, 1), (4, 256, 1), (8, 256, 1), (4, 1024, 1), (8, 1024, 2), (4, 2048, 3),
                                    (4, 128, 11), (4, 256, 11), (8, 256, 11), (4, 1024, 11), (8, 1024, 12), (4, 2048, 13)])
def test_flash_attention(ops, D, N, tc):
    """softmax(QK^T/sqrt(d))V, heads=2, incl. ragged N (not a multiple of the 64-token tile).
    tc > 0: the tcgen05 / TMEM / TMA forward (hn_attn_fwd_tc); tc = 2, 3: with every 2nd / 3rd exponential pair on the
    FMA-pipe polynomial."""
    saved = (ops.ATTN_FWD_TC, ops.ATTN_FWD_TC_POLY, ops.ATTN_BWD_TC, ops.ATTN_BWD_TC_POLY)
    # tc == 0: the mma.sync kernels; tc >= 10: the tcgen05 backward (hn_attn_bwd_tc) as well
    ops.ATTN_FWD_TC, ops.ATTN_FWD_TC_POLY = tc > 0, (0 if tc % 10 <= 1 else tc % 10)
    ops.ATTN_BWD_TC, ops.ATTN_BWD_TC_POLY = tc >= 10, (0 if tc % 10 <= 1 else tc % 10)
    try:
        _flash_attention_case(ops, D, N)
    finally:
        ops.ATTN_FWD_TC, ops.ATTN_FWD_TC_POLY, ops.ATTN_BWD_TC, ops.ATTN_BWD_TC_POLY = saved


def _flash_attention_case(ops, D, N):
    B, heads = 2, 2
    C = heads * D
    q, k, v = [_rand(B, N, C, seed=s, scale=1.5).requires_grad_(True) for s in (1, 2, 3)]

    def split(t):
        return t.view(B, N, heads, D).transpose(1, 2)
    s = split(q) @ split(k).transpose(-1, -2) / math.sqrt(D)
    ref = (torch.softmax(s, -1) @ split(v)).transpose(1, 2).reshape(B, N, C)
    gy = _rand(B, N, C, seed=4)
    ref.backward(gy)
    want = [t.grad.clone() for t in (q, k, v)]
    for t in (q, k, v):
        t.grad = None

    class _Cast(torch.autograd.Function):          # fp32 [B,N,C] <-> bf16 [B,1,N,C] (test glue)
        @staticmethod
        def forward(ctx, t):
            return t.to(torch.bfloat16).view(B, 1, N, C)

        @staticmethod
        def backward(ctx, g):
            return g.float().view(B, N, C)
    o = ops.attention(_Cast.apply(q), _Cast.apply(k), _Cast.apply(v), heads)
    out = o.float().view(B, N, C)
    assert rel_l2(out, ref) < 6e-3
    out.backward(gy)
    for t, r in zip((q, k, v), want):
        assert rel_l2(t.grad, r) < 1.5e-2


def test_adamw_matches_torch(ops):
    from hybridnet import _lib
    n = 10007
    p = _rand(n, seed=1).float().contiguous()
    g = _rand(n, seed=2).float().contiguous()
    ref = torch.nn.Parameter(p.clone())
    opt = torch.optim.AdamW([ref], lr=1e-3)
    m, v = torch.zeros_like(p), torch.zeros_like(p)
    for step in (1, 2, 3):
        ref.grad = g.clone()
        opt.step()
        _lib.call("hn_adamw_step", p.data_ptr(), g.data_ptr(), m.data_ptr(), v.data_ptr(), n, 1e-3, 0.9, 0.999, 1e-8,
                  1e-2, step, 1.0, torch.cuda.current_stream().cuda_stream)
    assert torch.allclose(p, ref.detach(), atol=1e-6, rtol=1e-5)


def test_bad_arguments_raise(ops):
    with pytest.raises(RuntimeError):
        ops.attention(*[torch.zeros(1, 1, 8, 12, dtype=torch.bfloat16, device=_dev())] * 3, 2)   # D=6
    with pytest.raises(RuntimeError):
        ops.layernorm_c(torch.zeros(1, 2, 2, 24, dtype=torch.bfloat16, device=_dev()),
                        torch.ones(24, device=_dev()), torch.zeros(24, device=_dev()))


TC_CASES = [
    # Cin, Cin2, Cout, k, dil, B, H, W, stats
    (64, 0, 64, 3, 1, 2, 16, 16, False),
    (128, 0, 256, 3, 1, 1, 32, 32, True),
    (64, 64, 128, 3, 1, 2, 8, 8, False),        # virtual concat, 8x8 maps: one 128-pixel tile = 2 images
    (256, 0, 64, 1, 1, 1, 16, 128, False),      # 1x1, W = 128
    (64, 0, 192, 3, 2, 1, 64, 64, True),        # dilation 2, N tile 192
    (192, 0, 64, 3, 1, 1, 8, 256, False),       # W = 256 (two tiles per row)
]


@pytest.mark.parametrize("case", TC_CASES)
def test_conv_tcgen05_path(ops, case):
    """The tcgen05/TMEM/TMA kernel against PyTorch fp32 AND bit-for-bit-close to the mma.sync kernel."""
    from hybridnet import _lib
    Cin, Cin2, Cout, k, dil, B, H, W, stats = case
    pad = dil * (k - 1) // 2
    x = _rand(B, Cin, H, W, seed=2).requires_grad_(True)
    x2 = _rand(B, Cin2, H, W, seed=12).requires_grad_(True) if Cin2 else None
    w = _rand(Cout, Cin + Cin2, k, k, seed=3, scale=1.0 / math.sqrt((Cin + Cin2) * k * k)).requires_grad_(True)
    b = _rand(Cout, seed=4, scale=0.1).requires_grad_(True)
    xin = torch.cat([x, x2], 1) if Cin2 else x
    ref = F.relu(F.conv2d(xin, w, b, 1, pad, dil))
    gy = _rand(*ref.shape, seed=5)
    ref.backward(gy)
    want = [t.grad.clone() for t in (x, w, b)] + ([x2.grad.clone()] if Cin2 else [])
    for t in (x, w, b, x2):
        if t is not None:
            t.grad = None
    d = ops._desc(B=B, H=H, W=W, C1=Cin, C2=Cin2, ld1=Cin, ld2=Cin2, OH=H, OW=W, Cout=Cout, ldy=Cout, KH=k, KW=k,
                  stride=1, pad=pad, dil=dil, mode=0, ups=0, act=1, pshuf=0, ldr=0, Kpad=k * k * (Cin + Cin2))
    assert _lib.load().hn_conv_tc_supported(d) == 1
    assert ops.USE_TC
    calls = []
    _lib.set_profiler(lambda name, args: (calls.append(name), __import__("contextlib").nullcontext())[1])
    try:
        y = ops.conv2d(ops.to_nhwc(x), w, b, x2=ops.to_nhwc(x2) if Cin2 else None, pad=pad, dil=dil, act="relu")
        out = ops.to_nchw(y, Cout)
        out.backward(gy)
    finally:
        _lib.set_profiler(None)
    assert calls.count("hn_conv_tc_fwd") == 2, calls            # forward and (flipped-tap) input gradient
    assert calls.count("hn_conv_wgrad_tc") == 1, calls
    assert rel_l2(out, ref) < FWD_TOL
    got = [x.grad, w.grad, b.grad] + ([x2.grad] if Cin2 else [])
    for g, r in zip(got, want):
        assert rel_l2(g, r) < BWD_TOL
    # same op through the mma.sync kernel: identical up to fp32 summation order
    ops.USE_TC = False
    try:
        y2 = ops.conv2d(ops.to_nhwc(x.detach()), w.detach(), b.detach(), x2=ops.to_nhwc(x2.detach()) if Cin2 else None,
                        pad=pad, dil=dil, act="relu")
    finally:
        ops.USE_TC = True
    assert rel_l2(y.float(), y2.float()) < 2e-3
    if stats:
        s = [torch.zeros(Cout, device=_dev()) for _ in range(4)]
        wp, Kpad = ops._packed(w.detach(), "fwd", Cin + Cin2)
        z1 = ops._conv_raw(ops.to_nhwc(x.detach()), None, wp, Kpad, b.detach(), None, OH=H, OW=W, Cout=Cout, KH=k, KW=k,
                           stride=1, pad=pad, dil=dil, stats=(s[0], s[1]))
        ops.USE_TC = False
        try:
            z2 = ops._conv_raw(ops.to_nhwc(x.detach()), None, wp, Kpad, b.detach(), None, OH=H, OW=W, Cout=Cout, KH=k,
                               KW=k, stride=1, pad=pad, dil=dil, stats=(s[2], s[3]))
        finally:
            ops.USE_TC = True
        zf = z1.float()
        assert torch.allclose(s[0], zf.sum(dim=(0, 1, 2)), rtol=1e-3, atol=1e-2)
        assert torch.allclose(s[1], (zf * zf).sum(dim=(0, 1, 2)), rtol=1e-3, atol=1e-2)
        assert torch.allclose(s[0], s[2], rtol=2e-3, atol=5e-2) and torch.allclose(s[1], s[3], rtol=2e-3, atol=5e-2)


def test_conv_transpose_tcgen05(ops):
    from hybridnet import _lib
    B, H, cin, cout = 2, 16, 256, 384
    x = _rand(B, cin, H, H, seed=1).requires_grad_(True)
    w = _rand(cin, cout, 2, 2, seed=2, scale=1.0 / math.sqrt(cin)).requires_grad_(True)
    b = _rand(cout, seed=3, scale=0.1).requires_grad_(True)
    ref = F.conv_transpose2d(x, w, b, stride=2)
    gy = _rand(*ref.shape, seed=4)
    ref.backward(gy)
    want = [t.grad.clone() for t in (x, w, b)]
    for t in (x, w, b):
        t.grad = None
    calls = []
    _lib.set_profiler(lambda name, args: (calls.append(name), __import__("contextlib").nullcontext())[1])
    try:
        out = ops.to_nchw(ops.conv2d(ops.to_nhwc(x), w, b, stride=2, pad=0, transposed=True), cout)
        out.backward(gy)
    finally:
        _lib.set_profiler(None)
    assert "hn_conv_tc_fwd" in calls
    assert rel_l2(out, ref) < FWD_TOL
    for t, r in zip((x, w, b), want):
        assert rel_l2(t.grad, r) < BWD_TOL


@pytest.mark.parametrize("C,p", [(8, 0.0), (16, 0.0), (64, 0.0), (8, 0.1), (32, 0.1)])
def test_wide_focus_fused_glue(ops, C, p):
    """hn_wf_* (csrc/hn_wide_focus.cu): sum_i dropout(gelu(z_i + b_i)) and res + dropout(gelu(z + b)) against torch
    (p = 0), and for p > 0 the backward must re-derive exactly the forward's masks (keep scale 1/(1-p))."""
    B, H = 2, 16
    dev = _dev()
    zs = [_rand(B, C, H, H, seed=10 + i).requires_grad_(True) for i in range(3)]
    bs = [(_rand(C, seed=20 + i) * 0.3).float().requires_grad_(True) for i in range(3)]
    res = _rand(B, C, H, H, seed=30).requires_grad_(True)
    seeds = (11, 22, 33)
    y = ops._WfCombineFn.apply(*[ops.to_nhwc(z) for z in zs], *bs, p, seeds)
    out = ops._WfOutFn.apply(y, bs[0], ops.to_nhwc(res), p, 44)
    o = ops.to_nchw(out, C)
    gy = _rand(B, C, H, H, seed=31)
    o.backward(gy)
    got = [t.grad.clone() for t in (*zs, *bs, res)]
    for t in (*zs, *bs, res):
        t.grad = None
    if p == 0.0:
        yr = sum(bf16r(F.gelu(z + b.view(1, C, 1, 1))) for z, b in zip(zs, bs))
        yr = bf16r(yr)
        ref = res + bf16r(F.gelu(yr + bs[0].view(1, C, 1, 1)))
        assert rel_l2(o, ref) < FWD_TOL
        ref.backward(gy)
        for g, t in zip(got, (*zs, *bs, res)):
            assert rel_l2(g, t.grad) < 1.5e-2
        return
    # p > 0: out - res = gelu(y + b) * m with m in {0, 1/(1-p)}; backward dy = gy * m * gelu'(y + b)
    yn = ops.to_nchw(y, C).detach()
    act = bf16r(F.gelu(yn + bs[0].detach().view(1, C, 1, 1)))
    ratio = (o.detach() - res.detach()) / act
    big = act.abs() > 0.05
    m_fwd = (ratio[big] > 0.5 / (1 - p)).float()
    keep = float(m_fwd.mean())
    assert abs(keep - (1 - p)) < 0.03, keep
    kept = ratio[big][m_fwd > 0]
    assert float((kept - 1 / (1 - p)).abs().max()) < 0.08          # bf16 rounding of out and res
    assert torch.equal(got[-1], bf16r(gy))                          # the residual gradient passes through
    # the three branch gradients are zero exactly where the branch was dropped; each branch keeps ~1-p
    for gz in got[:3]:
        frac = float((gz != 0).float().mean())
        assert abs(frac - (1 - p) * (1 - p)) < 0.05, frac           # kept by its own mask AND by the out mask
    # same seeds -> same masks (forward is a pure function of (seed, step counter, index))
    y2 = ops._WfCombineFn.apply(*[ops.to_nhwc(z.detach()) for z in zs], *[b.detach() for b in bs], p, seeds)
    assert torch.equal(y2, y.detach())


def test_wide_focus_module_matches_unfused_reference(ops):
    """Wide_Focus through the fused path vs the same block written with plain torch ops (dropout off)."""
    import hybridnet
    torch.manual_seed(0)
    C = 16
    m = hybridnet.Wide_Focus(C, C).to(_dev()).eval()
    x = _rand(2, C, 24, 24, seed=5).requires_grad_(True)
    out = m(x)
    gy = _rand(2, C, 24, 24, seed=6)
    out.backward(gy)
    got = [x.grad.clone()] + [p.grad.clone() for p in m.parameters()]
    x.grad = None
    for p_ in m.parameters():
        p_.grad = None
    w = {k: bf16r(v.detach()).requires_grad_(True) if k.endswith("weight") else v.detach().clone().requires_grad_(True)
         for k, v in m.state_dict().items()}
    ys = [bf16r(F.gelu(bf16r(F.conv2d(x, w[f"conv{i}.weight"], None, 1, d, d)) + w[f"conv{i}.bias"].view(1, C, 1, 1)))
          for i, d in ((1, 1), (2, 2), (3, 3))]
    added = bf16r(ys[0] + ys[1] + ys[2])
    ref = bf16r(F.gelu(bf16r(F.conv2d(added, w["conv4.weight"], None, 1, 1, 1)) + w["conv4.bias"].view(1, C, 1, 1)))
    assert rel_l2(out, ref) < 6e-3
    ref.backward(gy)
    want = [x.grad] + [w[k].grad for k, _ in m.named_parameters()]
    for g, r in zip(got, want):
        assert rel_l2(g, r) < 2e-2
