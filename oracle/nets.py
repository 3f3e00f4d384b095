"""CPU oracle for the hybrid segmentation nets -- TEST INFRASTRUCTURE ONLY.

This file is a plain-PyTorch fp32 *functional* restatement of the reference's
hot path (forward passes; backward comes from torch autograd on CPU).  It is
NOT part of the product: only ``tests/``, ``__graft_entry__.smoke()`` and the
``cpu_baseline`` / ``--impl reference`` legs of ``bench.py`` may import it.  The
product path (package ``hybridnet``) never imports anything under ``oracle/``.

Every function works on a flat ``state_dict`` (name -> tensor) with exactly the
reference's key names, so a checkpoint saved by the reference can be fed here
unchanged.  Each function cites the reference lines it follows
(paths relative to the reference checkout).

Pinning: ``tests/golden/make_golden.py`` imported the *real* reference classes
(UNet.py, AutoEncoder.py, AE_Unet_Combined.py, FCT.py lines 1-255) in the build
container, loaded the state_dicts produced by :func:`init_state_dict` with
``strict=True`` and stored the reference outputs / gradients in
``tests/golden/*.pt``; ``tests/test_oracle_golden.py`` checks this file against
them.  The Autoencoder32K family has no source in the reference (bytecode only,
``__pycache__/AE_32K.cpython-38.pyc``); for it **parity is unpinned** beyond
the state_dict key list / parameter counts recorded in SURVEY.md App. A.
"""
from __future__ import annotations

import math
from collections import OrderedDict
from typing import Callable, Dict, Optional

import torch
import torch.nn.functional as F

SD = Dict[str, torch.Tensor]


# --------------------------------------------------------------------------
# small helpers
# --------------------------------------------------------------------------
class Ctx:
    """Per-call options: train/eval, dropout handling, BN running-stat sink."""

    def __init__(self, training: bool = True, dropout: str = "off",
                 masks: Optional[Dict[str, torch.Tensor]] = None,
                 bn_sink: Optional[SD] = None, generator=None):
        # dropout: "off" (identity), "on" (torch RNG), "masks" (injected keep
        # masks already scaled by 1/(1-p), keyed by site name)
        self.training = training
        self.dropout = dropout
        self.masks = masks or {}
        self.bn_sink = bn_sink
        self.generator = generator

    def drop(self, x, p, site, channelwise=False):
        if not self.training or self.dropout == "off" or p == 0.0:
            return x
        if self.dropout == "masks":
            return x * self.masks[site]
        if channelwise:
            return F.dropout2d(x, p, True)
        return F.dropout(x, p, True)


# ---- optional emulation of the product path's bf16 rounding points ------------
# With ``with nets.bf16_emulation():`` every conv/linear operand and every stored
# activation is rounded to bf16 (straight-through in backward) exactly where the
# CUDA path rounds, while all accumulation stays fp32.  Tests use it to separate
# "kernel is wrong" from "bf16 storage loses precision" (parity).
_EMU = [False]


class bf16_emulation:
    def __enter__(self):
        self.prev = _EMU[0]
        _EMU[0] = True

    def __exit__(self, *a):
        _EMU[0] = self.prev


class _RoundSTE(torch.autograd.Function):
    @staticmethod
    def forward(ctx, t):
        return t.to(torch.bfloat16).to(t.dtype)

    @staticmethod
    def backward(ctx, g):
        return g


def _q(t):
    return _RoundSTE.apply(t) if _EMU[0] else t


def _conv2d(x, w, b=None, stride=1, padding=0, dilation=1, groups=1, wq=True):
    """F.conv2d with bf16 operands under emulation (weights of the depthwise
    projections stay fp32 in the product, wq=False)."""
    return F.conv2d(_q(x), _q(w) if wq else w, b, stride, padding, dilation, groups)


def _conv(sd: SD, name: str, x, stride=1, padding=1, dilation=1, groups=1):
    return _conv2d(x, sd[name + ".weight"], sd.get(name + ".bias"), stride,
                   padding, dilation, groups, wq=(groups == 1))


def _ln_c(sd: SD, name: str, x):
    """LayerNorm over the channel axis of an NCHW tensor (FCT.py:43-45)."""
    c = x.shape[1]
    y = F.layer_norm(x.permute(0, 2, 3, 1), (c,), sd[name + ".weight"],
                     sd[name + ".bias"], 1e-5)
    return _q(y.permute(0, 3, 1, 2))


def _bn(sd: SD, name: str, x, ctx: Ctx):
    """BatchNorm2d, batch statistics in training (UNet.py:59)."""
    x = _q(x)
    rm, rv = sd[name + ".running_mean"], sd[name + ".running_var"]
    if ctx.training:
        rm, rv = rm.clone(), rv.clone()
    y = F.batch_norm(x, rm, rv, sd[name + ".weight"], sd[name + ".bias"],
                     ctx.training, 0.1, 1e-5)
    if ctx.training and ctx.bn_sink is not None:
        ctx.bn_sink[name + ".running_mean"] = rm
        ctx.bn_sink[name + ".running_var"] = rv
        ctx.bn_sink[name + ".num_batches_tracked"] = \
            sd[name + ".num_batches_tracked"] + 1
    return y


def _up2(x):
    return F.interpolate(x, scale_factor=2.0, mode="nearest")


# --------------------------------------------------------------------------
# FCT  (FCT.py:24-254)
# --------------------------------------------------------------------------
def mha(sd: SD, p: str, q, k, v, heads: int):
    """nn.MultiheadAttention(batch_first, need_weights=False) math, FCT.py:37,75.
    q,k,v: [B,N,C]."""
    B, N, C = q.shape
    d = C // heads
    w, b = sd[p + ".in_proj_weight"], sd[p + ".in_proj_bias"]
    qp = _q(F.linear(_q(q), _q(w[:C]), b[:C])).view(B, N, heads, d).transpose(1, 2)
    kp = _q(F.linear(_q(k), _q(w[C:2 * C]), b[C:2 * C])).view(B, N, heads, d).transpose(1, 2)
    vp = _q(F.linear(_q(v), _q(w[2 * C:]), b[2 * C:])).view(B, N, heads, d).transpose(1, 2)
    s = torch.matmul(qp, kp.transpose(-1, -2)) / math.sqrt(d)
    if _EMU[0]:   # flash form: un-normalised probabilities are rounded, the row sum is not
        e = torch.exp(s - s.amax(dim=-1, keepdim=True))
        o = torch.matmul(_q(e), vp) / e.sum(dim=-1, keepdim=True)
    else:
        o = torch.matmul(torch.softmax(s, dim=-1), vp)
    o = _q(o.transpose(1, 2).reshape(B, N, C))
    return _q(F.linear(o, _q(sd[p + ".out_proj.weight"]), sd[p + ".out_proj.bias"]))


def fct_attention(sd: SD, p: str, x, heads=2):
    """Attention.forward, FCT.py:41-79."""
    B, C, H, W = x.shape
    proj = []
    for t in "qkv":
        y = F.relu(_conv(sd, f"{p}.conv_{t}", x, groups=C))      # fused with LN in the product: no rounding here
        y = _ln_c(sd, f"{p}.layernorm_{t}", y)
        proj.append(y.reshape(B, C, H * W).permute(0, 2, 1))
    o = mha(sd, p + ".attention", proj[0], proj[1], proj[2], heads)
    return o.permute(0, 2, 1).reshape(B, C, H, W)


def fct_wide_focus(sd: SD, p: str, x, ctx: Ctx):
    """Wide_Focus.forward, FCT.py:117-132."""
    acc = None
    for i, dil in ((1, 1), (2, 2), (3, 3)):
        y = _q(F.gelu(_q(_conv(sd, f"{p}.conv{i}", x, padding=dil, dilation=dil))))
        y = ctx.drop(y, 0.1, f"{p}.drop{i}")
        acc = y if acc is None else acc + y
    y = _q(F.gelu(_q(_conv(sd, f"{p}.conv4", _q(acc)))))
    return ctx.drop(y, 0.1, f"{p}.drop4")


def fct_transformer(sd: SD, p: str, x, ctx: Ctx):
    """Transformer.forward, FCT.py:93-102."""
    a = fct_attention(sd, p + ".attention_output", x)
    x2 = _q(_conv(sd, p + ".conv1", a) + x)
    x3 = fct_wide_focus(sd, p + ".wide_focus", _ln_c(sd, p + ".layernorm", x2), ctx)
    return _q(x2 + x3)


def fct_enc_block(sd: SD, p: str, blk: str, x, scale_img, ctx: Ctx, with_trans=True):
    """Block_encoder_bottleneck.forward (FCT.py:149-162) and, with
    with_trans=False, EncoderBlock.forward (AutoEncoder.py:17-27)."""
    if blk in ("first", "bottleneck"):
        y = _q(F.relu(_conv(sd, p + ".conv1_a", x)))
        y = _q(F.relu(_conv(sd, p + ".conv2", y)))
    else:
        skip = _q(F.relu(_conv(sd, p + ".conv1_b", scale_img)))
        y = torch.cat([skip, x], dim=1)
        y = _q(F.relu(_conv(sd, p + ".conv2", y)))
        y = _q(F.relu(_conv(sd, p + ".conv3", y)))
    y = F.max_pool2d(ctx.drop(y, 0.3, p + ".drop"), 2)
    return fct_transformer(sd, p + ".trans", y, ctx) if with_trans else y


def fct_dec_block(sd: SD, p: str, x, skip, ctx: Ctx):
    """Block_decoder.forward, FCT.py:178-186."""
    y = _q(F.relu(_conv(sd, p + ".conv1", _up2(x))))
    y = torch.cat((skip, y), dim=1)
    y = _q(F.relu(_conv(sd, p + ".conv2", y)))
    y = _q(F.relu(_conv(sd, p + ".conv3", y)))
    y = ctx.drop(y, 0.3, p + ".drop")
    return fct_transformer(sd, p + ".trans", y, ctx)


def fct_forward(sd: SD, x, ctx: Optional[Ctx] = None, pre_sigmoid=False):
    """FCT.forward, FCT.py:235-254."""
    ctx = ctx or Ctx()
    x = _q(x)
    s2 = _q(F.avg_pool2d(x, 2))
    s3 = _q(F.avg_pool2d(s2, 2))
    s4 = _q(F.avg_pool2d(s3, 2))
    x1 = fct_enc_block(sd, "block_1", "first", x, None, ctx)
    x2 = fct_enc_block(sd, "block_2", "second", x1, s2, ctx)
    x3 = fct_enc_block(sd, "block_3", "third", x2, s3, ctx)
    x4 = fct_enc_block(sd, "block_4", "fourth", x3, s4, ctx)
    y = fct_enc_block(sd, "block_5", "bottleneck", x4, None, ctx)
    y = fct_dec_block(sd, "block_6", y, x4, ctx)
    y = fct_dec_block(sd, "block_7", y, x3, ctx)
    y = fct_dec_block(sd, "block_8", y, x2, ctx)
    y = fct_dec_block(sd, "block_9", y, x1, ctx)
    # DS_out.forward, FCT.py:201-206
    y = _q(F.relu(_conv(sd, "ds.conv1", _up2(y))))
    y = _q(F.relu(_conv(sd, "ds.conv2", y)))
    z = _conv(sd, "ds.conv3", y)
    return _q(z) if pre_sigmoid else _q(torch.sigmoid(z))


# --------------------------------------------------------------------------
# UNet  (UNet.py:6-66)
# --------------------------------------------------------------------------
def _unet_block(sd: SD, p: str, name: str, x, ctx: Ctx):
    """UNet._block, UNet.py:54-66: [conv3x3(no bias) -> BN -> ReLU] x2."""
    for i in (1, 2):
        x = _conv2d(x, sd[f"{p}.{name}conv{i}.weight"], None, 1, 1)
        x = _q(F.relu(_bn(sd, f"{p}.{name}norm{i}", x, ctx)))
    return x


def unet_forward(sd: SD, x, ctx: Optional[Ctx] = None, pre_sigmoid=False, prefix=""):
    """UNet.forward, UNet.py:32-51."""
    ctx = ctx or Ctx()
    P = prefix
    x = _q(x)
    e1 = _unet_block(sd, P + "encoder1", "enc1", x, ctx)
    e2 = _unet_block(sd, P + "encoder2", "enc2", F.max_pool2d(e1, 2), ctx)
    e3 = _unet_block(sd, P + "encoder3", "enc3", F.max_pool2d(e2, 2), ctx)
    e4 = _unet_block(sd, P + "encoder4", "enc4", F.max_pool2d(e3, 2), ctx)
    y = _unet_block(sd, P + "bottleneck", "bottleneck", F.max_pool2d(e4, 2), ctx)
    for lvl, skip in ((4, e4), (3, e3), (2, e2), (1, e1)):
        y = _q(F.conv_transpose2d(_q(y), _q(sd[f"{P}upconv{lvl}.weight"]),
                                  sd[f"{P}upconv{lvl}.bias"], stride=2))
        y = torch.cat((y, skip), dim=1)
        y = _unet_block(sd, f"{P}decoder{lvl}", f"dec{lvl}", y, ctx)
    z = _conv2d(y, sd[P + "conv.weight"], sd[P + "conv.bias"])
    return _q(z) if pre_sigmoid else _q(torch.sigmoid(z))


# --------------------------------------------------------------------------
# AutoEncoder / Image2Image2Mask  (AutoEncoder.py, AE_Unet_Combined.py)
# --------------------------------------------------------------------------
def autoencoder_forward(sd: SD, x, ctx: Optional[Ctx] = None, prefix=""):
    """AutoEncoder.forward (AutoEncoder.py:131-134) -> (latent, output)."""
    ctx = ctx or Ctx()
    E = prefix + "encoder."
    x = _q(x)
    s2 = _q(F.avg_pool2d(x, 2))
    s3 = _q(F.avg_pool2d(s2, 2))
    s4 = _q(F.avg_pool2d(s3, 2))          # scale_img_5 (AutoEncoder.py:92) is dead
    x1 = fct_enc_block(sd, E + "block_1", "first", x, None, ctx, False)
    x2 = fct_enc_block(sd, E + "block_2", "second", x1, s2, ctx, False)
    x3 = fct_enc_block(sd, E + "block_3", "third", x2, s3, ctx, False)
    x4 = fct_enc_block(sd, E + "block_4", "fourth", x3, s4, ctx, False)
    latent = fct_enc_block(sd, E + "block_5", "bottleneck", x4, None, ctx, False)
    y = latent
    D = prefix + "decoder."
    for b in (4, 3, 2, 1):            # DecoderBlock.forward, AutoEncoder.py:42-48
        p = f"{D}block_{b}"
        y = _q(F.relu(_conv(sd, p + ".conv1", _up2(y))))
        y = _q(F.relu(_conv(sd, p + ".conv2", y)))
        y = _q(F.relu(_conv(sd, p + ".conv3", y)))
        y = ctx.drop(y, 0.3, p + ".drop")
    # DeepSupervisionBlock.forward, AutoEncoder.py:63-68 (ReLU head)
    y = _q(F.relu(_conv(sd, D + "ds.conv1", _up2(y))))
    y = _q(F.relu(_conv(sd, D + "ds.conv2", y)))
    y = _q(F.relu(_conv(sd, D + "ds.conv3", y)))
    return latent, y


def i2i2m_forward(sd: SD, x, ctx: Optional[Ctx] = None):
    """Image2Image2Mask.forward, AE_Unet_Combined.py:14-17."""
    ctx = ctx or Ctx()
    latent, recon = autoencoder_forward(sd, x, ctx, "image2imageAE.")
    mask = unet_forward(sd, recon, ctx, prefix="unet.")
    return latent, recon, mask


# --------------------------------------------------------------------------
# Autoencoder32K family (bytecode only: __pycache__/AE_32K.cpython-38.pyc;
# source line numbers from its lnotab)
# --------------------------------------------------------------------------
def _bottleneck(sd: SD, p: str, x, stride: int, ctx: Ctx):
    """Bottleneck.forward, pyc:AE_32K:L34-52."""
    y = _q(F.relu(_bn(sd, p + ".bn1", _conv2d(x, sd[p + ".conv1.weight"]), ctx)))
    y = _q(F.relu(_bn(sd, p + ".bn2",
                   _conv2d(y, sd[p + ".conv2.weight"], None, stride, 1), ctx)))
    y = _bn(sd, p + ".bn3", _conv2d(y, sd[p + ".conv3.weight"]), ctx)
    if p + ".downsample.0.weight" in sd:
        x = _q(_bn(sd, p + ".downsample.1",
                   _conv2d(x, sd[p + ".downsample.0.weight"], None, stride), ctx))
    return _q(F.relu(y + x))


def encoder32k_forward(sd: SD, x, ctx: Ctx, p=""):
    """Encoder_32K.forward, pyc:AE_32K:L104-115 -> [B,8,H/4,W/4]."""
    y = _q(F.relu(_bn(sd, p + "bn1", _conv2d(x, sd[p + "conv1.weight"], None, 2, 3), ctx)))
    for i in range(3):
        y = _bottleneck(sd, f"{p}layer1.{i}", y, 1, ctx)
    y = ctx.drop(y, 0.3, p + "drop1", True)
    for i in range(4):
        y = _bottleneck(sd, f"{p}layer2.{i}", y, 2 if i == 0 else 1, ctx)
    y = ctx.drop(y, 0.3, p + "drop2", True)
    for i in (2, 3, 4):
        y = _q(F.relu(_bn(sd, f"{p}bn{i}", _conv(sd, f"{p}conv{i}", y), ctx)))
    y = ctx.drop(_conv(sd, p + "conv5", y), 0.3, p + "drop3", True)  # L113 order
    return _q(F.relu(_bn(sd, p + "bn5", y, ctx)))


def decoder32k_forward(sd: SD, x, ctx: Ctx, p="", pre_activation=False):
    """Decoder_32K.forward, pyc:AE_32K:L162-168."""
    y = x
    for i in (1, 2, 3, 4):
        y = _q(F.relu(_bn(sd, f"{p}bn{i}", _conv(sd, f"{p}conv{i}", y), ctx)))
    y = _q(F.conv_transpose2d(_q(y), _q(sd[p + "transConv1.weight"]), sd[p + "transConv1.bias"], stride=2))
    y = _q(F.relu(_bn(sd, p + "dbn2", y, ctx)))
    y = _q(F.conv_transpose2d(_q(y), _q(sd[p + "transConv2.weight"]), sd[p + "transConv2.bias"], stride=2))
    y = _q(F.relu(_bn(sd, p + "dbn3", y, ctx)))
    y = _q(F.relu(_bn(sd, p + "bn5", _conv(sd, p + "conv5", y), ctx)))
    y = _q(F.relu(_bn(sd, p + "bn6", _conv(sd, p + "conv6", y), ctx)))
    z = _bn(sd, p + "outputDeterminerNorm", _conv(sd, p + "outputDeterminerConv", y), ctx)
    if pre_activation:
        return z
    is_mask = sd[p + "outputDeterminerConv.weight"].shape[0] == 1
    return _q(torch.sigmoid(z) if is_mask else F.relu(z))


def ae32k_forward(sd: SD, x, ctx: Optional[Ctx] = None, enc="encoder.", dec="decoder."):
    """Autoencoder32K.forward (pyc:AE_32K:L183-186); with enc="0.", dec="1."
    it is the latent-to-mask Sequential (pyc:AE_32K:L241-244)."""
    ctx = ctx or Ctx()
    return decoder32k_forward(sd, encoder32k_forward(sd, x, ctx, enc), ctx, dec)


# --------------------------------------------------------------------------
# losses  (Metrics.py:5-22; nn.MSELoss)
# --------------------------------------------------------------------------
def dice_loss(y_pred, y_true, smooth=1.0):
    """DiceLoss.forward, Metrics.py:14-22 (channel 0 only)."""
    p = y_pred[:, 0].reshape(-1)
    t = y_true[:, 0].reshape(-1)
    inter = (p * t).sum()
    return 1.0 - (2.0 * inter + smooth) / (p.sum() + t.sum() + smooth)


def jaccard(y_pred, y_true):
    """JaccardScore.forward, Metrics.py:31-42 (threshold 0.5 masks)."""
    p = y_pred[:, 0].reshape(-1) > 0.5
    t = y_true[:, 0].reshape(-1) > 0.5
    union = (p | t).sum().item()
    return 1.0 if union == 0 else (p & t).sum().item() / union


# --------------------------------------------------------------------------
# deterministic parameter specs + init  (shapes follow the ctor lines cited)
# --------------------------------------------------------------------------
def _conv_spec(spec, name, cout, cin, k, bias=True, groups=1):
    spec[name + ".weight"] = ("conv_w", (cout, cin // groups, k, k))
    if bias:
        spec[name + ".bias"] = ("conv_b", (cout,), (cin // groups) * k * k)


def _ln_spec(spec, name, c):
    spec[name + ".weight"] = ("ones", (c,))
    spec[name + ".bias"] = ("zeros", (c,))


def _bn_spec(spec, name, c):
    spec[name + ".weight"] = ("ones", (c,))
    spec[name + ".bias"] = ("zeros", (c,))
    spec[name + ".running_mean"] = ("zeros", (c,))
    spec[name + ".running_var"] = ("ones", (c,))
    spec[name + ".num_batches_tracked"] = ("count", ())


def _trans_spec(spec, p, c):
    a = p + ".attention_output"
    for t in "qkv":                                   # FCT.py:31-36
        if t == "q":
            pass
        _conv_spec(spec, f"{a}.conv_{t}", c, c, 3, True, groups=c)
        _ln_spec(spec, f"{a}.layernorm_{t}", c)
    spec[a + ".attention.in_proj_weight"] = ("xavier", (3 * c, c))
    spec[a + ".attention.in_proj_bias"] = ("zeros", (3 * c,))
    spec[a + ".attention.out_proj.weight"] = ("lin_w", (c, c))
    spec[a + ".attention.out_proj.bias"] = ("zeros", (c,))
    _conv_spec(spec, p + ".conv1", c, c, 3)           # FCT.py:88
    _ln_spec(spec, p + ".layernorm", c)
    for i in (1, 2, 3, 4):                            # FCT.py:110-113
        _conv_spec(spec, f"{p}.wide_focus.conv{i}", c, c, 3)


def _encblock_spec(spec, p, cin, cout, trans):
    _conv_spec(spec, p + ".conv1_a", cout, cin, 3)    # FCT.py:140-143
    _conv_spec(spec, p + ".conv1_b", cin, 3, 3)
    _conv_spec(spec, p + ".conv2", cout, cout, 3)
    _conv_spec(spec, p + ".conv3", cout, cout, 3)
    if trans:
        _trans_spec(spec, p + ".trans", cout)


def fct_spec():
    spec = OrderedDict()
    f = [8, 16, 32, 64, 128, 64, 32, 16, 8]           # FCT.py:215
    cin = 3
    for i in range(5):
        _encblock_spec(spec, f"block_{i + 1}", cin, f[i], True)
        cin = f[i]
    for i in range(5, 9):                             # FCT.py:170-174
        p = f"block_{i + 1}"
        _conv_spec(spec, p + ".conv1", f[i], f[i - 1], 3)
        _conv_spec(spec, p + ".conv2", f[i], 2 * f[i], 3)
        _conv_spec(spec, p + ".conv3", f[i], f[i], 3)
        _trans_spec(spec, p + ".trans", f[i])
    _conv_spec(spec, "ds.conv1", 8, 8, 3)             # FCT.py:194-197
    _conv_spec(spec, "ds.conv2", 8, 8, 3)
    _conv_spec(spec, "ds.conv3", 1, 8, 3)
    return spec


def unet_spec(prefix="", in_channels=3, out_channels=1, f=32):
    spec = OrderedDict()

    def block(p, name, cin, c):
        for i, ci in ((1, cin), (2, c)):
            _conv_spec(spec, f"{prefix}{p}.{name}conv{i}", c, ci, 3, False)
            _bn_spec(spec, f"{prefix}{p}.{name}norm{i}", c)

    block("encoder1", "enc1", in_channels, f)
    block("encoder2", "enc2", f, 2 * f)
    block("encoder3", "enc3", 2 * f, 4 * f)
    block("encoder4", "enc4", 4 * f, 8 * f)
    block("bottleneck", "bottleneck", 8 * f, 16 * f)
    for lvl, c in ((4, 8 * f), (3, 4 * f), (2, 2 * f), (1, f)):
        spec[f"{prefix}upconv{lvl}.weight"] = ("convT_w", (2 * c, c, 2, 2))
        spec[f"{prefix}upconv{lvl}.bias"] = ("conv_b", (c,), c * 4)
        block(f"decoder{lvl}", f"dec{lvl}", 2 * c, c)
    _conv_spec(spec, prefix + "conv", out_channels, f, 1)
    return spec


def autoencoder_spec(prefix=""):
    spec = OrderedDict()
    f = [8, 16, 32, 64, 128]
    cin = 3
    for i in range(5):
        _encblock_spec(spec, f"{prefix}encoder.block_{i + 1}", cin, f[i], False)
        cin = f[i]
    g = [128, 64, 32, 16, 8]
    for j, b in enumerate((4, 3, 2, 1)):
        p = f"{prefix}decoder.block_{b}"
        _conv_spec(spec, p + ".conv1", g[j + 1], g[j], 3)
        _conv_spec(spec, p + ".conv2", g[j + 1], g[j + 1], 3)
        _conv_spec(spec, p + ".conv3", g[j + 1], g[j + 1], 3)
    _conv_spec(spec, prefix + "decoder.ds.conv1", 8, 8, 3)
    _conv_spec(spec, prefix + "decoder.ds.conv2", 8, 8, 3)
    _conv_spec(spec, prefix + "decoder.ds.conv3", 3, 8, 3)
    return spec


def i2i2m_spec():
    spec = autoencoder_spec("image2imageAE.")
    spec.update(unet_spec("unet."))
    return spec


def encoder32k_spec(p=""):
    spec = OrderedDict()
    _conv_spec(spec, p + "conv1", 64, 3, 7, False)
    _bn_spec(spec, p + "bn1", 64)
    inpl = 64
    for layer, planes, n in (("layer1", 64, 3), ("layer2", 128, 4)):
        for i in range(n):
            q = f"{p}{layer}.{i}"
            _conv_spec(spec, q + ".conv1", planes, inpl, 1, False)
            _bn_spec(spec, q + ".bn1", planes)
            _conv_spec(spec, q + ".conv2", planes, planes, 3, False)
            _bn_spec(spec, q + ".bn2", planes)
            _conv_spec(spec, q + ".conv3", planes * 4, planes, 1, False)
            _bn_spec(spec, q + ".bn3", planes * 4)
            if i == 0:
                _conv_spec(spec, q + ".downsample.0", planes * 4, inpl, 1, False)
                _bn_spec(spec, q + ".downsample.1", planes * 4)
            inpl = planes * 4
    for i, (ci, co) in zip((2, 3, 4, 5), ((512, 128), (128, 64), (64, 16), (16, 8))):
        _conv_spec(spec, f"{p}conv{i}", co, ci, 3)
        _bn_spec(spec, f"{p}bn{i}", co)
    return spec


def decoder32k_spec(p="", output="image"):
    spec = OrderedDict()
    for i, (ci, co) in zip((1, 2, 3, 4), ((8, 16), (16, 64), (64, 128), (128, 256))):
        _conv_spec(spec, f"{p}conv{i}", co, ci, 3)
        _bn_spec(spec, f"{p}bn{i}", co)
    spec[p + "transConv1.weight"] = ("convT_w", (256, 384, 2, 2))
    spec[p + "transConv1.bias"] = ("conv_b", (384,), 384 * 4)
    _bn_spec(spec, p + "dbn2", 384)
    spec[p + "transConv2.weight"] = ("convT_w", (384, 192, 2, 2))
    spec[p + "transConv2.bias"] = ("conv_b", (192,), 192 * 4)
    _bn_spec(spec, p + "dbn3", 192)
    _conv_spec(spec, p + "conv5", 64, 192, 3)
    _bn_spec(spec, p + "bn5", 64)
    _conv_spec(spec, p + "conv6", 8, 64, 3)
    _bn_spec(spec, p + "bn6", 8)
    c = 3 if output.lower() == "image" else 1
    _conv_spec(spec, p + "outputDeterminerConv", c, 8, 3)
    _bn_spec(spec, p + "outputDeterminerNorm", c)
    return spec


def ae32k_spec(output="image", enc="encoder.", dec="decoder."):
    spec = encoder32k_spec(enc)
    spec.update(decoder32k_spec(dec, output))
    return spec


SPECS: Dict[str, Callable[[], OrderedDict]] = {
    "fct": fct_spec,
    "unet": unet_spec,
    "autoencoder": autoencoder_spec,
    "i2i2m": i2i2m_spec,
    "ae32k_image": lambda: ae32k_spec("image"),
    "ae32k_mask": lambda: ae32k_spec("mask"),
    "latent2mask": lambda: ae32k_spec("mask", "0.", "1."),
}


def init_state_dict(model: str, seed: int = 0, jitter: float = 0.0) -> SD:
    """Deterministic PyTorch-default-like init from one torch.Generator.
    ``jitter`` perturbs norm affine params / zero biases so that parity tests
    exercise them (default init has gamma=1, beta=0)."""
    g = torch.Generator().manual_seed(seed)
    sd: SD = OrderedDict()
    for name, s in SPECS[model]().items():
        kind, shape = s[0], s[1]
        if kind in ("conv_w", "convT_w", "lin_w"):
            fan_in = shape[1] * (shape[2] * shape[3] if len(shape) == 4 else 1)
            if kind == "convT_w":
                fan_in = shape[1] * shape[2] * shape[3]
            b = 1.0 / math.sqrt(fan_in)
            t = (torch.rand(shape, generator=g) * 2 - 1) * b
        elif kind == "conv_b":
            b = 1.0 / math.sqrt(s[2])
            t = (torch.rand(shape, generator=g) * 2 - 1) * b
        elif kind == "xavier":
            b = math.sqrt(6.0 / (shape[0] + shape[1]))
            t = (torch.rand(shape, generator=g) * 2 - 1) * b
        elif kind == "ones":
            t = torch.ones(shape) + jitter * (torch.rand(shape, generator=g) * 2 - 1)
        elif kind == "zeros":
            t = jitter * (torch.rand(shape, generator=g) * 2 - 1)
            if name.endswith("running_mean"):
                t = torch.zeros(shape)
        elif kind == "count":
            t = torch.zeros((), dtype=torch.long)
        else:  # pragma: no cover
            raise KeyError(kind)
        sd[name] = t
    return sd


FORWARDS = {
    "fct": fct_forward,
    "unet": unet_forward,
    "autoencoder": autoencoder_forward,
    "i2i2m": i2i2m_forward,
    "ae32k_image": ae32k_forward,
    "ae32k_mask": ae32k_forward,
    "latent2mask": lambda sd, x, ctx=None: ae32k_forward(sd, x, ctx, "0.", "1."),
}

# fwd / fwd+bwd algorithmic GFLOP per 256x256 frame
GFLOP_PER_FRAME = {
    "unet": (24.14, 72.31), "autoencoder": (1.164, 3.450), "i2i2m": (25.31, 75.88),
    "fct": (21.80, 65.85), "ae32k_image": (54.66, 163.68), "ae32k_mask": (54.64, 163.6),
    "latent2mask": (54.64, 117.0),
}
