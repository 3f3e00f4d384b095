"""Parameter-holding layers.  They subclass the torch.nn layer types ONLY to get
identical parameter names, shapes and default initialisation (hence identical
``state_dict`` layout and RNG consumption order as the reference); their compute
is re-routed to the sm_100a kernels (``ops``).  ``forward`` keeps the torch.nn
calling convention (NCHW fp32 in/out) so a sub-module can be used stand-alone;
the model-level forwards call the ``nhwc`` methods and stay in NHWC bf16.
"""
import torch
import torch.nn as nn

from . import ops


def _same_pad(k, dil):
    return dil * (k - 1) // 2


class Conv2d(nn.Conv2d):
    def _geom(self):
        k = self.kernel_size[0]
        dil = self.dilation[0]
        pad = _same_pad(k, dil) if isinstance(self.padding, str) else self.padding[0]
        return self.stride[0], pad, dil

    def nhwc(self, x, act="none", x2=None, res=None, ups=0):
        if self.groups != 1:
            raise RuntimeError("hybridnet: grouped conv only exists fused inside Attention")
        stride, pad, dil = self._geom()
        return ops.conv2d(x, self.weight, self.bias, x2=x2, res=res, stride=stride, pad=pad, dil=dil, ups=ups, act=act)

    def forward(self, x):
        return ops.to_nchw(self.nhwc(ops.to_nhwc(x)), self.out_channels)


class ConvTranspose2d(nn.ConvTranspose2d):
    def nhwc(self, x, act="none"):
        return ops.conv2d(x, self.weight, self.bias, stride=2, pad=0, act=act, transposed=True)

    def forward(self, x):
        return ops.to_nchw(self.nhwc(ops.to_nhwc(x)), self.out_channels)


class BatchNorm2d(nn.BatchNorm2d):
    def nhwc(self, x, act="none", res=None):
        training = self.training or self.running_mean is None
        if training:
            ops._fold_cache.pop(id(self), None)        # running statistics are about to move (raw-pointer update)
        if training and self.num_batches_tracked is not None:
            self.num_batches_tracked.add_(1)
        return ops.bn_act(x, self.weight, self.bias, self.running_mean, self.running_var, training, res=res,
                          momentum=self.momentum, eps=self.eps, act=act)

    def forward(self, x):
        return ops.to_nchw(self.nhwc(ops.to_nhwc(x)), self.num_features)


def conv_bn(conv, bn, x, act="relu", x2=None, res=None):
    """conv -> bn -> act with the batch statistics taken in the conv epilogue."""
    if isinstance(conv, nn.ConvTranspose2d):
        return ops.conv_bn_act(x, conv.weight, conv.bias, bn, stride=2, pad=0, act=act, transposed=True, res=res)
    stride, pad, dil = conv._geom()
    return ops.conv_bn_act(x, conv.weight, conv.bias, bn, x2=x2, res=res, stride=stride, pad=pad, dil=dil, act=act)


class LayerNorm(nn.LayerNorm):
    """LayerNorm over channels; ``nhwc`` takes [B,H,W,C] bf16."""

    def nhwc(self, x):
        return ops.layernorm_c(x, self.weight, self.bias, self.eps)

    def forward(self, x):  # [..., C] like nn.LayerNorm
        shp = x.shape
        y = self.nhwc(x.reshape(1, 1, -1, shp[-1]).to(torch.bfloat16).contiguous())
        return y.reshape(shp).to(x.dtype)


class MultiheadAttention(nn.MultiheadAttention):
    """nn.MultiheadAttention(batch_first=True) parameter layout (packed in_proj_weight [3C,C],
    in_proj_bias, out_proj.{weight,bias}) with the core replaced by the flash kernel."""

    def nhwc(self, q, k, v):
        C = self.embed_dim
        w = self.in_proj_weight.view(3, C, C, 1, 1)
        b = self.in_proj_bias.view(3, C) if self.in_proj_bias is not None else (None, None, None)
        qp = ops.conv2d(q, w[0], b[0], pad=0)
        kp = ops.conv2d(k, w[1], b[1], pad=0)
        vp = ops.conv2d(v, w[2], b[2], pad=0)
        o = ops.attention(qp, kp, vp, self.num_heads)
        return ops.conv2d(o, self.out_proj.weight.view(C, C, 1, 1), self.out_proj.bias, pad=0)

    def forward(self, query, key, value, need_weights=False, **kw):
        if need_weights or kw.get("attn_mask") is not None or kw.get("key_padding_mask") is not None:
            raise RuntimeError("hybridnet: attention weights / masks are not part of the reference path (FCT.py:75)")
        B, N, C = query.shape
        f = lambda t: t.reshape(B, 1, N, C).to(torch.bfloat16).contiguous()
        o = self.nhwc(f(query), f(key), f(value))
        return o.reshape(B, N, C).to(query.dtype), None
