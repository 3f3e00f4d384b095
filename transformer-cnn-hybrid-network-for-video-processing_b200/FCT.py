"""FCT conv-attention U-net on the sm_100a kernels -- drop-in for the reference's
``FCT.py`` model classes (FCT.py:24-254): same class names, constructor
signatures, attribute names (=> identical ``state_dict``) and forward returns.
All activations between the input pack and the final unpack are NHWC bf16.
"""
import numpy as np
import torch
import torch.nn as nn

from . import ops
from .layers import Conv2d, LayerNorm, MultiheadAttention


def _io(fn):
    """Public ``forward``: NCHW fp32 in/out around the NHWC bf16 implementation."""
    def wrapped(self, *xs):
        ys = fn(self, *[ops.to_nhwc(x) if (torch.is_tensor(x) or isinstance(x, ops.PackedFrames)) else x for x in xs])
        return ops.to_nchw(ys, self._out_channels)
    return wrapped


class Attention(nn.Module):
    """Depthwise-conv projected multi-head self-attention (FCT.py:24-79)."""

    def __init__(self, channels, num_heads, proj_drop=0.0, kernel_size=3, stride_kv=1, stride_q=1,
                 padding_kv="same", padding_q="same", attention_bias=True):
        super().__init__()
        if (kernel_size, stride_kv, stride_q) != (3, 1, 1):
            raise NotImplementedError("hybridnet Attention: the reference only ever uses k=3, stride 1 (FCT.py:86)")
        self.stride_kv, self.stride_q = stride_kv, stride_q
        self.num_heads, self.proj_drop = num_heads, proj_drop
        self._out_channels = channels
        # registration order follows FCT.py:31-37 so default init consumes the RNG identically
        self.conv_q = Conv2d(channels, channels, kernel_size, stride_q, padding_q, bias=attention_bias, groups=channels)
        self.layernorm_q = LayerNorm(channels, eps=1e-5)
        self.conv_k = Conv2d(channels, channels, kernel_size, stride_kv, stride_kv, bias=attention_bias, groups=channels)
        self.layernorm_k = LayerNorm(channels, eps=1e-5)
        self.conv_v = Conv2d(channels, channels, kernel_size, stride_kv, stride_kv, bias=attention_bias, groups=channels)
        self.layernorm_v = LayerNorm(channels, eps=1e-5)
        self.attention = MultiheadAttention(embed_dim=channels, bias=attention_bias, batch_first=True,
                                            num_heads=num_heads)
        self.relu = nn.ReLU()
        self.dropout = nn.Dropout()          # constructed but unused in the reference too (FCT.py:39,78)

    def forward_conv_nhwc(self, x):
        convs = (self.conv_q, self.conv_k, self.conv_v)
        lns = (self.layernorm_q, self.layernorm_k, self.layernorm_v)
        return ops.dwconv_relu_ln_qkv(x, tuple(c.weight for c in convs), tuple(c.bias for c in convs),
                                      tuple(l.weight for l in lns), tuple(l.bias for l in lns), 1e-5)

    def nhwc(self, x):
        q, k, v = self.forward_conv_nhwc(x)
        return self.attention.nhwc(q, k, v)

    # -- reference-shaped public API ------------------------------------------------
    def _build_projection(self, x, qkv):
        q, k, v = self.forward_conv(x)
        return {"q": q, "k": k, "v": v}[qkv]

    def forward_conv(self, x):
        q, k, v = self.forward_conv_nhwc(ops.to_nhwc(x))
        C = self._out_channels
        return ops.to_nchw(q, C), ops.to_nchw(k, C), ops.to_nchw(v, C)

    forward = _io(nhwc)


class Wide_Focus(nn.Module):
    """Three dilated 3x3 branches (d=1,2,3) -> GELU -> sum -> conv -> GELU (FCT.py:107-132)."""

    def __init__(self, in_channels, out_channels):
        super().__init__()
        self._out_channels = out_channels
        self.conv1 = Conv2d(in_channels, out_channels, 3, 1, padding="same")
        self.conv2 = Conv2d(in_channels, out_channels, 3, 1, padding="same", dilation=2)
        self.conv3 = Conv2d(in_channels, out_channels, 3, 1, padding="same", dilation=3)
        self.conv4 = Conv2d(in_channels, out_channels, 3, 1, padding="same")
        self.gelu = nn.GELU()
        self.dropout = nn.Dropout(0.1)

    def nhwc(self, x, res=None):
        # the three dilated convs share one autograd node; GELU, dropout, the 3-way sum, conv4's GELU/dropout and
        # the caller's residual add run as two fused passes each way (csrc/hn_wide_focus.cu)
        return ops.wide_focus(x, (self.conv1, self.conv2, self.conv3, self.conv4), res, self.dropout.p, self.training)

    forward = _io(nhwc)


class Transformer(nn.Module):
    """Attention -> conv -> +x -> LayerNorm -> Wide_Focus -> + (FCT.py:84-102)."""

    def __init__(self, in_channels, out_channels, num_heads, dpr=None, proj_drop=0.0, attention_bias=True,
                 padding_q="same", padding_kv="same", stride_kv=1, stride_q=1):
        super().__init__()
        self._out_channels = out_channels
        self.attention_output = Attention(channels=in_channels, num_heads=num_heads, proj_drop=proj_drop,
                                          padding_q=padding_q, padding_kv=padding_kv, stride_kv=stride_kv,
                                          stride_q=stride_q, attention_bias=attention_bias)
        self.conv1 = Conv2d(out_channels, out_channels, 3, 1, padding="same")
        self.layernorm = LayerNorm(self.conv1.out_channels, eps=1e-5)
        self.wide_focus = Wide_Focus(out_channels, out_channels)

    def nhwc(self, x):
        a = self.attention_output.nhwc(x)
        x2 = self.conv1.nhwc(a, res=x)                   # conv + residual in one epilogue
        return self.wide_focus.nhwc(self.layernorm.nhwc(x2), res=x2)

    forward = _io(nhwc)


class Block_encoder_bottleneck(nn.Module):
    """FCT.py:136-162."""

    def __init__(self, blk, in_channels, out_channels, att_heads, dpr):
        super().__init__()
        self.blk = blk
        self._out_channels = out_channels
        self.conv1_a = Conv2d(in_channels, out_channels, 3, 1, padding="same")
        self.conv1_b = Conv2d(3, in_channels, 3, 1, padding="same")
        self.conv2 = Conv2d(out_channels, out_channels, 3, 1, padding="same")
        self.conv3 = Conv2d(out_channels, out_channels, 3, 1, padding="same")
        self.trans = Transformer(in_channels=out_channels, out_channels=out_channels, num_heads=att_heads, dpr=dpr)
        self.relu = nn.ReLU()
        self.dropout = nn.Dropout(p=0.3)
        self.maxpool = nn.MaxPool2d(kernel_size=(2, 2))

    def convs_nhwc(self, x, scale_img="none"):
        if self.blk in ("first", "bottleneck"):
            x1 = self.conv1_a.nhwc(x, act="relu")
            x1 = self.conv2.nhwc(x1, act="relu")
        else:
            skip_x = self.conv1_b.nhwc(scale_img, act="relu")
            x1 = self.conv2.nhwc(skip_x, act="relu", x2=x)      # cat([skip_x, x]) is virtual
            x1 = self.conv3.nhwc(x1, act="relu")
        return ops.maxpool2(ops.dropout(x1, self.dropout.p, self.training))

    def nhwc(self, x, scale_img="none"):
        return self.trans.nhwc(self.convs_nhwc(x, scale_img))

    forward = _io(nhwc)


class Block_decoder(nn.Module):
    """FCT.py:167-186."""

    def __init__(self, in_channels, out_channels, att_heads, dpr):
        super().__init__()
        self._out_channels = out_channels
        self.upsample = nn.Upsample(scale_factor=2)
        self.conv1 = Conv2d(in_channels, out_channels, 3, 1, padding="same")
        self.conv2 = Conv2d(out_channels * 2, out_channels, 3, 1, padding="same")
        self.conv3 = Conv2d(out_channels, out_channels, 3, 1, padding="same")
        self.trans = Transformer(in_channels=out_channels, out_channels=out_channels, num_heads=att_heads, dpr=dpr)
        self.relu = nn.ReLU()
        self.dropout = nn.Dropout(0.3)

    def convs_nhwc(self, x, skip=None):
        x1 = self.conv1.nhwc(x, act="relu", ups=1)               # nearest x2 folded into the gather
        if skip is not None:
            x1 = self.conv2.nhwc(skip, act="relu", x2=x1)        # cat((skip, x1)) is virtual
        else:
            x1 = self.conv2.nhwc(x1, act="relu")
        x1 = self.conv3.nhwc(x1, act="relu")
        return ops.dropout(x1, self.dropout.p, self.training)

    def nhwc(self, x, skip):
        return self.trans.nhwc(self.convs_nhwc(x, skip))

    forward = _io(nhwc)


class DS_out(nn.Module):
    """FCT.py:191-206."""

    def __init__(self, in_channels, out_channels):
        super().__init__()
        self._out_channels = out_channels
        self.upsample = nn.Upsample(scale_factor=2)
        self.conv1 = Conv2d(in_channels, in_channels, 3, 1, padding="same")
        self.conv2 = Conv2d(in_channels, in_channels, 3, 1, padding="same")
        self.conv3 = Conv2d(in_channels, out_channels, 3, 1, padding="same")
        self.relu = nn.ReLU()
        self.sigmoid = nn.Sigmoid()

    def nhwc(self, x, final_act="sigmoid"):
        x1 = self.conv1.nhwc(x, act="relu", ups=1)
        x1 = self.conv2.nhwc(x1, act="relu")
        return self.conv3.nhwc(x1, act=final_act)

    forward = _io(nhwc)


class FCT(nn.Module):
    """FCT.py:210-254."""

    def __init__(self):
        super().__init__()
        att_heads = 2
        filters = [8, 16, 32, 64, 128, 64, 32, 16, 8]
        blocks = len(filters)
        stochastic_depth_rate = 0.0
        dpr = [x for x in np.linspace(0, stochastic_depth_rate, blocks)]
        self.drp_out = 0.3
        self._out_channels = 1
        self.scale_img = nn.AvgPool2d(2, 2)
        self.block_1 = Block_encoder_bottleneck("first", 3, filters[0], att_heads, dpr[0])
        self.block_2 = Block_encoder_bottleneck("second", filters[0], filters[1], att_heads, dpr[1])
        self.block_3 = Block_encoder_bottleneck("third", filters[1], filters[2], att_heads, dpr[2])
        self.block_4 = Block_encoder_bottleneck("fourth", filters[2], filters[3], att_heads, dpr[3])
        self.block_5 = Block_encoder_bottleneck("bottleneck", filters[3], filters[4], att_heads, dpr[4])
        self.block_6 = Block_decoder(filters[4], filters[5], att_heads, dpr[5])
        self.block_7 = Block_decoder(filters[5], filters[6], att_heads, dpr[6])
        self.block_8 = Block_decoder(filters[6], filters[7], att_heads, dpr[7])
        self.block_9 = Block_decoder(filters[7], filters[8], att_heads, dpr[8])
        self.ds = DS_out(filters[8], 1)

    def nhwc(self, x, final_act="sigmoid"):
        if x.shape[1] != x.shape[2] or x.shape[1] % 32:
            raise RuntimeError("FCT: square frames with side a multiple of 32 (FCT.py:77 reshapes by sqrt(N))")
        s2 = ops.avgpool2(x)
        s3 = ops.avgpool2(s2)
        s4 = ops.avgpool2(s3)
        x1 = self.block_1.nhwc(x)
        x2 = self.block_2.nhwc(x1, s2)
        x3 = self.block_3.nhwc(x2, s3)
        x4 = self.block_4.nhwc(x3, s4)
        y = self.block_5.nhwc(x4)
        y = self.block_6.nhwc(y, x4)
        y = self.block_7.nhwc(y, x3)
        y = self.block_8.nhwc(y, x2)
        y = self.block_9.nhwc(y, x1)
        return self.ds.nhwc(y, final_act)

    def forward(self, x):
        # sigmoid(conv3) (FCT.py:205) + the NCHW fp32 boundary in one pass; sigma' from the fp32 mask in backward
        return ops.head_nchw(self.nhwc(ops.to_nhwc(x), final_act="none"), 1, "sigmoid")
