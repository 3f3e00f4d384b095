"""Conv AutoEncoder (8K latent, 128x8x8 at 256^2) on the sm_100a kernels -- drop-in
for the reference ``AutoEncoder.py`` (AutoEncoder.py:5-134)."""
import torch.nn as nn

from . import ops
from .FCT import _io
from .layers import Conv2d


class EncoderBlock(nn.Module):
    """AutoEncoder.py:5-27 (Block_encoder_bottleneck without the Transformer)."""

    def __init__(self, blk, in_channels, out_channels):
        super().__init__()
        self.blk = blk
        self._out_channels = out_channels
        self.conv1_a = Conv2d(in_channels, out_channels, 3, 1, padding="same")
        self.conv1_b = Conv2d(3, in_channels, 3, 1, padding="same")
        self.conv2 = Conv2d(out_channels, out_channels, 3, 1, padding="same")
        self.conv3 = Conv2d(out_channels, out_channels, 3, 1, padding="same")
        self.relu = nn.ReLU()
        self.dropout = nn.Dropout(p=0.3)
        self.maxpool = nn.MaxPool2d(kernel_size=(2, 2))

    def nhwc(self, x, scale_img="none"):
        if self.blk in ("first", "bottleneck"):
            x1 = self.conv1_a.nhwc(x, act="relu")
            x1 = self.conv2.nhwc(x1, act="relu")
        else:
            skip_x = self.conv1_b.nhwc(scale_img, act="relu")
            x1 = self.conv2.nhwc(skip_x, act="relu", x2=x)
            x1 = self.conv3.nhwc(x1, act="relu")
        return ops.maxpool2(ops.dropout(x1, self.dropout.p, self.training))

    forward = _io(nhwc)


class DecoderBlock(nn.Module):
    """AutoEncoder.py:32-48: Upsample x2 -> 3 x (conv3x3 + ReLU) -> Dropout; no skip."""

    def __init__(self, in_channels, out_channels):
        super().__init__()
        self._out_channels = out_channels
        self.upsample = nn.Upsample(scale_factor=2)
        self.conv1 = Conv2d(in_channels, out_channels, 3, 1, padding="same")
        self.conv2 = Conv2d(out_channels, out_channels, 3, 1, padding="same")
        self.conv3 = Conv2d(out_channels, out_channels, 3, 1, padding="same")
        self.relu = nn.ReLU()
        self.dropout = nn.Dropout(0.3)

    def nhwc(self, x):
        x1 = self.conv1.nhwc(x, act="relu", ups=1)
        x1 = self.conv2.nhwc(x1, act="relu")
        x1 = self.conv3.nhwc(x1, act="relu")
        return ops.dropout(x1, self.dropout.p, self.training)

    forward = _io(nhwc)


class DeepSupervisionBlock(nn.Module):
    """AutoEncoder.py:53-68; the head activation is ReLU (AutoEncoder.py:67)."""

    def __init__(self, in_channels, out_channels):
        super().__init__()
        self._out_channels = out_channels
        self.upsample = nn.Upsample(scale_factor=2)
        self.conv1 = Conv2d(in_channels, in_channels, 3, 1, padding="same")
        self.conv2 = Conv2d(in_channels, in_channels, 3, 1, padding="same")
        self.conv3 = Conv2d(in_channels, out_channels, 3, 1, padding="same")
        self.relu = nn.ReLU()
        self.sigmoid = nn.Sigmoid()

    def nhwc(self, x):
        x1 = self.conv1.nhwc(x, act="relu", ups=1)
        x1 = self.conv2.nhwc(x1, act="relu")
        return self.conv3.nhwc(x1, act="relu")

    forward = _io(nhwc)


class Encoder(nn.Module):
    """AutoEncoder.py:73-99."""

    def __init__(self):
        super().__init__()
        filters = [8, 16, 32, 64, 128]
        self.drp_out = 0.3
        self._out_channels = filters[-1]
        self.scale_img = nn.AvgPool2d(2, 2)
        self.block_1 = EncoderBlock("first", 3, filters[0])
        self.block_2 = EncoderBlock("second", filters[0], filters[1])
        self.block_3 = EncoderBlock("third", filters[1], filters[2])
        self.block_4 = EncoderBlock("fourth", filters[2], filters[3])
        self.block_5 = EncoderBlock("bottleneck", filters[3], filters[4])

    def nhwc(self, x):
        if x.shape[1] % 32 or x.shape[2] % 32:
            raise RuntimeError("AutoEncoder: H and W must be multiples of 32 (five 2x2 poolings)")
        s2 = ops.avgpool2(x)
        s3 = ops.avgpool2(s2)
        s4 = ops.avgpool2(s3)          # the reference's scale_img_5 is dead code (AutoEncoder.py:92)
        x1 = self.block_1.nhwc(x)
        x2 = self.block_2.nhwc(x1, s2)
        x3 = self.block_3.nhwc(x2, s3)
        x4 = self.block_4.nhwc(x3, s4)
        return self.block_5.nhwc(x4)

    forward = _io(nhwc)


class Decoder(nn.Module):
    """AutoEncoder.py:103-121."""

    def __init__(self):
        super().__init__()
        filters = [128, 64, 32, 16, 8]
        self.drp_out = 0.3
        self._out_channels = 3
        self.block_4 = DecoderBlock(filters[0], filters[1])
        self.block_3 = DecoderBlock(filters[1], filters[2])
        self.block_2 = DecoderBlock(filters[2], filters[3])
        self.block_1 = DecoderBlock(filters[3], filters[4])
        self.ds = DeepSupervisionBlock(filters[4], 3)

    def nhwc(self, x):
        x = self.block_4.nhwc(x)
        x = self.block_3.nhwc(x)
        x = self.block_2.nhwc(x)
        x = self.block_1.nhwc(x)
        return self.ds.nhwc(x)

    forward = _io(nhwc)


class AutoEncoder(nn.Module):
    """AutoEncoder.py:125-134: returns the (latent, output) tuple."""

    def __init__(self):
        super().__init__()
        self.encoder = Encoder()
        self.decoder = Decoder()

    def nhwc(self, x):
        latent = self.encoder.nhwc(x)
        return latent, self.decoder.nhwc(latent)

    def forward(self, x):
        latent, output = self.nhwc(ops.to_nhwc(x))
        return ops.to_nchw(latent, 128), ops.to_nchw(output, 3)
