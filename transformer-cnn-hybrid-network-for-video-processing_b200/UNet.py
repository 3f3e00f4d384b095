"""BatchNorm U-Net on the sm_100a kernels -- drop-in for the reference ``UNet.py``
(UNet.py:6-66): same ctor ``UNet(in_channels=3, out_channels=1, init_features=32)``,
same OrderedDict child names (``encoder1.enc1conv1.weight`` ...), same forward.
"""
from collections import OrderedDict

import torch.nn as nn

from . import ops
from .layers import BatchNorm2d, Conv2d, ConvTranspose2d, conv_bn


class UNet(nn.Module):
    def __init__(self, in_channels=3, out_channels=1, init_features=32):
        super().__init__()
        f = init_features
        self._out_channels = out_channels
        # construction order == UNet.py:11-28 (default-init RNG parity)
        self.encoder1 = UNet._block(in_channels, f, name="enc1")
        self.pool1 = nn.MaxPool2d(kernel_size=2, stride=2)
        self.encoder2 = UNet._block(f, f * 2, name="enc2")
        self.pool2 = nn.MaxPool2d(kernel_size=2, stride=2)
        self.encoder3 = UNet._block(f * 2, f * 4, name="enc3")
        self.pool3 = nn.MaxPool2d(kernel_size=2, stride=2)
        self.encoder4 = UNet._block(f * 4, f * 8, name="enc4")
        self.pool4 = nn.MaxPool2d(kernel_size=2, stride=2)
        self.bottleneck = UNet._block(f * 8, f * 16, name="bottleneck")
        self.upconv4 = ConvTranspose2d(f * 16, f * 8, kernel_size=2, stride=2)
        self.decoder4 = UNet._block((f * 8) * 2, f * 8, name="dec4")
        self.upconv3 = ConvTranspose2d(f * 8, f * 4, kernel_size=2, stride=2)
        self.decoder3 = UNet._block((f * 4) * 2, f * 4, name="dec3")
        self.upconv2 = ConvTranspose2d(f * 4, f * 2, kernel_size=2, stride=2)
        self.decoder2 = UNet._block((f * 2) * 2, f * 2, name="dec2")
        self.upconv1 = ConvTranspose2d(f * 2, f, kernel_size=2, stride=2)
        self.decoder1 = UNet._block(f * 2, f, name="dec1")
        self.conv = Conv2d(in_channels=f, out_channels=out_channels, kernel_size=1)

    @staticmethod
    def _block(in_channels, features, name):
        return nn.Sequential(OrderedDict([
            (name + "conv1", Conv2d(in_channels=in_channels, out_channels=features, kernel_size=3, padding=1, bias=False)),
            (name + "norm1", BatchNorm2d(num_features=features)),
            (name + "relu1", nn.ReLU(inplace=True)),
            (name + "conv2", Conv2d(in_channels=features, out_channels=features, kernel_size=3, padding=1, bias=False)),
            (name + "norm2", BatchNorm2d(num_features=features)),
            (name + "relu2", nn.ReLU(inplace=True)),
        ]))

    @staticmethod
    def _run_block(block, x, x2=None):
        """[conv3x3 -> BN(batch stats) -> ReLU] x2 with BN statistics from the conv epilogue."""
        c1, n1, _, c2, n2, _ = block
        y = conv_bn(c1, n1, x, act="relu", x2=x2)
        return conv_bn(c2, n2, y, act="relu")

    def nhwc(self, x, final_act="sigmoid"):
        if x.shape[1] % 16 or x.shape[2] % 16:
            raise RuntimeError("UNet: H and W must be multiples of 16 (four 2x2 poolings)")
        rb = UNet._run_block
        enc1 = rb(self.encoder1, x)
        enc2 = rb(self.encoder2, ops.maxpool2(enc1))
        enc3 = rb(self.encoder3, ops.maxpool2(enc2))
        enc4 = rb(self.encoder4, ops.maxpool2(enc3))
        y = rb(self.bottleneck, ops.maxpool2(enc4))
        for up, dec, skip in ((self.upconv4, self.decoder4, enc4), (self.upconv3, self.decoder3, enc3),
                              (self.upconv2, self.decoder2, enc2), (self.upconv1, self.decoder1, enc1)):
            y = rb(dec, up.nhwc(y), x2=skip)              # cat((up, skip), 1) is virtual
        return self.conv.nhwc(y, act=final_act)

    def forward(self, x):
        # sigmoid + NHWC bf16 -> NCHW fp32 in one pass; sigma' in backward from the fp32 mask (ops.head_nchw)
        return ops.head_nchw(self.nhwc(ops.to_nhwc(x), final_act="none"), self._out_channels, "sigmoid")
