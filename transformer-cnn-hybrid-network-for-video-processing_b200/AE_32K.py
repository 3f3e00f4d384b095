"""Autoencoder32K family on the sm_100a kernels.  The reference ships these nets only as
bytecode (``__pycache__/AE_32K.cpython-38.pyc``; line numbers below are its lnotab source
lines): ResNet-bottleneck encoder -> [B,8,H/4,W/4] latent ("32K" at
256^2) -> conv / ConvTranspose decoder with an image (3ch, ReLU) or mask (1ch, sigmoid) head.
Attribute names are the bytecode's STORE_ATTR targets, i.e. the state_dict keys of
``saved_model/EncoderDecoder_road_image2image32K.tar`` / ``latent_to_mask.tar``.
"""
import torch.nn as nn

from . import ops
from .layers import BatchNorm2d, Conv2d, ConvTranspose2d, conv_bn


class Bottleneck(nn.Module):
    """pyc:AE_32K:L20-53 (torchvision-style, expansion 4)."""
    expansion = 4

    def __init__(self, in_channels, out_channels, stride=1, downsample=None):
        super().__init__()
        self.conv1 = Conv2d(in_channels, out_channels, kernel_size=1, bias=False)
        self.bn1 = BatchNorm2d(out_channels)
        self.conv2 = Conv2d(out_channels, out_channels, kernel_size=3, stride=stride, padding=1, bias=False)
        self.bn2 = BatchNorm2d(out_channels)
        self.conv3 = Conv2d(out_channels, out_channels * 4, kernel_size=1, bias=False)
        self.bn3 = BatchNorm2d(out_channels * 4)
        self.relu = nn.ReLU(inplace=True)
        self.downsample = downsample
        self.stride = stride
        self._out_channels = out_channels * 4

    def nhwc(self, x):
        out = conv_bn(self.conv1, self.bn1, x, act="relu")
        out = conv_bn(self.conv2, self.bn2, out, act="relu")
        residual = x
        if self.downsample is not None:
            residual = conv_bn(self.downsample[0], self.downsample[1], x, act="none")
        return conv_bn(self.conv3, self.bn3, out, act="relu", res=residual)   # relu(bn3(conv3) + residual)

    def forward(self, x):
        return ops.to_nchw(self.nhwc(ops.to_nhwc(x)), self._out_channels)


class Encoder_32K(nn.Module):
    """pyc:AE_32K:L57-115."""

    def __init__(self, block, layers):
        self.inplanes = 64
        super().__init__()
        self.conv1 = Conv2d(3, 64, kernel_size=7, stride=2, padding=3, bias=False)
        self.bn1 = BatchNorm2d(64)
        self.layer1 = self._make_layer(block, 64, layers[0])
        self.layer2 = self._make_layer(block, 128, layers[1], stride=2)
        self.conv2 = Conv2d(512, 128, kernel_size=3, stride=1, padding=1)
        self.bn2 = BatchNorm2d(128)
        self.conv3 = Conv2d(128, 64, kernel_size=3, stride=1, padding=1)
        self.bn3 = BatchNorm2d(64)
        self.conv4 = Conv2d(64, 16, kernel_size=3, stride=1, padding=1)
        self.bn4 = BatchNorm2d(16)
        self.conv5 = Conv2d(16, 8, kernel_size=3, stride=1, padding=1)
        self.bn5 = BatchNorm2d(8)
        self.relu = nn.ReLU(inplace=True)
        self.dropout = nn.Dropout2d(0.3)
        self.scale_img = nn.AvgPool2d(2, 2)        # unused in the reference too

    def _make_layer(self, block, planes, blocks, stride=1):
        downsample = None
        if stride != 1 or self.inplanes != planes * block.expansion:
            downsample = nn.Sequential(
                Conv2d(self.inplanes, planes * block.expansion, kernel_size=1, stride=stride, bias=False),
                BatchNorm2d(planes * block.expansion))
        layers = [block(self.inplanes, planes, stride, downsample)]
        self.inplanes = planes * block.expansion
        for _ in range(1, blocks):
            layers.append(block(self.inplanes, planes))
        return nn.Sequential(*layers)

    def _drop(self, x):
        return ops.dropout(x, self.dropout.p, self.training, channelwise=True)

    def nhwc(self, x):
        if x.shape[1] % 4 or x.shape[2] % 4:
            raise RuntimeError("Encoder_32K: H and W must be multiples of 4")
        x = conv_bn(self.conv1, self.bn1, x, act="relu")
        for blk in self.layer1:
            x = blk.nhwc(x)
        x = self._drop(x)
        for blk in self.layer2:
            x = blk.nhwc(x)
        x = self._drop(x)
        x = conv_bn(self.conv2, self.bn2, x, act="relu")
        x = conv_bn(self.conv3, self.bn3, x, act="relu")
        x = conv_bn(self.conv4, self.bn4, x, act="relu")
        if self.training and self.dropout.p > 0:      # relu(bn5(dropout(conv5(x)))), pyc L113
            return self.bn5.nhwc(self._drop(self.conv5.nhwc(x)), act="relu")
        return conv_bn(self.conv5, self.bn5, x, act="relu")

    def forward(self, x):
        return ops.to_nchw(self.nhwc(ops.to_nhwc(x)), 8)


class Decoder_32K(nn.Module):
    """pyc:AE_32K:L121-168."""

    def __init__(self, outputDeterminer):
        super().__init__()
        self.conv1 = Conv2d(8, 16, kernel_size=3, stride=1, padding=1)
        self.bn1 = BatchNorm2d(16)
        self.conv2 = Conv2d(16, 64, kernel_size=3, stride=1, padding=1)
        self.bn2 = BatchNorm2d(64)
        self.conv3 = Conv2d(64, 128, kernel_size=3, stride=1, padding=1)
        self.bn3 = BatchNorm2d(128)
        self.conv4 = Conv2d(128, 256, kernel_size=3, stride=1, padding=1)
        self.bn4 = BatchNorm2d(256)
        self.transConv1 = ConvTranspose2d(256, 384, kernel_size=2, stride=2, padding=0)
        self.dbn2 = BatchNorm2d(384)
        self.transConv2 = ConvTranspose2d(384, 192, kernel_size=2, stride=2, padding=0)
        self.dbn3 = BatchNorm2d(192)
        self.conv5 = Conv2d(192, 64, kernel_size=3, padding=1)
        self.bn5 = BatchNorm2d(64)
        self.conv6 = Conv2d(64, 8, kernel_size=3, padding=1)
        self.bn6 = BatchNorm2d(8)
        if outputDeterminer.lower() == "image":
            self.outputDeterminerConv = Conv2d(8, 3, kernel_size=3, padding=1)
            self.outputDeterminerNorm = BatchNorm2d(3)
            self.finalactivation = nn.ReLU()
        elif outputDeterminer.lower() == "mask":
            self.outputDeterminerConv = Conv2d(8, 1, kernel_size=3, padding=1)
            self.outputDeterminerNorm = BatchNorm2d(1)
            self.finalactivation = nn.Sigmoid()
        else:
            raise ValueError("outputDeterminer must be 'image' or 'mask'")
        self.relu = nn.ReLU()
        self._out_channels = self.outputDeterminerConv.out_channels

    def nhwc(self, x, final_act=None):
        x = conv_bn(self.conv1, self.bn1, x)
        x = conv_bn(self.conv2, self.bn2, x)
        x = conv_bn(self.conv3, self.bn3, x)
        x = conv_bn(self.conv4, self.bn4, x)
        x = conv_bn(self.transConv1, self.dbn2, x)
        x = conv_bn(self.transConv2, self.dbn3, x)
        x = conv_bn(self.conv5, self.bn5, x)
        x = conv_bn(self.conv6, self.bn6, x)
        act = final_act or ("sigmoid" if isinstance(self.finalactivation, nn.Sigmoid) else "relu")
        return conv_bn(self.outputDeterminerConv, self.outputDeterminerNorm, x, act=act)

    def _final_act(self):
        return "sigmoid" if isinstance(self.finalactivation, nn.Sigmoid) else "relu"

    def out_nchw(self, x):
        """NHWC latent -> NCHW fp32 output; the final activation rides the layout pass (ops.head_nchw)."""
        return ops.head_nchw(self.nhwc(x, final_act="none"), self._out_channels, self._final_act())

    def forward(self, x):
        return self.out_nchw(ops.to_nhwc(x))


class Autoencoder32K(nn.Module):
    """pyc:AE_32K:L174-188: ``forward`` returns the decoded tensor only."""

    def __init__(self, outputType):
        super().__init__()
        self.encoder = Encoder_32K(Bottleneck, [3, 4])
        self.decoder = Decoder_32K("image" if outputType.lower() == "image" else "mask")

    def nhwc(self, x):
        return self.decoder.nhwc(self.encoder.nhwc(x))

    def forward(self, x):
        return self.decoder.out_nchw(self.encoder.nhwc(ops.to_nhwc(x)))


class LatentToMask(nn.Sequential):
    """The latent-to-mask net of pyc:AE_32K:L232-245: ``nn.Sequential(frozen Encoder_32K,
    Decoder_32K('mask'))`` -- keys ``0.*`` / ``1.*`` as in ``saved_model/latent_to_mask.tar``.
    Backward stops at the latent because the encoder's parameters do not require grad."""

    def __init__(self, encoder=None, decoder=None):
        encoder = encoder if encoder is not None else Encoder_32K(Bottleneck, [3, 4])
        decoder = decoder if decoder is not None else Decoder_32K("mask")
        super().__init__(encoder, decoder)
        for p in self[0].parameters():
            p.requires_grad = False

    def nhwc(self, x):
        return self[1].nhwc(self[0].nhwc(x))

    def forward(self, x):
        return self[1].out_nchw(self[0].nhwc(ops.to_nhwc(x)))
