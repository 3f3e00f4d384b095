"""Image -> latent -> reconstruction -> mask pipeline -- drop-in for the reference
``AE_Unet_Combined.py`` (AE_Unet_Combined.py:6-17)."""
import torch.nn as nn

from . import ops
from .AutoEncoder import AutoEncoder
from .UNet import UNet


class Image2Image2Mask(nn.Module):
    def __init__(self):
        super().__init__()
        self.image2imageAE = AutoEncoder()
        self.unet = UNet()

    def nhwc(self, x):
        latent, recon = self.image2imageAE.nhwc(x)
        # recon is [B,H,W,8] with channels 3..7 zero: exactly the padded RGB layout UNet's first conv takes
        return latent, recon, self.unet.nhwc(recon, final_act="none")

    def forward(self, x):
        latent, recon, z = self.nhwc(ops.to_nhwc(x))
        return ops.to_nchw(latent, 128), ops.to_nchw(recon, 3), ops.head_nchw(z, 1, "sigmoid")
