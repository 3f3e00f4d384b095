"""hybridnet: B200-native hot path of the hybrid segmentation nets."""
from . import _lib, data, ops, optim  # noqa: F401
from .AE_32K import Autoencoder32K, Bottleneck, Decoder_32K, Encoder_32K, LatentToMask  # noqa: F401
from .AE_Unet_Combined import Image2Image2Mask  # noqa: F401
from .AutoEncoder import (AutoEncoder, Decoder, DecoderBlock, DeepSupervisionBlock, Encoder,  # noqa: F401
                          EncoderBlock)
from .FCT import (FCT, Attention, Block_decoder, Block_encoder_bottleneck, DS_out, Transformer,  # noqa: F401
                  Wide_Focus)
from .Metrics import DiceLoss, JaccardScore, MSELoss  # noqa: F401
from .UNet import UNet  # noqa: F401

MODELS = {
    "fct": FCT, "unet": UNet, "autoencoder": AutoEncoder, "i2i2m": Image2Image2Mask,
    "ae32k_image": lambda: Autoencoder32K("image"), "ae32k_mask": lambda: Autoencoder32K("mask"),
    "latent2mask": LatentToMask,
}
