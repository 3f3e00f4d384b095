"""Device-side input pipeline: what the reference does on the host right before
``model(x)`` -- ``ToTensor`` (Dataloader.py:16-24: uint8 HWC -> float CHW / 255) and the trainer's
augmentations (Model.py:42-53: batch hflip, N(0, 0.05^2) noise, up to three 16x16 blackout patches) -- as ONE
kernel (``hn_augment_pack``) that writes the batch straight into the first conv's operand layout (NHWC bf16).

The random *decisions* are drawn on the host with ``random`` in the reference's order, so a run seeded like the
reference takes the same branches; the noise *values* come from a counter-based generator on the device.
"""
import ctypes
import random

import torch

from . import ops
from ._lib import call


def draw_augmentation(height, width, rng=random):
    """The branch decisions of Model.py:44-53, in the reference's order of ``random`` calls.
    Returns (flip, noise_std, patches) with patches = [(x, y), ...] (x indexes H, y indexes W)."""
    flip = rng.random() > 0.5
    noise_std, patches = 0.0, []
    if rng.random() > 0.8:
        noise_std = 0.05
        if rng.random() > 0.5:
            for _ in range(rng.randint(0, 3)):
                x = rng.randint(0, height - 16)
                y = rng.randint(0, width - 16)
                patches.append((x, y))
    return flip, noise_std, patches


def augment_pack(frames, flip=False, noise_std=0.0, patches=(), seed=0, want_nchw=False, patch_size=16):
    """frames: uint8 [B,H,W,3] (decoded images as the loader holds them) or fp32 [B,3,H,W] in [0,1], on the GPU.
    Returns ``ops.PackedFrames`` (accepted by every model forward) and, with ``want_nchw``, also the augmented
    batch as fp32 [B,3,H,W] (what the reference's trainer calls ``aug_image``)."""
    if not frames.is_cuda:
        raise RuntimeError("hybridnet.data.augment_pack: CUDA tensors only (there is no CPU fallback)")
    frames = frames.contiguous()
    if frames.dtype == torch.uint8:
        B, H, W, C = frames.shape
        u8 = 1
    elif frames.dtype == torch.float32:
        B, C, H, W = frames.shape
        u8 = 0
    else:
        raise RuntimeError(f"hybridnet.data.augment_pack: uint8 HWC or fp32 CHW frames, got {frames.dtype}")
    if C != 3:
        raise RuntimeError("hybridnet.data.augment_pack: RGB frames")
    patches = list(patches)
    if len(patches) > 4:
        raise RuntimeError("hybridnet.data.augment_pack: at most 4 patches (the reference draws 0..3)")
    nhwc = torch.empty((B, H, W, 8), dtype=torch.bfloat16, device=frames.device)
    nchw = torch.empty((B, 3, H, W), dtype=torch.float32, device=frames.device) if want_nchw else None
    px = (ctypes.c_int * 4)(*([p[0] for p in patches] + [0] * (4 - len(patches))))
    py = (ctypes.c_int * 4)(*([p[1] for p in patches] + [0] * (4 - len(patches))))
    ctr = ops.step_counter(frames.device)
    call("hn_augment_pack", frames.data_ptr(), u8, nhwc.data_ptr(), nchw.data_ptr() if want_nchw else 0, B, H, W, 8,
         int(bool(flip)), float(noise_std), int(seed) & ((1 << 62) - 1), ctr.data_ptr(), len(patches), px, py,
         int(patch_size), torch.cuda.current_stream().cuda_stream)
    packed = ops.PackedFrames(nhwc, 3)
    return (packed, nchw) if want_nchw else packed
