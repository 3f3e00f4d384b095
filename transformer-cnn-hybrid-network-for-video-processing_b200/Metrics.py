"""Losses / metrics of the reference (Metrics.py:5-42, Model.py:30,57) on the sm_100a kernels
(csrc/hn_loss.cu): one launch per forward (grid-wide sums, last CTA writes the device scalar), one launch
per backward (the dL/dy seed of the whole backward pass).  Same class names, constructor arguments, call
signature and return type (a 0-d tensor) as the reference, so ``loss = DiceLoss()(y_pred, y)``,
``loss.item()``, ``loss.backward()`` (FCT.py:333-337) work unchanged.  ``MixedLoss`` needs the third-party
``pytorch_msssim`` and stays out of scope; its L2 term is :class:`MSELoss`.
"""
import torch
import torch.nn as nn

from ._lib import call

_ws = {}


def _workspace(device):
    """float[16] scratch of the loss kernels (zero once; the kernels leave it zero)."""
    w = _ws.get(device)
    if w is None:
        w = _ws[device] = torch.zeros(16, dtype=torch.float32, device=device)
    return w


def _stream():
    return torch.cuda.current_stream().cuda_stream


def _check(y_pred, y_true):
    assert y_pred.size() == y_true.size()
    if not (y_pred.is_cuda and y_true.is_cuda):
        raise RuntimeError("hybridnet.Metrics: CUDA tensors only (there is no CPU fallback)")
    return y_pred.contiguous().float(), y_true.contiguous().float()


class _DiceFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, y_pred, y_true, smooth):
        B, C, H, W = y_pred.shape
        out = torch.empty(4, dtype=torch.float32, device=y_pred.device)     # {loss, sum(p t), sum p + sum t + smooth}
        call("hn_dice_fwd", y_pred.data_ptr(), y_true.data_ptr(), B, C, y_true.shape[1], H * W, smooth,
             _workspace(y_pred.device).data_ptr(), out.data_ptr(), _stream())
        ctx.smooth, ctx.shape = smooth, (B, C, H, W)
        ctx.save_for_backward(y_true, out)
        return out[0]

    @staticmethod
    def backward(ctx, g):
        y_true, out = ctx.saved_tensors
        B, C, H, W = ctx.shape
        d = torch.empty((B, C, H, W), dtype=torch.float32, device=y_true.device)
        g = g.contiguous().float()
        call("hn_dice_bwd", y_true.data_ptr(), out.data_ptr(), g.data_ptr(), d.data_ptr(), B, C, y_true.shape[1], H * W,
             ctx.smooth, _stream())
        return d, None, None


class DiceLoss(nn.Module):
    """Metrics.py:5-22."""

    def __init__(self, num_classes=8):
        super().__init__()
        self.smooth = 1.0
        self.classes = 3
        self.ignore_index = None
        self.eps = 1e-7

    def forward(self, y_pred, y_true):
        y_pred, y_true = _check(y_pred, y_true)
        if y_pred.shape[2] * y_pred.shape[3] % 4:
            raise RuntimeError("hybridnet.DiceLoss: H*W must be a multiple of 4")
        return _DiceFn.apply(y_pred, y_true, float(self.smooth))


class JaccardScore(nn.Module):
    """Metrics.py:26-39: logical and/or of the RAW tensors (any non-zero value is True), returns a tensor."""

    def forward(self, y_pred, y_true):
        y_pred, y_true = _check(y_pred.detach(), y_true.detach())
        B, C, H, W = y_pred.shape
        out = torch.empty((), dtype=torch.float32, device=y_pred.device)
        call("hn_jaccard", y_pred.data_ptr(), y_true.data_ptr(), B, C, y_true.shape[1], H * W,
             _workspace(y_pred.device).data_ptr(), out.data_ptr(), _stream())
        return out


class _MSEFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, a, b):
        out = torch.empty(2, dtype=torch.float32, device=a.device)
        call("hn_mse_fwd", a.data_ptr(), b.data_ptr(), a.numel(), _workspace(a.device).data_ptr(), out.data_ptr(), _stream())
        ctx.save_for_backward(a, b)
        psnr = out[1]
        ctx.mark_non_differentiable(psnr)
        return out[0], psnr

    @staticmethod
    def backward(ctx, g, _gp):
        a, b = ctx.saved_tensors
        da = torch.empty_like(a)
        g = g.contiguous().float()
        call("hn_mse_bwd", a.data_ptr(), b.data_ptr(), a.numel(), g.data_ptr(), da.data_ptr(), _stream())
        return da, None


class MSELoss(nn.Module):
    """nn.MSELoss() (the L2 term of MixedLoss, Metrics.py:53; Model.py:57).  ``last_psnr`` holds the PSNR of the same
    pass (Model.py:30,62-64: the trainer recomputes F.mse_loss for it)."""

    def __init__(self):
        super().__init__()
        self.last_psnr = None

    def forward(self, y_pred, y_true):
        y_pred, y_true = _check(y_pred, y_true)
        loss, self.last_psnr = _MSEFn.apply(y_pred, y_true.detach())
        return loss


def mask_iou(y_pred, y_true, threshold=0.5):
    """IoU of thresholded masks (what the parity tests mean by "mask IoU"); NOT the reference's JaccardScore."""
    p = y_pred[:, 0].reshape(-1) > threshold
    t = y_true[:, 0].reshape(-1) > threshold
    union = torch.logical_or(p, t).sum().item()
    return 1.0 if union == 0 else torch.logical_and(p, t).sum().item() / union
