"""Losses / metrics of the reference (Metrics.py:5-42) -- tiny global reductions kept as
torch ops around the hot path; MixedLoss needs pytorch_msssim and is
out of scope."""
import torch
import torch.nn as nn


class DiceLoss(nn.Module):
    def __init__(self, num_classes=8):
        super().__init__()
        self.smooth = 1.0
        self.classes = 3
        self.ignore_index = None
        self.eps = 1e-7

    def forward(self, y_pred, y_true):
        assert y_pred.size() == y_true.size()
        p = y_pred[:, 0].reshape(-1)
        t = y_true[:, 0].reshape(-1)
        inter = torch.dot(p, t)
        return 1.0 - (2.0 * inter + self.smooth) / (p.sum() + t.sum() + self.smooth)


class JaccardScore(nn.Module):
    def forward(self, y_pred, y_true):
        assert y_pred.size() == y_true.size()
        p = y_pred[:, 0].reshape(-1) > 0.5
        t = y_true[:, 0].reshape(-1) > 0.5
        union = torch.logical_or(p, t).sum().item()
        return 1.0 if union == 0 else torch.logical_and(p, t).sum().item() / union
