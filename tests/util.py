import torch


def rel_l2(a, b):
    a, b = a.double().flatten(), b.double().flatten()
    return float((a - b).norm() / (b.norm() + 1e-12))


def max_abs(a, b):
    return float((a.double() - b.double()).abs().max())


def cosine(a, b):
    a, b = a.double().flatten(), b.double().flatten()
    return float(torch.dot(a, b) / (a.norm() * b.norm() + 1e-30))


def bf16r(t):
    """Round to bf16 and back: makes operands exactly representable so that the check
    measures the kernel (fp32 accumulation) and not the input quantisation."""
    return t.to(torch.bfloat16).float()


def nhwc_bf16(x_nchw):
    """NCHW fp32 -> NHWC bf16 with channels padded to 8 (pure torch, test side only)."""
    B, C, H, W = x_nchw.shape
    Cp = (C + 7) // 8 * 8
    out = torch.zeros(B, H, W, Cp, dtype=torch.bfloat16, device=x_nchw.device)
    out[..., :C] = x_nchw.permute(0, 2, 3, 1).to(torch.bfloat16)
    return out


def nchw_f32(x_nhwc, C):
    return x_nhwc[..., :C].float().permute(0, 3, 1, 2).contiguous()


def iou(a, b, thr=0.5):
    a, b = a > thr, b > thr
    u = (a | b).sum().item()
    return 1.0 if u == 0 else (a & b).sum().item() / u
