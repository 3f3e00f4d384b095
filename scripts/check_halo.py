"""Parity of the tcgen05 halo conv variant against torch.nn.functional.conv2d (fp32 on bf16-rounded operands)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch, torch.nn.functional as F
import hybridnet
from hybridnet import ops, _lib
lib = _lib.load()
dev = torch.device("cuda")
torch.manual_seed(0)
ok = True
for (B, H, W, C1, C2, N, dil) in [(2, 16, 128, 32, 0, 32, 1), (2, 8, 256, 32, 32, 32, 1), (1, 16, 128, 64, 0, 64, 1), (2, 8, 128, 32, 0, 64, 1),
                                   (1, 8, 128, 64, 0, 128, 1), (1, 8, 256, 64, 0, 64, 2)]:
    x1 = torch.randn(B, H, W, C1, device=dev).bfloat16()
    x2 = torch.randn(B, H, W, C2, device=dev).bfloat16() if C2 else None
    w = (torch.randn(N, C1 + C2, 3, 3, device=dev) * 0.1)
    bias = torch.randn(N, device=dev)
    wp, Kpad = ops._packed(w, "fwd", C1 + C2)
    names = []
    _lib.set_profiler(lambda name, args: (names.append(name), __import__("contextlib").nullcontext())[1])
    s0, s1 = torch.zeros(N, device=dev), torch.zeros(N, device=dev)
    y = ops._conv_raw(x1, x2, wp, Kpad, bias, None, OH=H, OW=W, Cout=N, KH=3, KW=3, stride=1, pad=dil, dil=dil, stats=(s0, s1))
    _lib.set_profiler(None)
    torch.cuda.synchronize()
    xin = torch.cat([x1, x2], 3) if C2 else x1
    ref = F.conv2d(xin.float().permute(0, 3, 1, 2), w.bfloat16().float(), bias, padding=dil, dilation=dil).permute(0, 2, 3, 1)
    err = float((y.float() - ref).abs().max()); rel = float((y.float() - ref).norm() / ref.norm())
    serr = float((s0 - ref.bfloat16().float().sum((0, 1, 2))).abs().max() / (ref.abs().sum((0, 1, 2)).max()))
    good = rel < 5e-3 and serr < 1e-3
    ok &= good
    print(f"B{B} {H}x{W} C{C1}+{C2}->{N} d{dil}: {names[-1]} halo={lib.hn_conv_tc_halo_supported(ops._desc(B=B,H=H,W=W,C1=C1,C2=C2,ld1=C1,ld2=C2,OH=H,OW=W,Cout=N,ldy=N,KH=3,KW=3,stride=1,pad=dil,dil=dil,mode=0,ups=0,act=0,pshuf=0,ldr=0,Kpad=Kpad))} max|err|={err:.3e} rel={rel:.3e} stats_rel={serr:.2e} {'OK' if good else 'FAIL'}")
print("ALL OK" if ok else "FAILED")
