"""Micro-benchmark of the fused q/k/v projection kernels (depthwise 3x3 + ReLU + LayerNorm, FCT.py:41-57) through the C
ABI at the FCT B=32 shapes: forward, LN/ReLU backward, depthwise data+weight gradient.
  python scripts/dwln_bench.py [--lib path.so] [--reps N]
Prints sums of the results too, so that two libraries can be compared on the same seeded inputs."""
import argparse, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

SHAPES = [(32, 128, 128, 8), (32, 64, 64, 16), (32, 32, 32, 32), (32, 16, 16, 64), (32, 8, 8, 128)]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--lib", default="")
    ap.add_argument("--reps", type=int, default=20)
    a = ap.parse_args()
    from hybridnet import _lib, ops
    if a.lib:
        _lib.LIB_PATH = os.path.abspath(a.lib)
        print("library:", a.lib)
    _lib.load()
    call, p3 = ops.call, ops._p3
    dev = torch.device("cuda")
    torch.manual_seed(0)
    print(f"{'shape':18s} {'fwd us':>9s} {'bwd_pre us':>11s} {'bwd_dw us':>10s} {'MB moved (fwd/pre/dw)':>24s}")
    for B, H, W, C in SHAPES:
        byt = 2.0 * B * H * W * C
        nbuf = max(2, int(300e6 // (7 * byt)) + 1)
        xs = [torch.randn(B, H, W, C, device=dev).bfloat16() for _ in range(nbuf)]
        dqkv = [torch.randn(3, B, H, W, C, device=dev).bfloat16() for _ in range(nbuf)]
        qkv = [torch.empty(3, B, H, W, C, device=dev, dtype=torch.bfloat16) for _ in range(nbuf)]
        dpre = [torch.empty(3, B, H, W, C, device=dev, dtype=torch.bfloat16) for _ in range(nbuf)]
        dx = [torch.empty(B, H, W, C, device=dev, dtype=torch.bfloat16) for _ in range(nbuf)]
        w = [torch.randn(C, 1, 3, 3, device=dev) * 0.3 for _ in range(3)]
        b = [torch.randn(C, device=dev) * 0.1 for _ in range(3)]
        g = [torch.rand(C, device=dev) + 0.5 for _ in range(3)]
        be = [torch.randn(C, device=dev) * 0.1 for _ in range(3)]
        dw = [torch.zeros(C, 1, 3, 3, device=dev) for _ in range(3)]
        db = [torch.zeros(C, device=dev) for _ in range(3)]
        dg = [torch.zeros(C, device=dev) for _ in range(3)]
        dbe = [torch.zeros(C, device=dev) for _ in range(3)]
        st = ops._stream()

        def fwd(j):
            call("hn_dwconv_relu_ln_fwd", xs[j].data_ptr(), qkv[j][0].data_ptr(), qkv[j][1].data_ptr(), qkv[j][2].data_ptr(),
                 p3(*w), p3(*b), p3(*g), p3(*be), B, H, W, C, C, C, 1e-5, st)

        def pre(j):
            call("hn_dwconv_relu_ln_bwd_pre", xs[j].data_ptr(), dqkv[j][0].data_ptr(), dqkv[j][1].data_ptr(),
                 dqkv[j][2].data_ptr(), dpre[j][0].data_ptr(), dpre[j][1].data_ptr(), dpre[j][2].data_ptr(), p3(*w), p3(*b),
                 p3(*g), p3(*dg), p3(*dbe), p3(*db), B, H, W, C, C, C, 1e-5, st)

        def bdw(j):
            call("hn_dwconv_bwd_data_weight", xs[j].data_ptr(), dpre[j][0].data_ptr(), dpre[j][1].data_ptr(),
                 dpre[j][2].data_ptr(), dx[j].data_ptr(), p3(*w), p3(*dw), B, H, W, C, C, C, C, st)

        res = []
        for f in (fwd, pre, bdw):
            for i in range(nbuf):
                f(i)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for i in range(a.reps):
                f(i % nbuf)
            e1.record()
            torch.cuda.synchronize()
            res.append(e0.elapsed_time(e1) / a.reps * 1e3)
        mb = (4 * byt / 1e6, 7 * byt / 1e6, 5 * byt / 1e6)
        print(f"{str((B, H, W, C)):18s} {res[0]:9.1f} {res[1]:11.1f} {res[2]:10.1f}   {mb[0]:6.1f} {mb[1]:6.1f} {mb[2]:6.1f}",
              flush=True)
        # fingerprints of the results (compare two libraries by eye: bf16 outputs are deterministic, the atomics' sums to ~1e-6)
        print("   sums:", " ".join(f"{t.float().sum().item():.6g}" for t in (qkv[0], dpre[0], dx[0])),
              " dw/dg:", f"{dw[0].double().sum().item():.6g} {dg[0].double().sum().item():.6g}", flush=True)


if __name__ == "__main__":
    main()
