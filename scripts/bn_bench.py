"""Micro-benchmark of the BatchNorm passes through the C ABI (each launch misses L2: buffers are rotated).
  python scripts/bn_bench.py [--reps N]"""
import argparse, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import hybridnet
from hybridnet import _lib

GEOMS = [("ae_192@256", 16 * 256 * 256, 192), ("ae_64@256", 16 * 256 * 256, 64), ("ae_384@128", 16 * 128 * 128, 384),
         ("ae_256@128", 16 * 128 * 128, 256), ("ae_64@128", 16 * 128 * 128, 64), ("ae_512@64", 16 * 64 * 64, 512),
         ("un_32@256", 32 * 256 * 256, 32), ("un_512@16", 32 * 16 * 16, 512)]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--reps", type=int, default=10)
    ap.add_argument("--only", default="")
    a = ap.parse_args()
    dev = torch.device("cuda")
    st = torch.cuda.current_stream().cuda_stream
    print(f"{'geometry':12s} {'kernel':24s} {'us':>8s} {'GB/s':>7s}")
    for tag, P, C in GEOMS:
        if a.only and a.only not in tag:
            continue
        nbuf = max(2, int(500e6 // (2 * P * C)) + 1)
        xs = [torch.randn(P, C, device=dev).bfloat16() for _ in range(nbuf)]
        dys = [torch.randn(P, C, device=dev).bfloat16() for _ in range(nbuf)]
        out = torch.empty(P, C, device=dev, dtype=torch.bfloat16)
        mean, rstd, gamma, beta = (torch.rand(C, device=dev) + 0.5 for _ in range(4))
        mean -= 1.0
        s1, s2 = torch.zeros(C, device=dev), torch.zeros(C, device=dev)
        p = lambda t: t.data_ptr()
        runs = {
            "hn_bn_stats": (1, lambda i: _lib.call("hn_bn_stats", p(xs[i]), P, C, C, p(s1), p(s2), st)),
            "hn_bn_apply": (2, lambda i: _lib.call("hn_bn_apply", p(xs[i]), None, p(out), p(mean), p(rstd), p(gamma), p(beta), P, C, C, 0, C, 1, st)),
            "hn_bn_relu_bwd_reduce": (2, lambda i: _lib.call("hn_bn_relu_bwd_reduce", p(dys[i]), p(xs[i]), p(mean), p(rstd), p(gamma), p(beta), P, C, C, C, p(s1), p(s2), st)),
            "hn_bn_relu_bwd_apply": (3, lambda i: _lib.call("hn_bn_relu_bwd_apply", p(dys[i]), p(xs[i]), p(out), p(mean), p(rstd), p(gamma), p(beta), p(s1), p(s2), P, C, C, C, C, 1, st)),
            "hn_bn_bwd_reduce": (3, lambda i: _lib.call("hn_bn_bwd_reduce", p(dys[i]), None, p(xs[i]), None, p(mean), p(rstd), P, C, C, C, C, C, 0, p(s1), p(s2), st)),
        }
        for name, (passes, fn) in runs.items():
            if name == "hn_bn_bwd_reduce":
                passes = 2
            for i in range(2):
                fn(i % nbuf)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize()
            e0.record()
            for i in range(a.reps):
                fn(i % nbuf)
            e1.record()
            torch.cuda.synchronize()
            us = e0.elapsed_time(e1) * 1e3 / a.reps
            print(f"{tag:12s} {name:24s} {us:8.1f} {2.0 * P * C * passes / us / 1e3:7.0f}")


if __name__ == "__main__":
    main()
