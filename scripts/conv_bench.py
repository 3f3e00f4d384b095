"""Micro-benchmark of single conv launches through the C ABI (forward-type calls and weight gradients).
  python scripts/conv_bench.py [--only SUBSTR[,SUBSTR...]] [--reps N]
Rotates over enough input/output buffers that every launch misses the 126 MB L2."""
import argparse, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import hybridnet
from hybridnet import ops, _lib

# (tag, B, H, W, C1, C2, Cout, k, stride, dil, pshuf, stats)
GEOMS = [
    ("ae_1x1_64_256", 16, 128, 128, 64, 0, 256, 1, 1, 1, 0, 1),
    ("ae_1x1_256_64", 16, 128, 128, 256, 0, 64, 1, 1, 1, 0, 1),
    ("ae_3x3_64_64", 16, 128, 128, 64, 0, 64, 3, 1, 1, 0, 1),
    ("ae_1x1_128_512", 16, 64, 64, 128, 0, 512, 1, 1, 1, 0, 1),
    ("ae_3x3_128_128", 16, 64, 64, 128, 0, 128, 3, 1, 1, 0, 1),
    ("ae_convT_384_192", 16, 128, 128, 384, 0, 192, 1, 1, 1, 192, 1),
    ("ae_3x3_192_64_256", 16, 256, 256, 192, 0, 64, 3, 1, 1, 0, 1),
    ("ae_3x3_64_192_256", 16, 256, 256, 64, 0, 192, 3, 1, 1, 0, 0),
    ("ae_k2s2_192_384", 16, 256, 256, 192, 0, 384, 2, 2, 1, 0, 0),
    ("ae_3x3_8_64_256", 16, 256, 256, 8, 0, 64, 3, 1, 1, 0, 0),
    ("ae_3x3_64_8_256", 16, 256, 256, 64, 0, 8, 3, 1, 1, 0, 1),
    ("un_3x3_32_32", 32, 256, 256, 32, 0, 32, 3, 1, 1, 0, 1),
    ("un_3x3_32+32_32", 32, 256, 256, 32, 32, 32, 3, 1, 1, 0, 1),
    ("un_3x3_64_64", 32, 128, 128, 64, 0, 64, 3, 1, 1, 0, 1),
    ("un_3x3_128_128", 32, 64, 64, 128, 0, 128, 3, 1, 1, 0, 1),
    ("un_3x3_256_256", 32, 32, 32, 256, 0, 256, 3, 1, 1, 0, 1),
    ("un_3x3_512_512", 32, 16, 16, 512, 0, 512, 3, 1, 1, 0, 1),
    ("fct_3x3_8_8", 32, 256, 256, 8, 0, 8, 3, 1, 1, 0, 0),
    ("fct_3x3_16_16", 32, 128, 128, 16, 0, 16, 3, 1, 1, 0, 0),
    ("fct_3x3_8_8_d3", 32, 128, 128, 8, 0, 8, 3, 1, 3, 0, 0),
    ("fct_1x1_8_8", 32, 128, 128, 8, 0, 8, 1, 1, 1, 0, 0),
    ("fct_3x3_32_32", 32, 64, 64, 32, 0, 32, 3, 1, 1, 0, 0),
    ("fct_3x3_8_1", 32, 256, 256, 8, 0, 1, 3, 1, 1, 0, 0),
    ("fct_s32_3x3_32_32", 32, 32, 32, 32, 0, 32, 3, 1, 1, 0, 0),
    ("fct_s16_3x3_64_64", 32, 16, 16, 64, 0, 64, 3, 1, 1, 0, 0),
]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--only", default="")
    ap.add_argument("--reps", type=int, default=20)
    ap.add_argument("--wgrad", action="store_true")
    ap.add_argument("--nostats", action="store_true")
    ap.add_argument("--graph", action="store_true", help="time the launches inside one CUDA graph")
    ap.add_argument("--nosc", action="store_true", help="skip the small-channel kernels (what the other paths make of the same layer)")
    ap.add_argument("--lib", default="", help="a variant library built by scripts/attn_variants.sh")
    a = ap.parse_args()
    if a.lib:
        _lib.LIB_PATH = os.path.abspath(a.lib)
        print("library:", a.lib)
    lib = _lib.load()
    if a.nosc:
        ops.USE_SC = False
    dev = torch.device("cuda")
    torch.manual_seed(0)
    print(f"{'geometry':22s} {'kernel':18s} {'us':>9s} {'TFLOP/s':>9s} {'GB/s':>8s}")
    for tag, B, H, W, C1, C2, Cout, k, s, dil, pshuf, st in GEOMS:
        if a.only and not any(o in tag for o in a.only.split(",")):
            continue
        OH, OW = H // s, W // s
        st = st and not a.nostats
        n = 4 * pshuf if pshuf else Cout
        pad = 0 if (k == 1 or s == 2) else dil
        Kpad = lib.hn_conv_kpad(k, k, C1 + C2)
        wp = (torch.randn(lib.hn_conv_npad(n), Kpad, device=dev) * 0.05).bfloat16()
        byt = 2.0 * B * (H * W * (C1 + C2) + OH * OW * n)
        nbuf = max(2, int(400e6 // byt) + 1)
        xs = [torch.randn(B, H, W, C1, device=dev).bfloat16() for _ in range(nbuf)]
        x2 = [torch.randn(B, H, W, C2, device=dev).bfloat16() for _ in range(nbuf)] if C2 else None
        oh, ow, oc = (2 * OH, 2 * OW, pshuf) if pshuf else (OH, OW, (Cout + 7) // 8 * 8)
        outs = [torch.empty(B, oh, ow, oc, device=dev, dtype=torch.bfloat16) for _ in range(nbuf)]
        stats = (torch.zeros(max(oc, 8), device=dev), torch.zeros(max(oc, 8), device=dev)) if st else None
        flops = 2.0 * B * OH * OW * n * k * k * (C1 + C2)
        names = []
        prof = lambda name, args: (names.append(name), __import__("contextlib").nullcontext())[1]
        _lib.set_profiler(prof)

        def run(i):
            j = i % nbuf
            if a.wgrad:
                dw = torch.zeros(Cout, C1 + C2, k, k, device=dev)
                ops._wgrad_raw(xs[j], x2[j] if x2 else None, outs[j], dw, Creal=C1 + C2, OH=OH, OW=OW, Cout=Cout, KH=k, KW=k,
                               stride=s, pad=pad, dil=dil, ups=0, strides=dw.stride())
            else:
                ops._conv_raw(xs[j], x2[j] if x2 else None, wp, Kpad, None, None, OH=OH, OW=OW, Cout=Cout, KH=k, KW=k, stride=s,
                              pad=pad, dil=dil, pshuf=pshuf, stats=stats, out=outs[j])
        if a.wgrad and pshuf:
            continue
        for i in range(3):
            run(i)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        if a.wgrad and a.graph:    # accumulate into one dW inside a CUDA graph (no zero-fill, no host launch floor)
            dwg = torch.zeros(Cout, C1 + C2, k, k, device=dev)
            gr = torch.cuda.CUDAGraph()
            with torch.cuda.graph(gr):
                for i in range(a.reps):
                    j = i % nbuf
                    ops._wgrad_raw(xs[j], x2[j] if x2 else None, outs[j], dwg, Creal=C1 + C2, OH=OH, OW=OW, Cout=Cout, KH=k,
                                   KW=k, stride=s, pad=pad, dil=dil, ups=0, strides=dwg.stride())
            gr.replay()
            torch.cuda.synchronize()
            e0.record()
            gr.replay()
            e1.record()
            torch.cuda.synchronize()
            us = e0.elapsed_time(e1) / a.reps * 1e3
        elif a.wgrad:     # exclude the zero-fill: time each launch separately
            tot = 0.0
            for i in range(a.reps):
                j = i % nbuf
                dw = torch.zeros(Cout, C1 + C2, k, k, device=dev)
                e0.record()
                ops._wgrad_raw(xs[j], x2[j] if x2 else None, outs[j], dw, Creal=C1 + C2, OH=OH, OW=OW, Cout=Cout, KH=k, KW=k,
                               stride=s, pad=pad, dil=dil, ups=0, strides=dw.stride())
                e1.record()
                torch.cuda.synchronize()
                tot += e0.elapsed_time(e1)
            us = tot / a.reps * 1e3
        elif a.graph:     # the launches of one CUDA graph: no host launch floor (~20 us per call from Python) between them
            gr = torch.cuda.CUDAGraph()
            with torch.cuda.graph(gr):
                for i in range(a.reps):
                    run(i)
            gr.replay()
            torch.cuda.synchronize()
            e0.record()
            gr.replay()
            e1.record()
            torch.cuda.synchronize()
            us = e0.elapsed_time(e1) / a.reps * 1e3
        else:
            e0.record()
            for i in range(a.reps):
                run(i)
            e1.record()
            torch.cuda.synchronize()
            us = e0.elapsed_time(e1) / a.reps * 1e3
        _lib.set_profiler(None)
        print(f"{tag:22s} {names[-1][3:]:18s} {us:9.1f} {flops / us / 1e6:9.1f} {byt / us / 1e3:8.0f}", flush=True)


if __name__ == "__main__":
    main()
