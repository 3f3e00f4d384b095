"""Time hn_attn_fwd / hn_attn_bwd alone on the FCT shapes (B=32): CUDA events, L2 flushed between iterations.
Usage: python scripts/attn_bench.py [--kb 64|128]"""
import argparse
import math
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from hybridnet import _lib  # noqa: E402
from hybridnet._lib import call  # noqa: E402

MUFU = 4.647e12


TC = {"on": False, "poly": 0, "bwd": False}


def run(B, N, heads, D, iters=5):
    C = heads * D
    dev = torch.device("cuda")
    q, k, v, do = [(torch.randn(B, N, C, device=dev) * 0.7).to(torch.bfloat16) for _ in range(4)]
    o = torch.empty_like(q)
    lse = torch.empty(B, heads, N, dtype=torch.float32, device=dev)
    dq, dk, dv = torch.empty_like(q), torch.empty_like(q), torch.empty_like(q)
    nws = _lib.load().hn_attn_bwd_workspace(B, N, heads, D)
    ws = torch.empty(nws, dtype=torch.float32, device=dev)
    flush = torch.empty(160 << 20, dtype=torch.uint8, device=dev)
    st = torch.cuda.current_stream().cuda_stream
    scale = 1.0 / math.sqrt(D)

    def fwd():
        if TC["on"] and _lib.load().hn_attn_fwd_tc_supported(N, D, C, C, C, C):
            call("hn_attn_fwd_tc", q.data_ptr(), k.data_ptr(), v.data_ptr(), o.data_ptr(), lse.data_ptr(), B, N, heads, D, C, C, C, C,
                 scale, TC["poly"], st)
            return
        call("hn_attn_fwd", q.data_ptr(), k.data_ptr(), v.data_ptr(), o.data_ptr(), lse.data_ptr(), B, N, heads, D, C, C, C, C, scale, st)

    nws_tc = _lib.load().hn_attn_bwd_tc_workspace(B, N, heads, D)
    ws_tc = torch.empty(nws_tc, dtype=torch.float32, device=dev)

    def bwd():
        if TC["bwd"] and _lib.load().hn_attn_bwd_tc_supported(N, D, C, C, C, C, C, C, C, C):
            call("hn_attn_bwd_tc", q.data_ptr(), k.data_ptr(), v.data_ptr(), o.data_ptr(), do.data_ptr(), lse.data_ptr(), ws_tc.data_ptr(), nws_tc,
                 dq.data_ptr(), dk.data_ptr(), dv.data_ptr(), B, N, heads, D, C, C, C, C, C, C, C, C, scale, TC["poly"], st)
            return
        call("hn_attn_bwd", q.data_ptr(), k.data_ptr(), v.data_ptr(), o.data_ptr(), do.data_ptr(), lse.data_ptr(), ws.data_ptr(), nws,
             dq.data_ptr(), dk.data_ptr(), dv.data_ptr(), B, N, heads, D, C, C, C, C, C, C, C, C, scale, st)
    res = {}
    for name, fn in (("fwd", fwd), ("bwd", bwd)):
        fn()
        torch.cuda.synchronize()
        ms = []
        for _ in range(iters):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            fn()
            e1.record()
            torch.cuda.synchronize()
            ms.append(e0.elapsed_time(e1))
        ms.sort()
        res[name] = ms[len(ms) // 2]
    elems = B * heads * N * N
    print(f"B={B} N={N} D={D}: fwd {res['fwd']:.3f} ms ({elems / res['fwd'] / 1e-3 / MUFU:.3f} of MUFU)  "
          f"bwd {res['bwd']:.3f} ms ({elems * (1 if D <= 8 else 2) / res['bwd'] / 1e-3 / MUFU:.3f} of MUFU)", flush=True)
    return res


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--kb", type=int, default=0)
    ap.add_argument("--poly", type=int, default=-1)
    ap.add_argument("--tc", type=int, default=-1, help="tcgen05 forward with this polynomial share (0, 2, 3, 4)")
    ap.add_argument("--tcbwd", action="store_true", help="tcgen05 backward as well (with --tc)")
    ap.add_argument("--lib", default="", help="a variant library built by scripts/attn_variants.sh")
    ap.add_argument("--big-only", action="store_true", help="only the two exponential-bound shapes")
    a = ap.parse_args()
    if a.lib:
        _lib.LIB_PATH = os.path.abspath(a.lib)
        print("library:", a.lib)
    if a.tc >= 0:
        TC["on"], TC["poly"], TC["bwd"] = True, a.tc, a.tcbwd
        print("tcgen05 forward, polynomial share: every", a.tc)
    if a.poly >= 0:
        call("hn_attn_tune_fwd", a.poly)
        print("forward polynomial share: every", a.poly, "th tile")
    if a.kb:
        call("hn_attn_tune", a.kb)
        print("fused backward keys per CTA:", a.kb)
    tot = {"fwd": 0.0, "bwd": 0.0}
    shapes = ((16384, 4, 2), (4096, 8, 2), (1024, 16, 2), (256, 32, 2), (64, 64, 1))
    for N, D, mult in (shapes[:2] if a.big_only else shapes):
        r = run(32, N, 2, D)
        for kk in tot:
            tot[kk] += mult * r[kk]
    print(f"FCT B=32 per step (9 attention layers): fwd {tot['fwd']:.2f} ms  bwd {tot['bwd']:.2f} ms  total {tot['fwd'] + tot['bwd']:.2f} ms")
