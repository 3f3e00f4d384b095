"""One small launch of each attention kernel for `ncu --set full` (keeps the replay count low).
Usage: python scripts/ncu_attn.py [D] [tc] [B] [N]   (4 tc 32 16384 and 8 tc 32 4096 are the bench shapes of FCT blocks 1 / 9 and 2 / 8)"""
import math
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from hybridnet import _lib  # noqa: E402
from hybridnet._lib import call  # noqa: E402

B, N, heads, D = (int(sys.argv[3]) if len(sys.argv) > 3 else 4), (int(sys.argv[4]) if len(sys.argv) > 4 else 16384), 2, int(sys.argv[1]) if len(sys.argv) > 1 else 4
C = heads * D
dev = torch.device("cuda")
q, k, v, do = [(torch.randn(B, N, C, device=dev) * 0.7).to(torch.bfloat16) for _ in range(4)]
o = torch.empty_like(q)
lse = torch.empty(B, heads, N, dtype=torch.float32, device=dev)
dq, dk, dv = torch.empty_like(q), torch.empty_like(q), torch.empty_like(q)
nws = _lib.load().hn_attn_bwd_workspace(B, N, heads, D)
ws = torch.empty(nws, dtype=torch.float32, device=dev)
st = torch.cuda.current_stream().cuda_stream
scale = 1.0 / math.sqrt(D)
TC = len(sys.argv) > 2 and sys.argv[2] == "tc"
nws_tc = _lib.load().hn_attn_bwd_tc_workspace(B, N, heads, D)
ws_tc = torch.empty(nws_tc, dtype=torch.float32, device=dev)
for _ in range(2):
    if TC:
        call("hn_attn_fwd_tc", q.data_ptr(), k.data_ptr(), v.data_ptr(), o.data_ptr(), lse.data_ptr(), B, N, heads, D, C, C, C, C, scale, 3, st)
        call("hn_attn_bwd_tc", q.data_ptr(), k.data_ptr(), v.data_ptr(), o.data_ptr(), do.data_ptr(), lse.data_ptr(), ws_tc.data_ptr(), nws_tc,
             dq.data_ptr(), dk.data_ptr(), dv.data_ptr(), B, N, heads, D, C, C, C, C, C, C, C, C, scale, 3, st)
        continue
    call("hn_attn_fwd", q.data_ptr(), k.data_ptr(), v.data_ptr(), o.data_ptr(), lse.data_ptr(), B, N, heads, D, C, C, C, C, scale, st)
    call("hn_attn_bwd", q.data_ptr(), k.data_ptr(), v.data_ptr(), o.data_ptr(), do.data_ptr(), lse.data_ptr(), ws.data_ptr(), nws,
         dq.data_ptr(), dk.data_ptr(), dv.data_ptr(), B, N, heads, D, C, C, C, C, C, C, C, C, scale, st)
torch.cuda.synchronize()
print("ok")
