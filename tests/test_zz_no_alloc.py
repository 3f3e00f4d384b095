"""hn_attn_bwd_tc_static: the tcgen05 attention backward with its scratch in a file-scope device array (no allocation
inside the call).  Collected last on purpose (the file name sorts after every other test file).
Checked against the existing entry point (caller workspace) on two calls in a row with different inputs --
the second call finds the first call's panels and dQ accumulator in the shared array."""
import math

import pytest
import torch

from tests.util import rel_l2

pytestmark = pytest.mark.gpu
BWD_TOL = 1.5e-2          # the attention backward tolerance of tests/test_ops_gpu.py
MAX_FLOATS = 2 * 1024 * 1024          # HN_ATTN_BWD_TC_STATIC_MAX_FLOATS (include/hybridnet.h)


def _run_pair(B, N, D, seed):
    from hybridnet import _lib
    from hybridnet._lib import call
    heads = 2
    C = heads * D
    dev = torch.device("cuda")
    g = torch.Generator().manual_seed(seed)
    q, k, v, do = [(torch.randn(B, N, C, generator=g) * 0.7).to(torch.bfloat16).to(dev) for _ in range(4)]
    o = torch.empty_like(q)
    lse = torch.empty(B, heads, N, dtype=torch.float32, device=dev)
    scale = 1.0 / math.sqrt(D)
    st = torch.cuda.current_stream().cuda_stream
    lib = _lib.load()
    assert lib.hn_attn_fwd_tc_supported(N, D, C, C, C, C) and lib.hn_attn_bwd_tc_supported(N, D, C, C, C, C, C, C, C, C)
    call("hn_attn_fwd_tc", q.data_ptr(), k.data_ptr(), v.data_ptr(), o.data_ptr(), lse.data_ptr(), B, N, heads, D, C, C, C, C,
         scale, 3, st)
    nws = lib.hn_attn_bwd_tc_workspace(B, N, heads, D)
    assert nws <= MAX_FLOATS
    ws = torch.empty(nws, dtype=torch.float32, device=dev)
    want = [torch.empty_like(q) for _ in range(3)]
    got = [torch.full_like(q, float("nan")) for _ in range(3)]          # outputs allocated BEFORE the call under test
    call("hn_attn_bwd_tc", q.data_ptr(), k.data_ptr(), v.data_ptr(), o.data_ptr(), do.data_ptr(), lse.data_ptr(), ws.data_ptr(), nws,
         want[0].data_ptr(), want[1].data_ptr(), want[2].data_ptr(), B, N, heads, D, C, C, C, C, C, C, C, C, scale, 3, st)
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated()
    call("hn_attn_bwd_tc_static", q.data_ptr(), k.data_ptr(), v.data_ptr(), o.data_ptr(), do.data_ptr(), lse.data_ptr(),
         got[0].data_ptr(), got[1].data_ptr(), got[2].data_ptr(), B, N, heads, D, C, C, C, C, C, C, C, C, scale, 3, st)
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated() == before
    for name, a, b in zip(("dq", "dk", "dv"), got, want):
        assert torch.isfinite(a.float()).all(), name
        assert float(b.float().norm()) > 0, name
        assert rel_l2(a.float(), b.float()) < BWD_TOL, (name, rel_l2(a.float(), b.float()))


def test_static_scratch_matches_caller_workspace_two_calls_in_a_row():
    _run_pair(2, 1024, 8, seed=11)          # 2 * 1024 * 32 floats of the array
    _run_pair(3, 512, 4, seed=12)           # a different shape and different data over the first call's leftovers
    _run_pair(2, 1024, 8, seed=13)          # and the first shape again


def test_static_scratch_rejects_sizes_beyond_the_array():
    from hybridnet._lib import call
    B, N, D, heads = 8, 16384, 8, 2          # 8 * 16384 * 32 = 4 Mi floats > 2 Mi
    C = heads * D
    t = torch.zeros(16, dtype=torch.bfloat16, device="cuda")          # never touched: the call must fail before any launch
    f = torch.zeros(16, dtype=torch.float32, device="cuda")
    with pytest.raises(RuntimeError, match="built-in array holds"):
        call("hn_attn_bwd_tc_static", t.data_ptr(), t.data_ptr(), t.data_ptr(), t.data_ptr(), t.data_ptr(), f.data_ptr(),
             t.data_ptr(), t.data_ptr(), t.data_ptr(), B, N, heads, D, C, C, C, C, C, C, C, C, 1.0 / math.sqrt(D), 3,
             torch.cuda.current_stream().cuda_stream)
