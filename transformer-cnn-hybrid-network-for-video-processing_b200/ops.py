"""Autograd-visible operators of the hot path: every forward/backward below is a
call into libhybridnet_sm100.so (no ATen compute kernels; torch only allocates
device memory and tracks the autograd graph).

Activations are NHWC bf16 tensors ``[B,H,W,Cp]`` with unit channel stride, where
Cp is the channel count rounded up to 8 (only the RGB input and the 1/3-channel
heads are actually padded).  A channel slice ``t[..., a:b]`` of such a tensor is
itself a valid operand (pixel stride = ld), which is how concat costs nothing.
"""
import contextlib
import math
import os
import weakref

import torch

from . import _lib
from ._lib import ACT_GELU, ACT_NONE, ACT_RELU, ACT_SIGMOID, HnConv, call

ACTS = {"none": ACT_NONE, None: ACT_NONE, "relu": ACT_RELU, "gelu": ACT_GELU, "sigmoid": ACT_SIGMOID}


def pad8(c):
    return (c + 7) // 8 * 8


def _stream():
    return torch.cuda.current_stream().cuda_stream


def _nhwc_ok(t):
    if t.dtype != torch.bfloat16 or t.dim() != 4 or not t.is_cuda:
        return False
    B, H, W, C = t.shape
    ld = t.stride(2) if W > 1 else None
    if t.stride(3) != 1 and C > 1:
        return False
    if ld is None:
        ld = t.stride(1) if H > 1 else (t.stride(0) if B > 1 else C)
        if ld % 8 or ld < C:
            return False
        return (H == 1 or t.stride(1) == W * ld) and (B == 1 or t.stride(0) == H * W * ld)
    if ld % 8 or ld < C:
        return False
    if H > 1 and t.stride(1) != W * ld:
        return False
    if B > 1 and t.stride(0) != H * W * ld:
        return False
    return t.data_ptr() % 16 == 0


def nhwc(t):
    """Return (tensor, ld) with the tensor in a layout the kernels accept."""
    if not _nhwc_ok(t):
        t = t.contiguous()
        if not _nhwc_ok(t):
            raise RuntimeError(f"hybridnet: expected NHWC bf16 CUDA tensor with C%8==0, got {t.dtype} {tuple(t.shape)}")
    B, H, W, C = t.shape
    if W > 1:
        ld = t.stride(2)
    elif H > 1:
        ld = t.stride(1)
    elif B > 1:
        ld = t.stride(0)
    else:
        ld = C
    return t, ld


def _new(B, H, W, C, like, zero=False):
    f = torch.zeros if zero else torch.empty
    return f((B, H, W, C), dtype=torch.bfloat16, device=like.device)


# Zero-initialised fp32 scratch (BatchNorm sums, bias/weight-gradient accumulators: everything the kernels
# add into with atomics).  Eagerly each is its own torch fill launch -- ~250 two-microsecond launches per
# Autoencoder32K step.  A training loop that calls ``begin_step`` once per step gets them as slices of ONE
# pool cleared by ONE memset instead.  Contract: tensors handed out from the pool are dead by the next
# ``begin_step`` (gradients are accumulated into ``p.grad`` / the GradArena before the step ends); without
# ``begin_step`` nothing changes (plain torch.zeros).
_zero_pool = {"buf": None, "off": 0, "live": False}


def begin_step(device, capacity=None):
    """Clear the zero pool (one memset).  capacity (floats) defaults to the previous size; the first call
    must give it (e.g. total parameter count + 64k)."""
    zp = _zero_pool
    device = torch.device(device)
    if zp["buf"] is None or zp["buf"].device != device or (capacity and zp["buf"].numel() < capacity):
        if capacity is None:
            raise RuntimeError("hybridnet.ops.begin_step: first call needs a capacity")
        zp["buf"] = torch.empty(int(capacity), dtype=torch.float32, device=device)
    zp["buf"].zero_()
    zp["off"], zp["live"] = 0, True
    join_deferred()          # a backward pass that raised half-way never ran its end-of-pass callback


def end_zero_pool():
    """Back to one torch.zeros per request (e.g. before code that keeps gradients across steps)."""
    _zero_pool["live"] = False


def _f32(n, like, zero=True):
    if zero:
        zp = _zero_pool
        shape = (n,) if isinstance(n, int) else tuple(n)
        cnt = 1
        for d in shape:
            cnt *= d
        if zp["live"] and zp["buf"].device == like.device:
            end = zp["off"] + (cnt + 3) // 4 * 4            # 16-byte aligned slices (vector reductions)
            if end <= zp["buf"].numel():
                t = zp["buf"][zp["off"]:zp["off"] + cnt].view(shape)
                zp["off"] = end
                return t
        return torch.zeros(shape, dtype=torch.float32, device=like.device)
    return torch.empty(n, dtype=torch.float32, device=like.device)


def _ptr(t):
    return 0 if t is None else t.data_ptr()


# ----------------------------------------------------------------------------
# layout at the module boundary
# ----------------------------------------------------------------------------
class _ToNHWC(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        x = x.contiguous().float()
        B, C, H, W = x.shape
        ctx.C = C
        y = _new(B, H, W, pad8(C), x)
        call("hn_nchw_f32_to_nhwc_bf16", x.data_ptr(), y.data_ptr(), B, C, H, W, pad8(C), pad8(C), _stream())
        return y

    @staticmethod
    def backward(ctx, dy):
        dy, ld = nhwc(dy)
        B, H, W, _ = dy.shape
        dx = torch.empty((B, ctx.C, H, W), dtype=torch.float32, device=dy.device)
        call("hn_nhwc_bf16_to_nchw_f32", dy.data_ptr(), dx.data_ptr(), B, ctx.C, H, W, ld, _stream())
        return dx


class _ToNCHW(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, C):
        x, ld = nhwc(x)
        B, H, W, Cp = x.shape
        ctx.Cp = Cp
        y = torch.empty((B, C, H, W), dtype=torch.float32, device=x.device)
        call("hn_nhwc_bf16_to_nchw_f32", x.data_ptr(), y.data_ptr(), B, C, H, W, ld, _stream())
        return y

    @staticmethod
    def backward(ctx, dy):
        dy = dy.contiguous().float()
        B, C, H, W = dy.shape
        dx = _new(B, H, W, ctx.Cp, dy)
        call("hn_nchw_f32_to_nhwc_bf16", dy.data_ptr(), dx.data_ptr(), B, C, H, W, ctx.Cp, ctx.Cp, _stream())
        return dx, None


class PackedFrames:
    """A batch already in the first conv's operand layout (NHWC bf16, channels padded to 8) -- what
    ``hybridnet.data`` produces on the device; the model forwards accept it in place of the NCHW fp32 tensor."""
    __slots__ = ("nhwc", "channels")

    def __init__(self, nhwc_tensor, channels=3):
        self.nhwc, self.channels = nhwc_tensor, channels

    @property
    def shape(self):
        B, H, W, _ = self.nhwc.shape
        return torch.Size((B, self.channels, H, W))


def to_nhwc(x):
    """[B,C,H,W] fp32 -> [B,H,W,pad8(C)] bf16 (x.to(device) boundary, FCT.py:326)."""
    if isinstance(x, PackedFrames):
        return x.nhwc
    return _ToNHWC.apply(x)


def to_nchw(x, C):
    return _ToNCHW.apply(x, C)


class _HeadFn(torch.autograd.Function):
    """NCHW fp32 output = act(z) from the NHWC bf16 pre-activation: final activation + layout boundary in one pass;
    backward takes act' from the fp32 output (exact sigma' even where the mask saturates)."""

    @staticmethod
    def forward(ctx, z, C, act):
        z, ld = nhwc(z)
        B, H, W, Cp = z.shape
        out = torch.empty((B, C, H, W), dtype=torch.float32, device=z.device)
        call("hn_head_fwd", z.data_ptr(), out.data_ptr(), B, C, H, W, ld, act, _stream())
        ctx.act, ctx.Cp = act, Cp
        ctx.save_for_backward(out)
        return out

    @staticmethod
    def backward(ctx, dout):
        (out,) = ctx.saved_tensors
        dout = dout.contiguous().float()
        B, C, H, W = out.shape
        dz = _new(B, H, W, ctx.Cp, out)
        call("hn_head_bwd", dout.data_ptr(), out.data_ptr(), dz.data_ptr(), B, C, H, W, ctx.Cp, ctx.act, _stream())
        return dz, None, None


def head_nchw(z, C, act="sigmoid"):
    """act(z) as the module's NCHW fp32 output (FCT.py:205, UNet.py:51, pyc:AE_32K L168)."""
    return _HeadFn.apply(z, C, ACTS[act])


class _AvgPool2Fn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        x, ld = nhwc(x)
        B, H, W, C = x.shape
        y = _new(B, H // 2, W // 2, C, x)
        call("hn_avgpool2_fwd", x.data_ptr(), y.data_ptr(), B, H, W, C, ld, C, _stream())
        return y

    @staticmethod
    def backward(ctx, dy):
        dy, ld = nhwc(dy)
        B, H, W, C = dy.shape
        dx = _new(B, 2 * H, 2 * W, C, dy)
        call("hn_avgpool2_bwd", dy.data_ptr(), dx.data_ptr(), B, H, W, C, ld, C, _stream())
        return dx


def avgpool2(x):
    """AvgPool2d(2,2) of the input pyramid (FCT.py:238-240).  Differentiable like nn.AvgPool2d; the backward
    pass is only recorded when the frames require grad (they do not in the reference's trainers)."""
    return _AvgPool2Fn.apply(x)


# ----------------------------------------------------------------------------
# packed-weight cache (derived bf16 copies; the fp32 Parameter stays the truth)
# ----------------------------------------------------------------------------
USE_SC = os.environ.get("HN_DISABLE_SC", "0") != "1"     # small-channel direct conv (halo in smem)
USE_TC = os.environ.get("HN_DISABLE_TC", "0") != "1"     # tcgen05 path for the geometries it supports
_pack_cache = {}          # id(owner tensor) -> entry; dropped by a weakref callback


def _packed(w, kind, Cpad):
    """kind: 'fwd'  conv OIHW as [Cout][(kh,kw,ci)]
             'dgrad' conv OIHW as [Cin][(kh,kw,co)]          (input gradient)
             'tfwd' convT IOHW(k2) as [(tap,co)][ci]          (1x1 GEMM + pixel shuffle)
             'tdgrad' convT IOHW as [Cin][(kh,kw,co)]         (k2 s2 conv of dy)"""
    owner = w._base if w._base is not None else w
    key = (kind, Cpad, w.storage_offset(), tuple(w.shape))
    oid = id(owner)
    ent = _pack_cache.get(oid)
    ver = (owner._version, owner.data_ptr())
    if ent is not None and ent["ref"]() is owner and ent.get("ver") != ver and ent["ver"][1] == ver[1]:
        # the weights changed in place (optimizer step, load_state_dict): refresh the EXISTING buffers.  Their
        # device addresses may be baked into a captured CUDA graph and into the hn_pack_weight_multi job table,
        # so the allocations are never replaced while the parameter's storage stays where it is.
        for job in ent.get("jobs", {}).values():
            _run_pack_job(job)
        ent["ver"] = ver
    elif ent is None or ent["ref"]() is not owner or ent.get("ver") != ver:
        ent = {"ver": ver, "ref": weakref.ref(owner, lambda _r, oid=oid: _pack_cache.pop(oid, None))}
        _pack_cache[oid] = ent
    if key in ent:
        return ent[key]
    src = w.detach()
    assert src.dtype == torch.float32 and src.is_contiguous()
    d0, d1, KH, KW = src.shape
    if kind == "s2dgrad":       # stride-2 input gradient as a stride-1 GEMM + pixel shuffle (hn_pack_weight_s2dgrad)
        lib = _lib.load()
        win = 2 if KH == 3 else 1
        Kpad = lib.hn_conv_kpad(win, win, d0)
        buf = torch.empty((lib.hn_conv_npad(4 * d1), Kpad), dtype=torch.bfloat16, device=w.device)
        call("hn_pack_weight_s2dgrad", src.data_ptr(), buf.data_ptr(), d0, d1, KH, _stream())
        ent[key] = (buf, Kpad)
        ent.setdefault("jobs", {})[key] = [src.data_ptr(), buf.data_ptr(), d0, buf.shape[0], d1, 0, 0, 0, Kpad, 1, 0, 0, 0, 0, 0, KH]
        return ent[key]
    assert kind != "dgradf" or src.is_contiguous()
    big = 1 << 40
    if kind == "fwd":
        N, Creal, kh, kw = d0, d1, KH, KW
        args = (big, 0, d1 * KH * KW, KH * KW, KW, 1)
    elif kind == "dgrad":
        N, Creal, kh, kw = d1, d0, KH, KW
        args = (big, 0, KH * KW, d1 * KH * KW, KW, 1)
    elif kind == "dgradf":      # stride-1 input gradient as a forward conv: taps flipped
        N, Creal, kh, kw = d1, d0, KH, KW
        args = (big, 0, KH * KW, d1 * KH * KW, -KW, -1)
        src = src.reshape(-1)[KH * KW - 1:]
    elif kind == "tfwd":        # weight [Cin, Cout, 2, 2]; n = tap*Cout + co ; c = ci
        N, Creal, kh, kw = 4 * d1, d0, 1, 1
        args = (d1, 1, 4, d1 * 4, 0, 0)
    elif kind == "tdgrad":      # n = ci, c = co, taps (kh,kw)
        N, Creal, kh, kw = d0, d1, KH, KW
        args = (big, 0, d1 * KH * KW, KH * KW, KW, 1)
    else:
        raise KeyError(kind)
    lib = _lib.load()
    Kpad = lib.hn_conv_kpad(kh, kw, Cpad)
    Npad = lib.hn_conv_npad(N)
    buf = torch.empty((Npad, Kpad), dtype=torch.bfloat16, device=w.device)
    call("hn_pack_weight", src.data_ptr(), buf.data_ptr(), N, Creal, Cpad, kh, kw, *args, _stream())
    ent[key] = (buf, Kpad)
    ent.setdefault("jobs", {})[key] = [src.data_ptr(), buf.data_ptr(), N, Npad, Creal, Cpad, kh, kw, Kpad, *args, 0]
    return ent[key]


def _run_pack_job(job):
    """Re-run one recorded pack (same source, same destination buffer)."""
    if job[15] and job[9] == 1 and job[5] == 0:        # hn_pack_weight_s2dgrad job: [src, buf, d0, Npad, d1, 0,0,0, Kpad, 1,0,0,0,0,0, KH]
        call("hn_pack_weight_s2dgrad", job[0], job[1], job[2], job[4], job[15], _stream())
    else:                                              # [src, buf, N, Npad, Creal, Cpad, kh, kw, Kpad, a0..a5, 0]
        call("hn_pack_weight", job[0], job[1], job[2], job[4], job[5], job[6], job[7], *job[9:15], _stream())


_repack_table = {"sig": None, "dev": None, "n": 0, "max": 1}


def repack_all():
    """Refresh EVERY packed bf16 weight copy made so far with one launch (``hn_pack_weight_multi``) and mark
    them current -- what a training loop calls after the optimizer step instead of ``invalidate_packed()`` (which
    makes each layer repack itself: one tiny launch per weight).  Packs created later (first use of a layer)
    are picked up by the next call.  Returns the number of packs refreshed."""
    jobs, dev = [], None
    for ent in _pack_cache.values():
        owner = ent["ref"]()
        if owner is None or "jobs" not in ent:
            continue
        if ent["ver"][1] != owner.data_ptr():          # storage moved (.to()): let the lazy path rebuild it
            continue
        ent["ver"] = (owner._version, owner.data_ptr())
        dev = owner.device
        jobs.extend(ent["jobs"].values())
    if not jobs:
        return 0
    sig = (len(jobs), hash(tuple(j[0] for j in jobs)), hash(tuple(j[1] for j in jobs)))
    if _repack_table["sig"] != sig:
        if torch.cuda.is_current_stream_capturing():
            raise RuntimeError("hybridnet.ops.repack_all: the set of packed weights changed inside a graph capture; "
                               "run one eager step first")
        _repack_table["dev"] = torch.tensor(jobs, dtype=torch.int64, device=dev)
        _repack_table["sig"], _repack_table["n"] = sig, len(jobs)
        _repack_table["max"] = max(j[3] * j[8] for j in jobs)
    call("hn_pack_weight_multi", _repack_table["dev"].data_ptr(), _repack_table["n"], _repack_table["max"], _stream())
    return len(jobs)


def invalidate_packed():
    """Mark every derived bf16 weight copy stale (call after an optimizer step that bypasses ``_version``, or to
    make a benchmark pay for the repack each step like real training does).  The buffers are kept and refilled in
    place on next use: a captured CUDA graph may hold their addresses."""
    for ent in _pack_cache.values():
        ent["ver"] = (None, ent["ver"][1])


def clear_packed():
    """Drop every derived bf16 weight copy and its buffers.  NOT safe while a captured CUDA graph or the
    hn_pack_weight_multi job table still refers to them: for tests and for tearing a model down."""
    _pack_cache.clear()
    _repack_table.update(sig=None, dev=None, n=0, max=1)


def _desc(**kw):
    d = HnConv()
    for k, v in kw.items():
        setattr(d, k, int(v))
    return d


def _conv_raw(x1, x2, wp, Kpad, bias, res, *, OH, OW, Cout, KH, KW, stride, pad, dil, mode=0, ups=0,
              act=ACT_NONE, pshuf=0, stats=None, out=None, H=None, W=None, tc_only=False):
    """One hn_conv2d_fwd launch.  x1/x2 already validated NHWC.  tc_only: the caller's formulation exists
    only on the tcgen05 path."""
    x1, ld1 = nhwc(x1)
    B = x1.shape[0]
    H = H if H is not None else x1.shape[1] << ups
    W = W if W is not None else x1.shape[2] << ups
    C1 = x1.shape[3]
    C2, ld2 = 0, 0
    if x2 is not None:
        x2, ld2 = nhwc(x2)
        C2 = x2.shape[3]
    if out is None:
        oc = pad8(pshuf if pshuf else Cout)
        oh, ow = (2 * OH, 2 * OW) if pshuf else (OH, OW)
        out = _new(B, oh, ow, oc, x1, zero=(oc != (pshuf if pshuf else Cout)))
    out_t, ldy = nhwc(out)
    assert out_t.data_ptr() == out.data_ptr()
    ldr = 0
    if res is not None:
        res, ldr = nhwc(res)
    d = _desc(B=B, H=H, W=W, C1=C1, C2=C2, ld1=ld1, ld2=ld2, OH=OH, OW=OW, Cout=Cout, ldy=ldy, KH=KH, KW=KW,
              stride=stride, pad=pad, dil=dil, mode=mode, ups=ups, act=act, pshuf=pshuf, ldr=ldr, Kpad=Kpad)
    s0 = s1 = None
    if stats is not None:
        s0, s1 = stats
    if tc_only:
        call("hn_conv_tc_fwd", d, x1.data_ptr(), _ptr(x2), wp.data_ptr(), _ptr(bias), out.data_ptr(), _ptr(s0),
             _ptr(s1), _stream())
        return out
    if res is None and USE_TC and _lib.load().hn_conv_tc_halo_supported(d):
        call("hn_conv_tc_fwd", d, x1.data_ptr(), _ptr(x2), wp.data_ptr(), _ptr(bias), out.data_ptr(), _ptr(s0),
             _ptr(s1), _stream())
        return out
    if USE_SC and _lib.load().hn_conv_sc_supported(d):
        if res is None:
            call("hn_conv_sc_fwd", d, x1.data_ptr(), _ptr(x2), wp.data_ptr(), _ptr(bias), out.data_ptr(), _ptr(s0),
                 _ptr(s1), _stream())
        else:
            call("hn_conv_sc_fwd_res", d, x1.data_ptr(), _ptr(x2), wp.data_ptr(), _ptr(bias), res.data_ptr(),
                 out.data_ptr(), _ptr(s0), _ptr(s1), _stream())
        return out
    if res is None and USE_TC and _lib.load().hn_conv_tc_supported(d):
        call("hn_conv_tc_fwd", d, x1.data_ptr(), _ptr(x2), wp.data_ptr(), _ptr(bias), out.data_ptr(), _ptr(s0),
             _ptr(s1), _stream())
        return out
    call("hn_conv2d_fwd", d, x1.data_ptr(), _ptr(x2), wp.data_ptr(), _ptr(bias), _ptr(res), out.data_ptr(),
         _ptr(s0), _ptr(s1), _stream())
    return out


def _wgrad_raw(x1, x2, dz, dw, *, Creal, OH, OW, Cout, KH, KW, stride, pad, dil, ups, strides, H=None, W=None):
    x1, ld1 = nhwc(x1)
    B = x1.shape[0]
    H = H if H is not None else x1.shape[1] << ups
    W = W if W is not None else x1.shape[2] << ups
    C2, ld2 = 0, 0
    if x2 is not None:
        x2, ld2 = nhwc(x2)
        C2 = x2.shape[3]
    dz, lddz = nhwc(dz)
    lib = _lib.load()
    d = _desc(B=B, H=H, W=W, C1=x1.shape[3], C2=C2, ld1=ld1, ld2=ld2, OH=OH, OW=OW, Cout=Cout, ldy=lddz, KH=KH,
              KW=KW, stride=stride, pad=pad, dil=dil, mode=0, ups=ups, act=0, pshuf=0, ldr=0,
              Kpad=lib.hn_conv_kpad(KH, KW, x1.shape[3] + C2))
    if USE_SC and lib.hn_conv_wgrad_sc_supported(d):
        name = "hn_conv_wgrad_sc"
    elif USE_TC and lib.hn_conv_wgrad_tc_supported(d):
        name = "hn_conv_wgrad_tc"
    else:
        name = "hn_conv2d_wgrad"
    if name == "hn_conv_wgrad_tc":
        nws = lib.hn_conv_wgrad_tc_workspace(d, strides[1])       # partial sums of the split-K items (3x3 weights)
        ws = torch.empty(nws, dtype=torch.float32, device=dz.device) if nws > 0 else None
        call(name, d, x1.data_ptr(), _ptr(x2), dz.data_ptr(), dw.data_ptr(), Creal, *strides, _ptr(ws), nws, _stream())
        return
    call(name, d, x1.data_ptr(), _ptr(x2), dz.data_ptr(), dw.data_ptr(), Creal, *strides, _stream())


# ----------------------------------------------------------------------------
# convolution (+bias +residual +activation), optionally followed by BatchNorm
# ----------------------------------------------------------------------------
def _act_bwd(dy, yz, act, want_dbias, bias=None):
    """dz = dy * act'(.) ; dbias = column sums of dz (fp32[Cp]).  With `bias` (the parameter) and a gradient sink whose
    slice has exactly the padded length, the sums are reduced straight into the arena: the function then returns
    dbias = None and has already told the sink (no zero-pool slice, no AccumulateGrad add launch)."""
    dy, lddy = nhwc(dy)
    B, H, W, C = dy.shape
    P = B * H * W
    dbias = None
    sunk = None
    if want_dbias:
        sunk = _sink_for(bias) if (bias is not None and bias.numel() == C and bias.dtype == torch.float32) else None
        dbias = sunk if sunk is not None else _f32(C, dy)
    if act == ACT_NONE:
        if want_dbias:
            call("hn_colsum", dy.data_ptr(), P, C, lddy, dbias.data_ptr(), _stream())
        dz = dy
    else:
        yz, ldy = nhwc(yz)
        dz = _new(B, H, W, C, dy)
        call("hn_act_bwd", dy.data_ptr(), yz.data_ptr(), dz.data_ptr(), P, C, lddy, ldy, C, act, _ptr(dbias), _stream())
    if sunk is not None:
        _sink_done(bias.data_ptr())
        dbias = None
    return dz, dbias


class _Geom:
    """Static geometry of one conv layer call (kept on ctx for backward)."""
    __slots__ = ("KH", "KW", "stride", "pad", "dil", "ups", "transposed", "Cin", "Cout", "H", "W", "OH", "OW")


def _geom(x1, x2, weight, stride, pad, dil, ups, transposed):
    g = _Geom()
    g.transposed = transposed
    g.stride, g.pad, g.dil, g.ups = stride, pad, dil, ups
    g.KH, g.KW = weight.shape[2], weight.shape[3]
    g.H, g.W = x1.shape[1] << ups, x1.shape[2] << ups
    if transposed:
        assert g.KH == 2 and g.KW == 2 and stride == 2 and pad == 0 and x2 is None and not ups
        g.Cin, g.Cout = weight.shape[0], weight.shape[1]
        g.OH, g.OW = g.H, g.W           # GEMM runs on the input grid; output is 2x
    else:
        g.Cout, g.Cin = weight.shape[0], weight.shape[1]
        g.OH = (g.H + 2 * pad - dil * (g.KH - 1) - 1) // stride + 1
        g.OW = (g.W + 2 * pad - dil * (g.KW - 1) - 1) // stride + 1
    return g


def _conv_forward(g, x1, x2, weight, bias, res, act, stats=None):
    Cp = x1.shape[3] + (x2.shape[3] if x2 is not None else 0)
    if g.transposed:
        wp, Kpad = _packed(weight, "tfwd", Cp)
        return _conv_raw(x1, None, wp, Kpad, bias, res, OH=g.OH, OW=g.OW, Cout=g.Cout, KH=1, KW=1, stride=1, pad=0,
                         dil=1, act=act, pshuf=g.Cout, stats=stats)
    wp, Kpad = _packed(weight, "fwd", Cp)
    return _conv_raw(x1, x2, wp, Kpad, bias, res, OH=g.OH, OW=g.OW, Cout=g.Cout, KH=g.KH, KW=g.KW, stride=g.stride,
                     pad=g.pad, dil=g.dil, ups=g.ups, act=act, stats=stats)


# Gradient sink: a training loop that owns the gradient storage (parallel.GradArena) registers
# ``sink(param_data_ptr, numel) -> fp32 accumulation target | None`` and ``done(param_data_ptr)``.  The weight- and
# BatchNorm-parameter gradients are then reduced by the kernels DIRECTLY into that storage (it is zeroed once per
# step) and autograd gets None for them: no per-tensor zero fill, no per-tensor AccumulateGrad add launch.
_grad_sink = {"sink": None, "done": None}


def set_grad_sink(sink, done):
    _grad_sink["sink"], _grad_sink["done"] = sink, done


def _sink_for(t):
    f = _grad_sink["sink"]
    return f(t.data_ptr(), t.numel()) if (f is not None and t is not None and t.is_contiguous()) else None


def _sink_for_ptr(ptr, numel):
    f = _grad_sink["sink"]
    return f(ptr, numel) if (f is not None and ptr) else None


def _sink_done(ptr):
    _grad_sink["done"](ptr)


def _s2dgrad_ok(g, dz):
    """Stride-2 k3p1 / k1p0 conv whose input gradient fits the tcgen05 pixel-shuffle formulation."""
    if not USE_TC or g.stride != 2 or g.dil != 1 or g.ups or g.KH != g.KW or g.KH not in (1, 3) or g.pad != g.KH // 2:
        return False
    if g.H != 2 * g.OH or g.W != 2 * g.OW or g.Cin % 32 or dz.shape[3] != g.Cout or g.Cout % 32:
        return False
    dz, ld = nhwc(dz)
    win = 2 if g.KH == 3 else 1
    d = _desc(B=dz.shape[0], H=g.OH, W=g.OW, C1=g.Cout, C2=0, ld1=ld, ld2=0, OH=g.OH, OW=g.OW, Cout=g.Cin, ldy=g.Cin,
              KH=win, KW=win, stride=1, pad=0, dil=1, mode=0, ups=0, act=0, pshuf=g.Cin, ldr=0, Kpad=win * win * g.Cout)
    return bool(_lib.load().hn_conv_tc_supported(d))


# The weight gradient and the input gradient of one layer are independent given dz.  The small layers (FCT: C = 8..64,
# maps down to 8x8) launch 16..64 CTAs each and leave most of the 148 SMs idle, so the weight gradient goes to a side
# stream and runs beside the input gradient; the streams join again before the function returns (every tensor the
# side stream touches is alive until then, and the gradient sink's `done` fires after the join).  Under CapturedStep the
# fork / join become graph edges.
WGRAD_OVERLAP = True
_side_streams = {}


def _side_stream(dev):
    st = _side_streams.get(dev.index)
    if st is None:
        st = _side_streams[dev.index] = torch.cuda.Stream(device=dev)
    return st


# A weight gradient that a kernel reduces straight into the gradient arena (the sink) has no consumer before the gradient
# exchange / the optimizer, so the side stream need not rejoin after every layer: it runs on beside the backward kernels
# of the layers that follow (most of them launch fewer CTAs than there are SMs) and rejoins ONCE, in a callback the
# autograd engine runs at the end of the backward pass.  The tensors a deferred kernel reads are kept alive until then;
# GradArena makes its all-reduce stream wait for the side stream as well (deferred_side_stream).
WGRAD_DEFER = True
_deferred = {"refs": [], "side": None, "main": None}


def _join_deferred():
    st = _deferred
    if st["side"] is not None:
        st["main"].wait_stream(st["side"])
    st["refs"].clear()
    st["side"] = st["main"] = None


def _defer_join(main, side, tensors):
    st = _deferred
    if st["side"] is None:
        try:
            torch.autograd.Variable._execution_engine.queue_callback(_join_deferred)
        except RuntimeError:          # not inside a backward pass
            return False
        st["side"], st["main"] = side, main
    elif st["side"] is not side or st["main"] != main:
        return False
    st["refs"].append(tensors)
    return True


def join_deferred():
    """Rejoin a side stream left over from a backward pass whose end-of-pass callback did not run (an exception inside
    backward).  Called at the start of every step (begin_step, GradArena.zero_grad); a no-op otherwise."""
    if _deferred["side"] is not None:
        _join_deferred()


def deferred_side_stream():
    """The side stream with weight-gradient kernels still to be joined (None when there is none)."""
    return _deferred["side"]


def _conv_backward(g, x1, x2, weight, dz, need_dx, need_dw, bias=None):
    """dz: gradient at the conv output (NHWC, padded channels zero). Returns dx1, dx2, dW.
    bias: (sink slice, parameter address) of an activation-free conv's bias whose gradient (column sums of dz) is still
    owed: the sums are then taken beside the weight gradient (on the side stream when there is one)."""
    if not (WGRAD_OVERLAP and need_dx and need_dw):
        return _conv_backward_on(g, x1, x2, weight, dz, need_dx, need_dw, None, bias)
    dz, _ = nhwc(dz)
    return _conv_backward_on(g, x1, x2, weight, dz, need_dx, need_dw, _side_stream(dz.device), bias)


def _conv_backward_on(g, x1, x2, weight, dz, need_dx, need_dw, side, bias=None):
    dx1 = dx2 = dw = None
    dz, lddz = nhwc(dz)
    B = x1.shape[0]
    C1 = x1.shape[3]
    C2 = x2.shape[3] if x2 is not None else 0
    sunk = None
    main = torch.cuda.current_stream() if side is not None else None
    if side is not None:
        side.wait_stream(main)
    with (torch.cuda.stream(side) if side is not None else contextlib.nullcontext()):
        if bias is not None:
            call("hn_colsum", dz.data_ptr(), dz.shape[0] * dz.shape[1] * dz.shape[2], dz.shape[3], lddz,
                 bias[0].data_ptr(), _stream())
        if need_dw:
            sunk = _sink_for(weight)
            dw = sunk.view(weight.shape) if sunk is not None else _f32(tuple(weight.shape), weight)
            if g.transposed:
                # dW[ci][co][kh][kw] = sum_p x[p,ci] * dy[2p+(kh,kw), co]: wgrad of the k2s2 conv dy -> x
                _wgrad_raw(dz, None, x1, dw, Creal=g.Cout, OH=g.H, OW=g.W, Cout=g.Cin, KH=2, KW=2, stride=2, pad=0,
                           dil=1, ups=0, strides=(g.Cout * 4, 4, 2, 1))
            else:
                _wgrad_raw(x1, x2, dz, dw, Creal=g.Cin, OH=g.OH, OW=g.OW, Cout=g.Cout, KH=g.KH, KW=g.KW,
                           stride=g.stride, pad=g.pad, dil=g.dil, ups=g.ups,
                           strides=(g.Cin * g.KH * g.KW, g.KH * g.KW, g.KW, 1), H=g.H, W=g.W)
    if need_dx:
        if g.transposed:
            wp, Kpad = _packed(weight, "tdgrad", dz.shape[3])
            dx1 = _conv_raw(dz, None, wp, Kpad, None, None, OH=g.H, OW=g.W, Cout=g.Cin, KH=2, KW=2, stride=2, pad=0,
                            dil=1)
        else:
            if g.stride == 1:      # forward-gather form with flipped taps (tensor-core path eligible)
                wp, Kpad = _packed(weight, "dgradf", dz.shape[3])
                full = _conv_raw(dz, None, wp, Kpad, None, None, OH=g.H, OW=g.W, Cout=g.Cin, KH=g.KH, KW=g.KW,
                                 stride=1, pad=g.dil * (g.KH - 1) - g.pad, dil=g.dil, mode=0, H=g.OH, W=g.OW)
            elif _s2dgrad_ok(g, dz):
                # every dx pixel parity is a small stride-1 conv of dz: one tcgen05 GEMM with a 2x2 (k3) / 1x1 (k1)
                # window, the four parities on N, pixel-shuffle store (no scatter, no zero-fill pass)
                wp, Kpad = _packed(weight, "s2dgrad", dz.shape[3])
                win = 2 if g.KH == 3 else 1
                full = _conv_raw(dz, None, wp, Kpad, None, None, OH=g.OH, OW=g.OW, Cout=g.Cin, KH=win, KW=win,
                                 stride=1, pad=0, dil=1, pshuf=g.Cin, tc_only=True)
            else:
                wp, Kpad = _packed(weight, "dgrad", dz.shape[3])
                full = _conv_raw(dz, None, wp, Kpad, None, None, OH=g.H, OW=g.W, Cout=g.Cin, KH=g.KH, KW=g.KW,
                                 stride=g.stride, pad=g.pad, dil=g.dil, mode=1, H=g.OH, W=g.OW)
            if full.shape[3] != C1 + C2:        # image input: Cin=3 padded to 8
                assert C2 == 0 and full.shape[3] == C1
            if g.ups:
                small = _new(B, g.H // 2, g.W // 2, full.shape[3], full)
                call("hn_upsample2_bwd", full.data_ptr(), small.data_ptr(), B, g.H // 2, g.W // 2, full.shape[3],
                     full.shape[3], full.shape[3], _stream())
                full = small
            dx1 = full[..., :C1] if C2 else full
            dx2 = full[..., C1:] if C2 else None
    if side is not None:
        if not (sunk is not None and WGRAD_DEFER and _defer_join(main, side, (dz, x1, x2))):
            main.wait_stream(side)
    if bias is not None:
        _sink_done(bias[1])
    if sunk is not None:
        _sink_done(weight.data_ptr())
        dw = None
    return dx1, dx2, dw


class _ConvFn(torch.autograd.Function):
    """y = act(conv(x1|x2) + bias [+ res])  -- nn.Conv2d / nn.ConvTranspose2d forward+backward."""

    @staticmethod
    def forward(ctx, x1, x2, weight, bias, res, stride, pad, dil, ups, act, transposed):
        x1, _ = nhwc(x1)
        if x2 is not None:
            x2, _ = nhwc(x2)
        g = _geom(x1, x2, weight, stride, pad, dil, ups, transposed)
        y = _conv_forward(g, x1, x2, weight, bias, res, act)
        ctx.g, ctx.act = g, act
        ctx.has_bias, ctx.has_res = bias is not None, res is not None
        ctx.bias = bias                       # the parameter itself (its address finds the gradient sink in backward)
        ctx.save_for_backward(x1, x2, weight, y if act != ACT_NONE else None)
        return y

    @staticmethod
    def backward(ctx, dy):
        x1, x2, weight, y = ctx.saved_tensors
        g = ctx.g
        need = ctx.needs_input_grad
        want_db = ctx.has_bias and need[3]
        late_bias = None
        if (want_db and ctx.act == ACT_NONE and need[2] and ctx.bias.dtype == torch.float32
                and ctx.bias.numel() == dy.shape[3]):
            # activation-free conv: dz = dy and the bias gradient is one column-sum launch.  With a slice in the gradient
            # sink it goes beside the weight gradient (the side stream under WGRAD_OVERLAP) instead of ahead of the input
            # gradient on this stream
            bs = _sink_for(ctx.bias)
            if bs is not None:
                late_bias, want_db = (bs, ctx.bias.data_ptr()), False
        dz, dbias = _act_bwd(dy, y, ctx.act, want_db, ctx.bias)
        dx1, dx2, dw = _conv_backward(g, x1, x2, weight, dz, need[0] or need[1], need[2], late_bias)
        if dbias is not None:
            dbias = dbias[: g.Cout]
        return (dx1 if need[0] else None, dx2 if need[1] else None, dw, dbias,
                dz if (ctx.has_res and need[4]) else None, None, None, None, None, None, None)


class _ActFn(torch.autograd.Function):
    """Stand-alone activation that keeps its input (GELU needs the pre-activation in backward)."""

    @staticmethod
    def forward(ctx, z, act):
        z, ld = nhwc(z)
        B, H, W, C = z.shape
        y = _new(B, H, W, C, z)
        call("hn_act_fwd", z.data_ptr(), y.data_ptr(), B * H * W, C, ld, C, act, _stream())
        ctx.act = act
        ctx.save_for_backward(z)
        return y

    @staticmethod
    def backward(ctx, dy):
        (z,) = ctx.saved_tensors
        dz, _ = _act_bwd(dy, z, ctx.act, False)
        return dz, None


def conv2d(x, weight, bias=None, *, x2=None, res=None, stride=1, pad=1, dil=1, ups=0, act="none", transposed=False):
    a = ACTS[act]
    if a == ACT_GELU:      # GELU' needs z: keep the pre-activation, apply GELU as its own pass
        z = _ConvFn.apply(x, x2, weight, bias, res, stride, pad, dil, ups, ACT_NONE, transposed)
        return _ActFn.apply(z, ACT_GELU)
    return _ConvFn.apply(x, x2, weight, bias, res, stride, pad, dil, ups, a, transposed)


def _pad_vec(v, n, fill=0.0):
    if v.numel() == n:
        return v.detach().float().contiguous()
    out = torch.full((n,), fill, dtype=torch.float32, device=v.device)
    out[: v.numel()] = v.detach()
    return out


class _BnActFn(torch.autograd.Function):
    """y = act(BatchNorm2d(z) [+ res]) with batch statistics in training mode.
    If stats=(sum,sqsum) is given they came for free from the producing conv's epilogue."""

    @staticmethod
    def forward(ctx, z, gamma, beta, res, running_mean, running_var, training, momentum, eps, act, stats, count_real):
        z, ldz = nhwc(z)
        B, H, W, Cp = z.shape
        P = B * H * W
        C = gamma.numel()
        gam, bet = _pad_vec(gamma, Cp), _pad_vec(beta, Cp)
        mean, rstd = _f32(Cp, z, False), _f32(Cp, z, False)
        if training:
            if stats is None:
                s0, s1 = _f32(Cp, z), _f32(Cp, z)
                call("hn_bn_stats", z.data_ptr(), P, Cp, ldz, s0.data_ptr(), s1.data_ptr(), _stream())
            else:
                s0, s1 = stats
            rm = rv = None
            if running_mean is not None:
                rm, rv = _pad_vec(running_mean, Cp), _pad_vec(running_var, Cp, 1.0)
        else:
            mean = _pad_vec(running_mean, Cp)
            rstd = torch.rsqrt(_pad_vec(running_var, Cp, 1.0) + eps)
        y = _new(B, H, W, Cp, z)
        ldr = 0
        if res is not None:
            res, ldr = nhwc(res)
        if training:
            # statistics -> mean / rstd / running statistics and the normalisation in one launch
            call("hn_bn_apply_train", z.data_ptr(), _ptr(res), y.data_ptr(), s0.data_ptr(), s1.data_ptr(), P, eps, momentum,
                 mean.data_ptr(), rstd.data_ptr(), _ptr(rm), _ptr(rv), _ptr(count_real), gam.data_ptr(), bet.data_ptr(),
                 P, Cp, ldz, ldr, Cp, act, _stream())
            if rm is not None and rm.data_ptr() != running_mean.data_ptr():      # padded copies only (C % 8 != 0)
                running_mean.copy_(rm[:C])
                running_var.copy_(rv[:C])
        else:
            call("hn_bn_apply", z.data_ptr(), _ptr(res), y.data_ptr(), mean.data_ptr(), rstd.data_ptr(), gam.data_ptr(),
                 bet.data_ptr(), P, Cp, ldz, ldr, Cp, act, _stream())
        ctx.act, ctx.training, ctx.C, ctx.has_res = act, training, C, res is not None
        direct = C == Cp and gamma.requires_grad and beta.requires_grad and gamma.dtype == torch.float32
        ctx.gptr, ctx.bptr = (gamma.data_ptr(), beta.data_ptr()) if direct else (0, 0)
        ctx.maskz = act == ACT_RELU and res is None      # ReLU mask re-derived from z in backward
        ctx.save_for_backward(z, y if (act != ACT_NONE and not ctx.maskz) else None, mean, rstd, gam,
                              bet if ctx.maskz else None)
        return y

    @staticmethod
    def backward(ctx, dy):
        z, y, mean, rstd, gam, bet = ctx.saved_tensors
        dy, lddy = nhwc(dy)
        z, ldz = nhwc(z)
        B, H, W, Cp = z.shape
        P = B * H * W
        # per-channel sums of the two reductions ARE dbeta / dgamma: reduce straight into the gradient sink
        k1 = _sink_for_ptr(ctx.bptr, Cp)
        k2 = _sink_for_ptr(ctx.gptr, Cp) if k1 is not None else None
        sunk = k1 is not None and k2 is not None
        s1, s2 = (k1, k2) if sunk else (_f32(Cp, z), _f32(Cp, z))
        if ctx.maskz:
            call("hn_bn_relu_bwd_reduce", dy.data_ptr(), z.data_ptr(), mean.data_ptr(), rstd.data_ptr(),
                 gam.data_ptr(), bet.data_ptr(), P, Cp, lddy, ldz, s1.data_ptr(), s2.data_ptr(), _stream())
            dx = _new(B, H, W, Cp, z)
            call("hn_bn_relu_bwd_apply", dy.data_ptr(), z.data_ptr(), dx.data_ptr(), mean.data_ptr(),
                 rstd.data_ptr(), gam.data_ptr(), bet.data_ptr(), s1.data_ptr(), s2.data_ptr(), P, Cp, lddy, ldz,
                 Cp, 1 if ctx.training else 0, _stream())
            C = ctx.C
            if sunk:
                _sink_done(ctx.gptr)
                _sink_done(ctx.bptr)
                return (dx, None, None, None, None, None, None, None, None, None, None, None)
            return (dx, s2[:C], s1[:C], None, None, None, None, None, None, None, None, None)
        need_dz_buf = ctx.act != ACT_NONE
        dz = _new(B, H, W, Cp, z) if need_dz_buf else dy
        lddz = Cp if need_dz_buf else lddy
        ldy = 0
        if y is not None:
            y, ldy = nhwc(y)
        call("hn_bn_bwd_reduce", dy.data_ptr(), _ptr(y), z.data_ptr(), dz.data_ptr() if need_dz_buf else 0,
             mean.data_ptr(), rstd.data_ptr(), P, Cp, lddy, ldy, ldz, lddz, ctx.act, s1.data_ptr(), s2.data_ptr(),
             _stream())
        dx = _new(B, H, W, Cp, z)
        call("hn_bn_bwd_apply", dz.data_ptr(), z.data_ptr(), dx.data_ptr(), mean.data_ptr(), rstd.data_ptr(),
             gam.data_ptr(), s1.data_ptr(), s2.data_ptr(), P, Cp, lddz, ldz, Cp, 1 if ctx.training else 0, _stream())
        C = ctx.C
        if sunk:
            _sink_done(ctx.gptr)
            _sink_done(ctx.bptr)
            return (dx, None, None, dz if ctx.has_res else None, None, None, None, None, None, None, None, None)
        return (dx, s2[:C], s1[:C], dz if ctx.has_res else None,
                None, None, None, None, None, None, None, None)


def bn_act(z, gamma, beta, running_mean, running_var, training, *, res=None, momentum=0.1, eps=1e-5, act="none",
           stats=None, num_batches_tracked=None):
    """num_batches_tracked: the module's int64 counter; when given (training) the finalize kernel bumps it."""
    return _BnActFn.apply(z, gamma, beta, res, running_mean, running_var, training, momentum, eps, ACTS[act], stats,
                          num_batches_tracked if training else None)


class _ConvStatsFn(torch.autograd.Function):
    """conv (+bias) whose epilogue also emits the per-channel sum / sum-of-squares BatchNorm needs."""

    @staticmethod
    def forward(ctx, x1, x2, weight, bias, stride, pad, dil, ups, transposed):
        x1, _ = nhwc(x1)
        if x2 is not None:
            x2, _ = nhwc(x2)
        g = _geom(x1, x2, weight, stride, pad, dil, ups, transposed)
        Cp = pad8(g.Cout)
        s0, s1 = _f32(Cp, x1), _f32(Cp, x1)
        y = _conv_forward(g, x1, x2, weight, bias, None, ACT_NONE, stats=(s0, s1))
        ctx.g, ctx.has_bias = g, bias is not None
        ctx.save_for_backward(x1, x2, weight)
        ctx.mark_non_differentiable(s0, s1)
        ctx.set_materialize_grads(False)      # or autograd fills two zero vectors per layer for the unused s0/s1 grads
        return y, s0, s1

    @staticmethod
    def backward(ctx, dy, _a, _b):
        x1, x2, weight = ctx.saved_tensors
        g = ctx.g
        need = ctx.needs_input_grad
        # a bias in front of train-mode BatchNorm has an exactly-zero gradient: dy here is the BN input
        # gradient, whose per-channel sum vanishes identically (sum_p [dz - mean(dz) - xhat*mean(dz*xhat)]
        # = 0); the reference's autograd gets fp32 round-off around 0.  No reduction pass is spent on it.
        dz, _ = _act_bwd(dy, None, ACT_NONE, False)
        dbias = _f32(g.Cout, dz) if (ctx.has_bias and need[3]) else None
        dx1, dx2, dw = _conv_backward(g, x1, x2, weight, dz, need[0] or need[1], need[2])
        return (dx1 if need[0] else None, dx2 if need[1] else None, dw, dbias, None, None, None, None, None)


FOLD_EVAL_BN = os.environ.get("HN_DISABLE_BN_FOLD", "0") != "1"
_fold_cache = {}          # id(bn) -> (signature, folded weight, folded bias); dropped with the module


def _folded_eval_bn(weight, bias, bn, transposed):
    ts = (weight, bias, bn.weight, bn.bias, bn.running_mean, bn.running_var)
    sig = tuple((t.data_ptr(), t._version) if t is not None else None for t in ts) + (transposed, bn.eps)
    ent = _fold_cache.get(id(bn))
    if ent is not None and ent[0] == sig and ent[3]() is bn:
        return ent[1], ent[2]
    with torch.no_grad():
        scale = bn.weight.float() * torch.rsqrt(bn.running_var.float() + bn.eps)
        shift = bn.bias.float() - bn.running_mean.float() * scale
        wf = (weight.float() * (scale.view(1, -1, 1, 1) if transposed else scale.view(-1, 1, 1, 1))).contiguous()
        bf = (shift if bias is None else bias.float() * scale + shift).contiguous()
    key = id(bn)
    _fold_cache[key] = (sig, wf, bf, weakref.ref(bn, lambda _r, key=key: _fold_cache.pop(key, None)))
    return wf, bf


def conv_bn_act(x, weight, bias, bn, *, x2=None, res=None, stride=1, pad=1, dil=1, ups=0, act="relu",
                transposed=False):
    """Conv2d/ConvTranspose2d -> BatchNorm2d -> activation (UNet.py:54-66; pyc:AE_32K).  In training
    the batch statistics ride the conv epilogue; ``bn`` is the BatchNorm2d module (params + buffers)."""
    training = bn.training or (bn.running_mean is None)
    if not training and res is None and not torch.is_grad_enabled() and FOLD_EVAL_BN:
        # inference: eval-mode BatchNorm is a per-channel affine map of the conv output -- fold it into the conv's
        # weights and bias (fp32, cached until a parameter / running statistic changes) and let the conv epilogue
        # apply the activation: no BN pass at all (Bottleneck's bn3 + residual keeps the bn_apply path)
        wf, bf = _folded_eval_bn(weight, bias, bn, transposed)
        return conv2d(x, wf, bf, x2=x2, stride=stride, pad=pad, dil=dil, ups=ups, act=act, transposed=transposed)
    if training:
        # hn_bn_finalize moves the running statistics through raw pointers (no ``_version`` bump): whatever was
        # folded for inference from the old statistics is stale from here on
        _fold_cache.pop(id(bn), None)
        z, s0, s1 = _ConvStatsFn.apply(x, x2, weight, bias, stride, pad, dil, ups, transposed)
        stats = (s0, s1)
    else:
        z = conv2d(x, weight, bias, x2=x2, stride=stride, pad=pad, dil=dil, ups=ups, transposed=transposed)
        stats = None
    return bn_act(z, bn.weight, bn.bias, bn.running_mean, bn.running_var, training, res=res, momentum=bn.momentum,
                  eps=bn.eps, act=act, stats=stats, num_batches_tracked=bn.num_batches_tracked)


# ----------------------------------------------------------------------------
# pooling / dropout / pointwise
# ----------------------------------------------------------------------------
class _MaxPool2Fn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        x, ld = nhwc(x)
        B, H, W, C = x.shape
        y = _new(B, H // 2, W // 2, C, x)
        call("hn_maxpool2_fwd", x.data_ptr(), y.data_ptr(), B, H, W, C, ld, C, _stream())
        ctx.save_for_backward(x)
        return y

    @staticmethod
    def backward(ctx, dy):
        (x,) = ctx.saved_tensors
        x, ld = nhwc(x)
        dy, lddy = nhwc(dy)
        B, H, W, C = x.shape
        dx = _new(B, H, W, C, x)
        call("hn_maxpool2_bwd", x.data_ptr(), dy.data_ptr(), dx.data_ptr(), B, H, W, C, ld, lddy, C, _stream())
        return dx


def maxpool2(x):
    """nn.MaxPool2d(2) (FCT.py:147; UNet.py:12)."""
    return _MaxPool2Fn.apply(x)


_dropout_seed = [0x5DEECE66D]
_step_counter = {}        # device -> u64 tensor read by hn_dropout (see bump_step_counter)


def manual_seed(seed):
    """Seed of the stateless dropout hash (per-rank under data parallelism)."""
    _dropout_seed[0] = int(seed) & ((1 << 62) - 1)


def step_counter(device):
    """Device-resident step counter mixed into every dropout seed by the kernel itself."""
    t = _step_counter.get(device)
    if t is None:
        t = _step_counter[device] = torch.zeros(1, dtype=torch.int64, device=device)
    return t


def bump_step_counter(device):
    """One in-place device add: inside a captured step it makes every graph replay draw new dropout
    masks although the host-side seeds were frozen at capture time."""
    step_counter(device).add_(1)


class _DropoutFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, p, channelwise, seed):
        x, ld = nhwc(x)
        B, H, W, C = x.shape
        y = _new(B, H, W, C, x)
        ctr = step_counter(x.device)
        call("hn_dropout", x.data_ptr(), y.data_ptr(), B, H * W, C, ld, C, p, seed, ctr.data_ptr(), int(channelwise), _stream())
        ctx.p, ctx.cw, ctx.seed, ctx.ctr = p, channelwise, seed, ctr
        return y

    @staticmethod
    def backward(ctx, dy):
        dy, ld = nhwc(dy)
        B, H, W, C = dy.shape
        dx = _new(B, H, W, C, dy)
        call("hn_dropout", dy.data_ptr(), dx.data_ptr(), B, H * W, C, ld, C, ctx.p, ctx.seed, ctx.ctr.data_ptr(), int(ctx.cw), _stream())
        return dx, None, None, None


class _MaskMulFn(torch.autograd.Function):
    """Injected dropout mask (parity tests): y = x * mask, mask already scaled by 1/(1-p)."""

    @staticmethod
    def forward(ctx, x, mask):
        ctx.save_for_backward(mask)
        return x * mask

    @staticmethod
    def backward(ctx, dy):
        (mask,) = ctx.saved_tensors
        return dy * mask, None


def dropout(x, p, training, channelwise=False, mask=None):
    """nn.Dropout / nn.Dropout2d.  ``mask`` (NHWC bf16, pre-scaled) overrides the hash for parity tests."""
    if mask is not None:
        return _MaskMulFn.apply(x, mask)
    if not training or p == 0.0:
        return x
    _dropout_seed[0] = (_dropout_seed[0] * 6364136223846793005 + 1442695040888963407) & ((1 << 62) - 1)
    return _DropoutFn.apply(x, float(p), channelwise, _dropout_seed[0])


class _AddFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, a, b):
        a, lda = nhwc(a)
        b, ldb = nhwc(b)
        B, H, W, C = a.shape
        y = _new(B, H, W, C, a)
        call("hn_add", a.data_ptr(), b.data_ptr(), 0, y.data_ptr(), B * H * W, C, lda, ldb, 0, C, _stream())
        return y

    @staticmethod
    def backward(ctx, dy):
        return dy, dy


def add(a, b):
    """torch.add residual (FCT.py:96,101)."""
    return _AddFn.apply(a, b)


class _Gelu3SumFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, z1, z2, z3):
        z1, z2, z3 = z1.contiguous(), z2.contiguous(), z3.contiguous()
        y = torch.empty_like(z1)
        call("hn_gelu3_sum_fwd", z1.data_ptr(), z2.data_ptr(), z3.data_ptr(), y.data_ptr(), z1.numel(), _stream())
        ctx.save_for_backward(z1, z2, z3)
        return y

    @staticmethod
    def backward(ctx, dy):
        z1, z2, z3 = ctx.saved_tensors
        dy = dy.contiguous()
        d1, d2, d3 = torch.empty_like(z1), torch.empty_like(z1), torch.empty_like(z1)
        call("hn_gelu3_sum_bwd", dy.data_ptr(), z1.data_ptr(), z2.data_ptr(), z3.data_ptr(), d1.data_ptr(),
             d2.data_ptr(), d3.data_ptr(), z1.numel(), _stream())
        return d1, d2, d3


def gelu3_sum(z1, z2, z3):
    """gelu(z1)+gelu(z2)+gelu(z3): the three Wide_Focus branches (FCT.py:118-129)."""
    return _Gelu3SumFn.apply(z1, z2, z3)


# ----------------------------------------------------------------------------
# Wide_Focus (FCT.py:107-132): three dilated convs of one input, fused GELU/dropout/sum glue
# ----------------------------------------------------------------------------
class _MultiConvFn(torch.autograd.Function):
    """z_i = conv(x; w_i, dilation d_i) for the three Wide_Focus branches (no bias: it is added by the glue kernel).
    One autograd node, so that the three input gradients are summed by ONE hn_add launch instead of autograd's
    two elementwise fan-in adds."""

    @staticmethod
    def forward(ctx, x, w1, w2, w3, dils):
        x, _ = nhwc(x)
        ws = (w1, w2, w3)
        gs = [_geom(x, None, w, 1, d * (w.shape[2] - 1) // 2, d, 0, False) for w, d in zip(ws, dils)]
        zs = [_conv_forward(g, x, None, w, None, None, ACT_NONE) for g, w in zip(gs, ws)]
        ctx.gs = gs
        ctx.save_for_backward(x, w1, w2, w3)
        return tuple(zs)

    @staticmethod
    def backward(ctx, d1, d2, d3):
        x, w1, w2, w3 = ctx.saved_tensors
        need = ctx.needs_input_grad
        dxs, dws = [], []
        for g, w, d, nw in zip(ctx.gs, (w1, w2, w3), (d1, d2, d3), need[1:4]):
            dx, _, dw = _conv_backward(g, x, None, w, d, need[0], nw)
            dxs.append(dx)
            dws.append(dw)
        dx = None
        if need[0]:
            a, lda = nhwc(dxs[0])
            b, ldb = nhwc(dxs[1])
            c, ldc = nhwc(dxs[2])
            B, H, W, C = a.shape
            dx = _new(B, H, W, C, a)
            call("hn_add", a.data_ptr(), b.data_ptr(), c.data_ptr(), dx.data_ptr(), B * H * W, C, lda, ldb, ldc, C, _stream())
        return dx, dws[0], dws[1], dws[2], None


def _next_seed():
    _dropout_seed[0] = (_dropout_seed[0] * 6364136223846793005 + 1442695040888963407) & ((1 << 62) - 1)
    return _dropout_seed[0]


def _bias_grad_target(b):
    """fp32 [C] accumulation target for a bias gradient: the gradient sink when there is one, else scratch."""
    sunk = _sink_for(b)
    return (sunk, True) if sunk is not None else (_f32(b.numel(), b), False)


class _WfCombineFn(torch.autograd.Function):
    """y = sum_i dropout(gelu(z_i + b_i)): one pass forward, one pass backward (dz_i and the three bias gradients)."""

    @staticmethod
    def forward(ctx, z1, z2, z3, b1, b2, b3, p, seeds):
        z1, ld = nhwc(z1)
        z2, ld2 = nhwc(z2)
        z3, ld3 = nhwc(z3)
        assert ld == ld2 == ld3
        B, H, W, C = z1.shape
        y = _new(B, H, W, C, z1)
        ctr = step_counter(z1.device)
        call("hn_wf_combine_fwd", z1.data_ptr(), z2.data_ptr(), z3.data_ptr(), b1.data_ptr(), b2.data_ptr(), b3.data_ptr(),
             y.data_ptr(), B * H * W, C, ld, C, p, seeds[0], seeds[1], seeds[2], ctr.data_ptr(), _stream())
        ctx.p, ctx.seeds, ctx.ctr = p, seeds, ctr
        ctx.save_for_backward(z1, z2, z3, b1, b2, b3)
        return y

    @staticmethod
    def backward(ctx, dy):
        z1, z2, z3, b1, b2, b3 = ctx.saved_tensors
        dy, lddy = nhwc(dy)
        z1, ld = nhwc(z1)
        B, H, W, C = z1.shape
        d = torch.empty((3, B, H, W, C), dtype=torch.bfloat16, device=z1.device)
        tg = [_bias_grad_target(b) for b in (b1, b2, b3)]
        call("hn_wf_combine_bwd", dy.data_ptr(), z1.data_ptr(), z2.data_ptr(), z3.data_ptr(), b1.data_ptr(), b2.data_ptr(),
             b3.data_ptr(), d[0].data_ptr(), d[1].data_ptr(), d[2].data_ptr(), B * H * W, C, lddy, ld, C, ctx.p,
             ctx.seeds[0], ctx.seeds[1], ctx.seeds[2], ctx.ctr.data_ptr(), tg[0][0].data_ptr(), tg[1][0].data_ptr(),
             tg[2][0].data_ptr(), _stream())
        dbs = []
        for b, (t, sunk) in zip((b1, b2, b3), tg):
            if sunk:
                _sink_done(b.data_ptr())
                dbs.append(None)
            else:
                dbs.append(t)
        return d[0], d[1], d[2], dbs[0], dbs[1], dbs[2], None, None


class _WfOutFn(torch.autograd.Function):
    """out = res + dropout(gelu(z + bias)) (conv4 of Wide_Focus and the Transformer's second residual, FCT.py:101,131)."""

    @staticmethod
    def forward(ctx, z, bias, res, p, seed):
        z, ld = nhwc(z)
        B, H, W, C = z.shape
        ldr = 0
        if res is not None:
            res, ldr = nhwc(res)
        out = _new(B, H, W, C, z)
        ctr = step_counter(z.device)
        call("hn_wf_out_fwd", z.data_ptr(), bias.data_ptr(), _ptr(res), out.data_ptr(), B * H * W, C, ld, ldr, C, p, seed,
             ctr.data_ptr(), _stream())
        ctx.p, ctx.seed, ctx.ctr, ctx.has_res = p, seed, ctr, res is not None
        ctx.save_for_backward(z, bias)
        return out

    @staticmethod
    def backward(ctx, dout):
        z, bias = ctx.saved_tensors
        dout, lddo = nhwc(dout)
        z, ld = nhwc(z)
        B, H, W, C = z.shape
        dz = _new(B, H, W, C, z)
        t, sunk = _bias_grad_target(bias)
        call("hn_wf_out_bwd", dout.data_ptr(), z.data_ptr(), bias.data_ptr(), dz.data_ptr(), B * H * W, C, lddo, ld, C,
             ctx.p, ctx.seed, ctx.ctr.data_ptr(), t.data_ptr(), _stream())
        if sunk:
            _sink_done(bias.data_ptr())
        return dz, (None if sunk else t), (dout if ctx.has_res else None), None, None


def wide_focus(x, convs, res, p, training):
    """Wide_Focus.forward (FCT.py:117-132) + the residual add of Transformer.forward (FCT.py:101).
    convs = (conv1, conv2, conv3, conv4) modules (3x3, dilation 1/2/3/1, bias)."""
    p = float(p) if training else 0.0
    c1, c2, c3, c4 = convs
    if any(c.bias is None for c in convs):
        raise RuntimeError("hybridnet wide_focus: the reference's Wide_Focus convolutions all have a bias (FCT.py:110-113)")
    z1, z2, z3 = _MultiConvFn.apply(x, c1.weight, c2.weight, c3.weight, tuple(c.dilation[0] for c in (c1, c2, c3)))
    seeds = (_next_seed(), _next_seed(), _next_seed(), _next_seed())
    added = _WfCombineFn.apply(z1, z2, z3, c1.bias, c2.bias, c3.bias, p, seeds[:3])
    z4 = _ConvFn.apply(added, None, c4.weight, None, None, 1, c4.dilation[0] * (c4.kernel_size[0] - 1) // 2, c4.dilation[0], 0,
                       ACT_NONE, False)
    return _WfOutFn.apply(z4, c4.bias, res, p, seeds[3])


class _LayerNormFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, gamma, beta, eps):
        x, ld = nhwc(x)
        B, H, W, C = x.shape
        y = _new(B, H, W, C, x)
        call("hn_layernorm_fwd", x.data_ptr(), y.data_ptr(), gamma.data_ptr(), beta.data_ptr(), B * H * W, C, ld, C,
             eps, _stream())
        ctx.eps = eps
        ctx.beta = beta
        ctx.save_for_backward(x, gamma)
        return y

    @staticmethod
    def backward(ctx, dy):
        x, gamma = ctx.saved_tensors
        x, ld = nhwc(x)
        dy, lddy = nhwc(dy)
        B, H, W, C = x.shape
        dx = _new(B, H, W, C, x)
        # dgamma / dbeta straight into the gradient sink when there is one
        kg = _sink_for(gamma) if gamma.dtype == torch.float32 else None
        kb = _sink_for(ctx.beta) if (kg is not None and ctx.beta.dtype == torch.float32) else None
        sunk = kg is not None and kb is not None
        dg, db = (kg, kb) if sunk else (_f32(C, x), _f32(C, x))
        call("hn_layernorm_bwd", x.data_ptr(), dy.data_ptr(), dx.data_ptr(), gamma.data_ptr(), dg.data_ptr(),
             db.data_ptr(), B * H * W, C, ld, lddy, C, ctx.eps, _stream())
        if sunk:
            _sink_done(gamma.data_ptr())
            _sink_done(ctx.beta.data_ptr())
            return dx, None, None, None
        return dx, dg, db, None


def layernorm_c(x, gamma, beta, eps=1e-5):
    """nn.LayerNorm(C) applied on the channel axis (FCT.py:97-99)."""
    return _LayerNormFn.apply(x, gamma, beta, eps)


# ----------------------------------------------------------------------------
# conv-projected attention (FCT.py:24-79)
# ----------------------------------------------------------------------------
def _p3(a, b, c):
    """HOST array of three device pointers (None -> null) for the per-projection parameter arguments."""
    import ctypes
    return (ctypes.c_void_p * 3)(*(t.data_ptr() if t is not None else None for t in (a, b, c)))


class _DwLnQKVFn(torch.autograd.Function):
    """q,k,v = LayerNorm_C(ReLU(depthwise3x3(x))) for the three projections, one pass over x.  The parameters are the
    reference's twelve separate tensors (conv_{q,k,v}.weight/.bias, layernorm_{q,k,v}.weight/.bias): the kernels take
    three pointers per kind, and each parameter gradient is reduced straight into the gradient sink when there is one."""

    @staticmethod
    def forward(ctx, x, eps, *params):
        # params: wq, wk, wv [C,1,3,3]; bq, bk, bv [C] or None; gq, gk, gv; beq, bek, bev  (fp32)
        x, ld = nhwc(x)
        B, H, W, C = x.shape
        w, b, g, be = params[0:3], params[3:6], params[6:9], params[9:12]
        qkv = torch.empty((3, B, H, W, C), dtype=torch.bfloat16, device=x.device)
        call("hn_dwconv_relu_ln_fwd", x.data_ptr(), qkv[0].data_ptr(), qkv[1].data_ptr(), qkv[2].data_ptr(),
             _p3(*w), _p3(*b), _p3(*g), _p3(*be), B, H, W, C, ld, C, eps, _stream())
        ctx.eps = eps
        ctx.params = params
        ctx.save_for_backward(x)
        return qkv[0], qkv[1], qkv[2]

    @staticmethod
    def backward(ctx, dq, dk, dv):
        (x,) = ctx.saved_tensors
        params = ctx.params
        w, b, g, be = params[0:3], params[3:6], params[6:9], params[9:12]
        x, ld = nhwc(x)
        B, H, W, C = x.shape
        dq, dk, dv = dq.contiguous(), dk.contiguous(), dv.contiguous()
        dpre = torch.empty((3, B, H, W, C), dtype=torch.bfloat16, device=x.device)
        # gradient targets: the sink's arena slice where there is one, zero-pool scratch otherwise
        tgt, sunk = [], []
        for t in params:
            if t is None:
                tgt.append(None); sunk.append(False)
                continue
            k = _sink_for(t) if (t.dtype == torch.float32 and t.requires_grad) else None
            sunk.append(k is not None)
            tgt.append(k if k is not None else _f32(t.numel(), x))
        dw, db, dg, dbe = tgt[0:3], tgt[3:6], tgt[6:9], tgt[9:12]
        call("hn_dwconv_relu_ln_bwd_pre", x.data_ptr(), dq.data_ptr(), dk.data_ptr(), dv.data_ptr(),
             dpre[0].data_ptr(), dpre[1].data_ptr(), dpre[2].data_ptr(), _p3(*w), _p3(*b), _p3(*g), _p3(*dg), _p3(*dbe),
             _p3(*db), B, H, W, C, ld, C, ctx.eps, _stream())
        dx = _new(B, H, W, C, x)
        call("hn_dwconv_bwd_data_weight", x.data_ptr(), dpre[0].data_ptr(), dpre[1].data_ptr(), dpre[2].data_ptr(),
             dx.data_ptr(), _p3(*w), _p3(*dw), B, H, W, C, ld, C, C, _stream())
        grads = []
        for t, k, s in zip(params, tgt, sunk):
            if t is None:
                grads.append(None)
            elif s:
                _sink_done(t.data_ptr())
                grads.append(None)
            else:
                grads.append(k.view(t.shape))
        return (dx, None, *grads)


def dwconv_relu_ln_qkv(x, weights, biases, gammas, betas, eps=1e-5):
    """weights / biases / gammas / betas: 3-tuples (q, k, v) of the modules' parameters (biases may hold None)."""
    return _DwLnQKVFn.apply(x, eps, *weights, *biases, *gammas, *betas)


# tcgen05 / TMEM / TMA attention for the exponential-bound shapes (D in {4, 8}, two heads, N % 128 == 0: FCT blocks
# 1, 2, 8, 9 = 99 % of the score elements); every other shape takes the mma.sync kernels.  POLY: every n-th pair of
# exponentials on the FMA pipe (measured on B200, scripts/attn_bench.py: n = 3 is the fastest for both directions).
ATTN_FWD_TC = True
ATTN_FWD_TC_POLY = 3
ATTN_BWD_TC = True
ATTN_BWD_TC_POLY = 3


class _AttnFn(torch.autograd.Function):
    """softmax(Q K^T / sqrt(d)) V per head, flash style; q,k,v: [B,H,W,heads*d] (N = H*W tokens)."""

    @staticmethod
    def forward(ctx, q, k, v, heads):
        q, ldq = nhwc(q)
        k, ldk = nhwc(k)
        v, ldv = nhwc(v)
        B, H, W, C = q.shape
        N, D = H * W, C // heads
        scale = 1.0 / math.sqrt(D)
        o = _new(B, H, W, C, q)
        lse = torch.empty((B, heads, N), dtype=torch.float32, device=q.device)
        if ATTN_FWD_TC and heads == 2 and _lib.load().hn_attn_fwd_tc_supported(N, D, ldq, ldk, ldv, C):
            call("hn_attn_fwd_tc", q.data_ptr(), k.data_ptr(), v.data_ptr(), o.data_ptr(), lse.data_ptr(), B, N, heads, D,
                 ldq, ldk, ldv, C, scale, ATTN_FWD_TC_POLY, _stream())
        else:
            call("hn_attn_fwd", q.data_ptr(), k.data_ptr(), v.data_ptr(), o.data_ptr(), lse.data_ptr(), B, N, heads, D,
                 ldq, ldk, ldv, C, scale, _stream())
        ctx.heads, ctx.scale = heads, scale
        ctx.save_for_backward(q, k, v, o, lse)
        return o

    @staticmethod
    def backward(ctx, do):
        q, k, v, o, lse = ctx.saved_tensors
        q, ldq = nhwc(q)
        k, ldk = nhwc(k)
        v, ldv = nhwc(v)
        do, lddo = nhwc(do)
        B, H, W, C = q.shape
        heads = ctx.heads
        N, D = H * W, C // heads
        dqkv = torch.empty((3, B, H, W, C), dtype=torch.bfloat16, device=q.device)
        lib = _lib.load()
        if ATTN_BWD_TC and heads == 2 and lib.hn_attn_bwd_tc_supported(N, D, ldq, ldk, ldv, C, lddo, C, C, C):
            nws = lib.hn_attn_bwd_tc_workspace(B, N, heads, D)
            ws = torch.empty(nws, dtype=torch.float32, device=q.device)
            call("hn_attn_bwd_tc", q.data_ptr(), k.data_ptr(), v.data_ptr(), o.data_ptr(), do.data_ptr(), lse.data_ptr(),
                 ws.data_ptr(), nws, dqkv[0].data_ptr(), dqkv[1].data_ptr(), dqkv[2].data_ptr(), B, N, heads, D, ldq, ldk,
                 ldv, C, lddo, C, C, C, ctx.scale, ATTN_BWD_TC_POLY, _stream())
            return dqkv[0], dqkv[1], dqkv[2], None
        nws = _lib.load().hn_attn_bwd_workspace(B, N, heads, D)
        ws = torch.empty(nws, dtype=torch.float32, device=q.device)
        call("hn_attn_bwd", q.data_ptr(), k.data_ptr(), v.data_ptr(), o.data_ptr(), do.data_ptr(), lse.data_ptr(),
             ws.data_ptr(), nws, dqkv[0].data_ptr(), dqkv[1].data_ptr(), dqkv[2].data_ptr(), B, N, heads, D, ldq, ldk,
             ldv, C, lddo, C, C, C, ctx.scale, _stream())
        return dqkv[0], dqkv[1], dqkv[2], None


def attention(q, k, v, heads):
    return _AttnFn.apply(q, k, v, heads)
