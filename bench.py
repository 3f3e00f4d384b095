#!/usr/bin/env python
"""Headline benchmark: train frames/sec (fwd+bwd, 256x256) of the hybrid segmentation nets.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload fct|unet|ae32k_image|...]
  python bench.py --impl reference ...      # the reference's own classes on the host cores (oracle/_ref)

Own arm: one process per GPU (torchrun sets RANK/LOCAL_RANK/WORLD_SIZE); a step is
zero_grad -> forward -> loss -> backward (-> NCCL gradient all-reduce when N>1) -> fused AdamW (+ weight repack)
over one batch of synthetic frames.  One JSON line on rank 0.  The default workload is FCT, batch 32 =
two 16-frame clips per GPU (BASELINE configs[2] / configs[4]); at N=1 the same line carries UNet B=32 and
Autoencoder32K B=16 under "extra", each with its own roofline / e2e / cpu_baseline, timed in the same run.
"""
import argparse
import contextlib
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

# workload -> (model key, per-GPU batch, loss, reference/BASELINE config string)
WORKLOADS = {
    "fct": ("fct", 32, "dice", "FCT conv-attention net, 2 x 16-frame clips (batch 32) per GPU, fwd+bwd (BASELINE configs[2], configs[4])"),
    "unet": ("unet", 32, "dice", "UNet fwd+bwd (BASELINE configs[0] shape on GPU)"),
    "ae32k_image": ("ae32k_image", 16, "mse", "Autoencoder32K('image') 256->32K-latent image2image, fwd+bwd (BASELINE configs[1])"),
    "autoencoder": ("autoencoder", 64, "mse", "AutoEncoder (8K latent) fwd+bwd"),
    "i2i2m": ("i2i2m", 16, "dice", "Image2Image2Mask (AE -> UNet) fwd+bwd (BASELINE configs[3] ii)"),
    "latent2mask": ("latent2mask", 16, "dice", "latent-to-mask: frozen Encoder_32K -> Decoder_32K('mask') training (BASELINE configs[3] i)"),
}
DEFAULT_WORKLOAD = "fct"
EXTRA_WORKLOADS = ("unet", "ae32k_image")     # carried under "extra" by the default N=1 line
CLIP_FRAMES = 16                              # BASELINE configs[4]: 16-frame clips
# frames per step of the CPU arm (bounded sample: FCT B=32 would need ~60 s and >10 GB per step on 16 cores)
CPU_FRAMES = {"fct": 1, "unet": 4, "ae32k_image": 2, "autoencoder": 8, "i2i2m": 2, "latent2mask": 2}
LR = 1e-4
MUFU_TEXP_S = 4.647        # ex2.approx results/s (x1e12), scripts/mufu_bench.cu on B200 @ 1965 MHz


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return {"hbm": d["hbm_gbs"], "tensor": d["bf16_tflops_sustained"], "tensor_burst": d["bf16_tflops"], "src": "measured"}
    return {"hbm": 6650.0, "tensor": 1400.0, "tensor_burst": 1590.0, "src": "fallback"}


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe).  One streaming
    `nvidia-smi -lms 20` process is started well before the timed region (its start-up costs ~100 ms); rows are
    time-stamped on arrival and the summary uses the rows that fell inside the region."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.t0, self.t1, self.proc = index, [], None, None, None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-i",
                                          str(index), "-lms", "20"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL,
                                         text=True, bufsize=1)
        except Exception:
            self.proc = None
        self.t = threading.Thread(target=self._run, daemon=True)
        self.t.start()

    def _run(self):
        if self.proc is None:
            return
        for line in self.proc.stdout:
            cells = [c.strip() for c in line.strip().split(",")]
            if len(cells) >= 6:
                self.rows.append((time.perf_counter(), cells))

    def __enter__(self):
        self.t0 = time.perf_counter()
        return self

    def __exit__(self, *a):
        self.t1 = time.perf_counter()
        time.sleep(0.03)                      # let the last in-region sample arrive
        if self.proc is not None:
            self.proc.terminate()
        self.t.join(timeout=3)

    def summary(self):
        inside = [c for ts, c in self.rows if self.t0 is not None and self.t0 <= ts <= (self.t1 or ts) + 0.03]
        rows = inside or [c for _, c in self.rows[-3:]]
        sm = sorted(int(r[0]) for r in rows if r[0].isdigit())
        mx = [int(r[1]) for r in rows if r[1].isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({n for r in rows for n, v in zip(names, r[2:6]) if v.lower().startswith("active")})
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(inside)}


def make_batch(B, seed, device, pinned=False):
    g = torch.Generator().manual_seed(seed)
    x = torch.rand(B, 3, 256, 256, generator=g)
    y = (torch.rand(B, 1, 256, 256, generator=g) > 0.5).float()
    if pinned:
        return x.pin_memory(), y.pin_memory()
    return x.to(device), y.to(device)


def loss_fn(kind, out, x, y):
    """Reference formulas in plain torch (CPU arm, torch-eager bar): nn.MSELoss / DiceLoss (Metrics.py:14-22)."""
    out = out[-1] if isinstance(out, tuple) else out
    if kind == "mse":
        return torch.nn.functional.mse_loss(out, x)
    p, t = out[:, 0].reshape(-1), y[:, 0].reshape(-1)
    return 1.0 - (2.0 * (p * t).sum() + 1.0) / (p.sum() + t.sum() + 1.0)


# ---------------------------------------------------------------------------------------------
# per-kernel roofline pass: CUDA events around every C-ABI call of one extra step
# ---------------------------------------------------------------------------------------------
def _conv_work(args):
    d = args[0]
    n = 4 * d.pshuf if d.pshuf else d.Cout
    flops = 2.0 * d.B * d.OH * d.OW * n * d.KH * d.KW * (d.C1 + d.C2)
    if d.mode == 1 and d.stride > 1:
        flops /= d.stride * d.stride
    byt = 2.0 * d.B * (d.H * d.W * (d.C1 + d.C2) / (4 if d.ups else 1) + d.OH * d.OW * n)
    return flops, byt


def _attn_work(args, bwd):
    if bwd:
        B, N, heads, D = args[11], args[12], args[13], args[14]
    else:
        B, N, heads, D = args[5], args[6], args[7], args[8]
    f = 4.0 * B * heads * N * N * D
    # softmax exponentials: the one-pass backward (D <= 8) recomputes P once, the two-kernel form (D >= 16) twice
    exps = float(B) * heads * N * N * ((1 if D <= 8 else 2) if bwd else 1)
    return (2.5 * f if bwd else f), 2.0 * B * N * heads * D * (8 if bwd else 4), exps


def _bn_work(P, C, passes):
    return 0.0, 2.0 * P * C * passes


WORK = {
    # BatchNorm passes (bf16 tensors of P x C): arguments as in include/hybridnet.h
    "hn_bn_apply": lambda a: _bn_work(a[7], a[8], 3 if a[1] else 2),
    "hn_bn_apply_train": lambda a: _bn_work(a[15], a[16], 3 if a[1] else 2),
    "hn_bn_relu_bwd_reduce": lambda a: _bn_work(a[6], a[7], 2),
    "hn_bn_relu_bwd_apply": lambda a: _bn_work(a[9], a[10], 3),
    "hn_bn_bwd_reduce": lambda a: _bn_work(a[6], a[7], 4 if a[3] else 3),
    "hn_bn_bwd_apply": lambda a: _bn_work(a[8], a[9], 3),
    "hn_conv2d_fwd": lambda a: _conv_work(a),
    "hn_conv_tc_fwd": lambda a: _conv_work(a),
    "hn_conv_sc_fwd": lambda a: _conv_work(a),
    "hn_conv2d_wgrad": lambda a: _conv_work(a),
    "hn_conv_wgrad_tc": lambda a: _conv_work(a),
    "hn_conv_wgrad_sc": lambda a: _conv_work(a),
    "hn_attn_fwd": lambda a: _attn_work(a, False),
    "hn_attn_bwd": lambda a: _attn_work(a, True),
    "hn_attn_fwd_tc": lambda a: _attn_work(a, False),          # same argument positions for B, N, heads, D
    "hn_attn_bwd_tc": lambda a: _attn_work(a, True),
}


class KernelProfiler:
    def __init__(self, detail=False):
        self.detail = detail
        self.events = []        # (name, start, end, flops, bytes)

    @contextlib.contextmanager
    def __call__(self, name, args):
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        w = WORK[name](args) if name in WORK else (0.0, 0.0)
        fl, by, ex = w[0], w[1], (w[2] if len(w) > 2 else 0.0)
        if self.detail and name.startswith("hn_conv"):
            d = args[0]
            name = (f"{name}[{d.B}x{d.H}x{d.W} C{d.C1}+{d.C2}->{4 * d.pshuf if d.pshuf else d.Cout} k{d.KH} s{d.stride} "
                    f"d{d.dil} m{d.mode}{' st' if (len(args) > 8 and args[-3]) else ''}]")
        s.record()
        yield
        e.record()
        self.events.append((name, s, e, fl, by, ex))

    def table(self):
        torch.cuda.synchronize()
        agg = {}
        for name, s, e, fl, by, ex in self.events:
            a = agg.setdefault(name, {"ms": 0.0, "n": 0, "flops": 0.0, "bytes": 0.0, "exps": 0.0})
            a["exps"] += ex
            a["ms"] += s.elapsed_time(e)
            a["n"] += 1
            a["flops"] += fl
            a["bytes"] += by
        return agg


def ncu_traffic(workload, kernel):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of `kernel`, from the committed ncu --set full
    capture of this workload (profiles/*_traffic.json, newest first); (None, None) when there is none."""
    import glob
    for f in sorted(glob.glob(os.path.join(os.path.dirname(os.path.abspath(__file__)), "profiles", "*_traffic.json")), reverse=True):
        try:
            d = json.load(open(f))
            return d[workload][kernel]["traffic_bytes_per_launch"], d["source"]
        except (KeyError, ValueError, OSError):
            continue
    return None, None


def roofline_from(agg, steps, peaks):
    if any("[" in k for k in agg):            # --kernel-detail rows: the roofline object is per kernel, not per geometry
        merged = {}
        for k, v in agg.items():
            m = merged.setdefault(k.split("[")[0], {"ms": 0.0, "n": 0, "flops": 0.0, "bytes": 0.0, "exps": 0.0})
            for f in m:
                m[f] += v.get(f, 0.0)
        agg = merged
    total = sum(a["ms"] for a in agg.values())
    name, a = max(agg.items(), key=lambda kv: kv[1]["ms"])
    sec = a["ms"] / 1e3
    tf = a["flops"] / sec / 1e12 if sec > 0 else 0.0
    gbs = a["bytes"] / sec / 1e9 if sec > 0 else 0.0
    tensor_frac, hbm_frac = tf / peaks["tensor"], gbs / peaks["hbm"]
    if a["flops"] > 0 and tensor_frac >= hbm_frac:
        r = {"bound": "tensor", "achieved": round(tf, 2), "peak": peaks["tensor"], "unit": "TFLOP/s", "frac": round(tensor_frac, 4)}
    else:
        r = {"bound": "hbm", "achieved": round(gbs, 1), "peak": peaks["hbm"], "unit": "GB/s", "frac": round(hbm_frac, 4)}
    r.update({"kernel": name, "traffic": None, "peak_source": peaks["src"] + (" (sustained)" if r["bound"] == "tensor" else ""),
              "launches_per_step": a["n"] // steps, "avg_launch_us": round(a["ms"] * 1e3 / a["n"], 1),
              "share_of_step": round(a["ms"] / total, 3),
              "algorithmic_per_launch": {"GFLOP": round(a["flops"] / a["n"] / 1e9, 3), "MB": round(a["bytes"] / a["n"] / 1e6, 2)}})
    if a.get("exps", 0) > 0:
        # attention with head dims 4..16 is bound by the softmax exponentials (MUFU ex2, 16 results/clk/SM =
        # 4.647e12/s measured on this pool's B200 with scripts/mufu_bench.cu), not by the tensor pipe
        r["xu"] = {"bound": "mufu_ex2", "achieved": round(a["exps"] / sec / 1e12, 3), "peak": MUFU_TEXP_S, "unit": "Texp/s",
                   "frac": round(a["exps"] / sec / 1e12 / MUFU_TEXP_S, 4)}
    top = sorted(agg.items(), key=lambda kv: -kv[1]["ms"])[:6]
    r["top_kernels"] = {k: {"share": round(v["ms"] / total, 3), "TFLOP/s": round(v["flops"] / (v["ms"] / 1e3) / 1e12, 1) if v["ms"] else 0,
                            "GB/s": round(v["bytes"] / (v["ms"] / 1e3) / 1e9, 0) if v["ms"] else 0} for k, v in top}
    return r


# ---------------------------------------------------------------------------------------------
# reference arm / cpu baseline: the oracle (CPU restatement of the reference's PyTorch path)
# ---------------------------------------------------------------------------------------------
def cpu_reference(workload, frames, steps, warmup, infer=False):
    """The same step on the host cores, fp32, all threads.  FCT / UNet / AutoEncoder / Image2Image2Mask: the
    reference's UNMODIFIED classes (oracle/_ref, staged by oracle/make_ref.py) with the reference's DiceLoss /
    nn.MSELoss and torch.optim.AdamW -> kind "reference".  Autoencoder32K family (bytecode only in the reference):
    the oracle port -> kind "port".  Returns (frames/s, s/step, cores, kind)."""
    from oracle import nets, ref_loader
    key, _, kind, _ = WORKLOADS[workload]
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    x, y = make_batch(frames, 1, "cpu")
    times = []
    if ref_loader.available() and key in ("fct", "unet", "autoencoder", "i2i2m"):
        classes, RefDice = ref_loader.load()
        torch.manual_seed(0)
        model = classes[key]()
        model.train(not infer)
        crit = RefDice() if kind == "dice" else torch.nn.MSELoss()
        opt = torch.optim.AdamW(model.parameters(), lr=LR)
        for i in range(warmup + steps):
            t0 = time.perf_counter()
            if infer:
                with torch.no_grad():
                    model(x)
            else:
                opt.zero_grad()
                out = model(x)
                out = out[-1] if isinstance(out, tuple) else out
                loss = crit(out, y if kind == "dice" else x)
                loss.backward()
                opt.step()
            if i >= warmup:
                times.append(time.perf_counter() - t0)
        sec = sum(times) / len(times)
        return frames / sec, sec, cores, "reference"
    sd = nets.init_state_dict(key, 0)
    frozen = (lambda k: k.startswith("0.")) if key == "latent2mask" else (lambda k: False)
    params = {k: (v.clone().requires_grad_(True) if v.is_floating_point() and "running" not in k and not frozen(k) else v)
              for k, v in sd.items()}
    opt = torch.optim.AdamW([p for p in params.values() if p.requires_grad], lr=LR)
    ctx = nets.Ctx(training=not infer, dropout="off" if infer else "on")
    for i in range(warmup + steps):
        t0 = time.perf_counter()
        if infer:
            with torch.no_grad():
                nets.FORWARDS[key](params, x, ctx)
        else:
            opt.zero_grad()
            out = nets.FORWARDS[key](params, x, ctx)
            loss_fn(kind, out, x, y).backward()
            opt.step()
        if i >= warmup:
            times.append(time.perf_counter() - t0)
    sec = sum(times) / len(times)
    return frames / sec, sec, cores, "port"


def cpu_sample_text(kind_impl, frames, steps, infer, sec):
    src = ("the reference's unmodified classes from oracle/_ref (torch fp32 CPU)" if kind_impl == "reference"
           else "oracle port of the reference's PyTorch path (torch fp32 CPU; the reference ships this family as bytecode only)")
    what = "fwd" if infer else "zero_grad+fwd+loss+bwd+AdamW"
    return f"{frames} frames x {steps} timed {what} steps of the same net, {src}, {sec:.2f} s/step"


def torch_eager_bar(key, kind, B, dev, steps=5, warmup=2):
    """The library bar: the same modules with torch eager on this B200 --
    cuDNN / cuBLAS / SDPA, fp32 and bf16 autocast + channels_last -- same batch, same step definition.  A stated
    baseline, never the product path.  frames/s per variant, or the error text."""
    from oracle import nets, ref_loader
    out = {}
    x, y = make_batch(B, 1, dev)
    for variant in ("fp32", "bf16_autocast_channels_last"):
        try:
            torch.manual_seed(0)
            if ref_loader.available() and key in ("fct", "unet", "autoencoder", "i2i2m"):
                classes, _ = ref_loader.load()
                model = classes[key]().to(dev).train()
                fwd = lambda xx: model(xx)
                params = list(model.parameters())
                if variant != "fp32":
                    model = model.to(memory_format=torch.channels_last)
            else:
                sd = nets.init_state_dict(key, 0)
                frozen = (lambda k: k.startswith("0.")) if key == "latent2mask" else (lambda k: False)
                pd = {k: (v.to(dev).requires_grad_(True) if v.is_floating_point() and "running" not in k and not frozen(k)
                          else v.to(dev)) for k, v in sd.items()}
                ctx = nets.Ctx(training=True, dropout="on")
                fwd = lambda xx: nets.FORWARDS[key](pd, xx, ctx)
                params = [p for p in pd.values() if p.requires_grad]
            opt = torch.optim.AdamW(params, lr=LR)
            xx = x if variant == "fp32" else x.contiguous(memory_format=torch.channels_last)

            def one():
                opt.zero_grad(set_to_none=True)
                if variant == "fp32":
                    o = fwd(xx)
                else:
                    with torch.autocast("cuda", dtype=torch.bfloat16):
                        o = fwd(xx)
                o = o[-1] if isinstance(o, tuple) else o
                loss_fn(kind, o.float(), x, y).backward()
                opt.step()
            for _ in range(warmup):
                one()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(steps):
                one()
            e1.record()
            torch.cuda.synchronize()
            out[variant] = round(B * steps / (e0.elapsed_time(e1) / 1e3), 1)
            del opt, params
        except Exception as e:      # e.g. out of memory for a materialised 16384^2 attention
            out[variant] = f"failed: {type(e).__name__}: {str(e)[:120]}"
        torch.cuda.empty_cache()
    return out


def run_ours(args, workload, B, rank, world, local, dev, steps, warmup, with_cpu, with_eager):
    """One workload through the product: returns the result dict on rank 0 (None elsewhere)."""
    import torch.distributed as dist
    import hybridnet
    from hybridnet import _lib, ops, parallel
    key, _, kind, cfg_name = WORKLOADS[workload]
    metric, unit = "train frames/sec (fwd+bwd, 256x256)", "frames/s"
    if args.infer:
        metric = "inference frames/sec (eval-mode fwd, 256x256)"
    config = make_config(args, workload, B, world)
    torch.manual_seed(0)
    model = hybridnet.MODELS[key]().to(dev).train()
    if args.infer:
        model.eval()
    if args.no_dropout:
        for m in model.modules():
            if isinstance(m, (torch.nn.Dropout, torch.nn.Dropout2d)):
                m.p = 0.0
    ops.manual_seed(1234 + rank)
    dead = parallel.used_parameter_names(key)
    used = {n for n, p in model.named_parameters() if p.requires_grad and n not in dead}
    arena = parallel.GradArena(model, used=used, buckets=4).direct()     # kernels reduce dW / BN grads straight into the arena
    opt = hybridnet.optim.AdamW(model.parameters(), arena, lr=LR)        # flat fused AdamW; folds the 1/world of the mean
    crit = hybridnet.DiceLoss() if kind == "dice" else hybridnet.MSELoss()
    x, y = make_batch(B, 1 + rank, dev)
    xh, yh = make_batch(B, 1 + rank, dev, pinned=True)
    zero_cap = sum(p.numel() for p in model.parameters()) + (1 << 18)

    def infer_step(xd, yd):
        with torch.no_grad():
            out = model(xd)
        return out[-1] if isinstance(out, tuple) else out

    def step(xd, yd):
        if args.infer:
            return infer_step(xd, yd)
        opt.zero_grad()                         # one memset of the gradient arena
        ops.begin_step(dev, zero_cap)           # all zero-initialised fp32 scratch of the step: one memset
        out = model(xd)
        out = out[-1] if isinstance(out, tuple) else out
        loss = crit(out, yd if kind == "dice" else xd)      # fused loss kernel (one launch), dL/dy seed kernel in backward
        loss.backward()
        arena.finish()                          # joins the bucketed NCCL all-reduce (sum)
        opt.step()                              # fused AdamW over the flat arenas (x 1/world) + one-launch weight repack
        return loss

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    clocks = ClockSampler(local)                # streaming nvidia-smi, started ahead of the timed region
    for _ in range(warmup):
        step(x, y)
    barrier()
    # the public API for a training loop: the step recorded once as a CUDA graph, replayed per batch
    # (hybridnet.parallel.CapturedStep); --no-graph times the eager dispatch instead
    run = step
    captured = None
    if not args.no_graph:
        captured = parallel.CapturedStep(step, x, y, warmup=1)
        run = captured
        for _ in range(warmup):
            run()
        barrier()
    # ---- timed region 1: inputs resident in HBM ------------------------------------------------
    _lib.stats["launches"] = 0
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with clocks:
        barrier()
        e0.record()
        for _ in range(steps):
            loss = run() if captured else run(x, y)     # captured: the batch already sits in the static buffers
        e1.record()
        barrier()
    launches = captured.launches * steps if captured else _lib.stats["launches"]
    ms = e0.elapsed_time(e1)
    # ---- timed region 2: end to end through the public API with host buffers ---------------------
    barrier()
    e2, e3 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e2.record()
    last = 0.0
    out_host = None
    if captured:
        # the batch of step i+1 crosses PCIe on the copy stream while step i computes (CapturedStep.prefetch);
        # every step's H2D copy and D2H read are inside the timed region
        captured.prefetch(xh, yh)
        prev = None
        for i in range(steps):
            out = captured()
            if i + 1 < steps:
                captured.prefetch(xh, yh)
            # the step's result comes back to pinned host memory on the download stream (--infer: the whole mask /
            # reconstruction; training: the loss) and is read here one step later, while the next step already runs:
            # every step's result is read inside the timed region, none of the reads stalls the device
            h = captured.download()
            if prev is not None:
                last = float(prev.wait().view(-1)[0])
            prev = h
        if prev is not None:
            out_host = prev.wait()
            last = float(out_host.view(-1)[0])
    else:
        for _ in range(steps):
            xd = xh.to(dev, non_blocking=True)
            yd = yh.to(dev, non_blocking=True)
            out = step(xd, yd)
            if args.infer:
                if out_host is None:
                    out_host = torch.empty(out.shape, dtype=out.dtype).pin_memory()
                out_host.copy_(out)
                last = float(out_host.view(-1)[0])
            else:
                last = float(out.item())
    e3.record()
    barrier()
    ms_e2e = e2.elapsed_time(e3)
    t = torch.tensor([ms, ms_e2e], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms, ms_e2e = float(t[0]), float(t[1])

    result = None
    if rank == 0:
        peaks = load_peaks()
        # ---- roofline pass (not part of either timed region) -----------------------------------
        prof = KernelProfiler(args.kernel_detail)
        _lib.set_profiler(prof)
        psteps = 2
        arena.exchange = False                  # rank-0-only pass: no collectives (the other ranks are at the barrier)
        # per-kernel durations are taken one kernel at a time: the weight gradients stay on the main stream here (in the
        # timed step they run on a side stream beside the input gradient, which would stretch both event intervals)
        overlap, ops.WGRAD_OVERLAP = ops.WGRAD_OVERLAP, False
        for _ in range(psteps):
            step(x, y)
        ops.WGRAD_OVERLAP = overlap
        arena.exchange = True
        _lib.set_profiler(None)
        agg = prof.table()
        roof = roofline_from(agg, psteps, peaks)
        roof["traffic"], roof["traffic_source"] = ncu_traffic(workload, roof["kernel"])
        if args.kernel_table:
            tot = sum(a["ms"] for a in agg.values())
            print(f"---- {workload} (eager per-call CUDA events; launches under ~25 us carry the event floor)", file=sys.stderr)
            for k, a in sorted(agg.items(), key=lambda kv: -kv[1]["ms"]):
                print(f"{k:34s} {a['ms'] / psteps:9.3f} ms/step {100 * a['ms'] / tot:5.1f}%  n={a['n'] // psteps:4d}  "
                      f"{a['flops'] / max(a['ms'], 1e-9) / 1e9:8.1f} TFLOP/s {a['bytes'] / max(a['ms'], 1e-9) / 1e6:8.0f} GB/s",
                      file=sys.stderr)
        from oracle.nets import GFLOP_PER_FRAME
        gf = GFLOP_PER_FRAME[key][0 if args.infer else 1]
        fps = B * world * steps / (ms / 1e3)
        fps_e2e = B * world * steps / (ms_e2e / 1e3)
        result = {"metric": metric, "value": round(fps, 2), "unit": unit, "n_gpus": world, "steps": steps,
                  "warmup": warmup, "ms_per_step": round(ms / steps, 3), "higher_is_better": True,
                  "scaling": "strong" if args.strong_clips else "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
                  "config": config, "dispatch": "cuda-graph replay of the captured step" if captured else "eager",
                  "clocks": clocks.summary(),
                  "e2e": {"value": round(fps_e2e, 2), "unit": unit,
                          "overlap": "H2D of batch i+1 and D2H of the result of step i-1 (loss; the whole output under --infer) on copy streams during step i" if captured else "none",
                          "h2d_bytes_per_step": int(xh.numel() * 4 + yh.numel() * 4),
                          "d2h_bytes_per_step": int(out_host.numel() * out_host.element_size()) if args.infer else 4,
                          "last_loss": round(last, 5)},
                  "gpu_launches": launches,
                  "model_tensor_roofline": {"GFLOP_per_frame_fwd" if args.infer else "GFLOP_per_frame_fwd_bwd": gf, "achieved_TFLOPs": round(fps / world * gf / 1e3, 2),
                                            "frac_of_sustained_bf16": round(fps / world * gf / 1e3 / peaks["tensor"], 4)},
                  "roofline": roof}
    # release this workload's device memory before the next one
    ops.set_grad_sink(None, None)
    ops.end_zero_pool()
    if world > 1:
        dist.barrier()
    # the captured graph holds NCCL's persistent plans: it must die before the communicator does
    del run, captured, step, infer_step, opt, arena, model, crit
    import gc
    gc.collect()
    torch.cuda.synchronize()
    torch.cuda.empty_cache()
    if rank == 0 and with_eager:
        result["torch_eager_b200"] = torch_eager_bar(key, kind, B, dev)
    if rank == 0 and with_cpu:
        frames = CPU_FRAMES[workload]
        cfps, sec, cores, kimpl = cpu_reference(workload, frames, 2, 1, args.infer)
        result["cpu_baseline"] = {"value": round(cfps, 4), "unit": unit, "cores": cores, "kind": kimpl,
                                  "sample": cpu_sample_text(kimpl, frames, 2, args.infer, sec)}
    return result


def make_config(args, workload, B, world):
    key, _, _, cfg_name = WORKLOADS[workload]
    cfg = {"workload": cfg_name, "frames_per_gpu": B, "global_batch": B * world, "frame": "3x256x256 fp32 in [0,1)",
           "parallelism": f"dp{world}" if world > 1 else "single",
           "step": "eval forward" if args.infer else "zero_grad, forward, loss, backward, gradient all-reduce (N>1), AdamW",
           "l2": "activations per step >> 126 MB L2 (no flush needed)"}
    if workload == "fct" and args.strong_clips:
        cfg["workload"] = (f"FCT conv-attention net, {args.strong_clips} x {CLIP_FRAMES}-frame clip(s) in total split by frame "
                           f"across the ranks ({B} frames per GPU), fwd+bwd (BASELINE configs[4])")
        cfg["clips"] = {"frames_per_clip": CLIP_FRAMES, "clips_total": args.strong_clips,
                        "sharding": f"{args.strong_clips} clip(s) split by frame across the ranks: {B} frames per rank"}
    elif workload == "fct" and B % CLIP_FRAMES == 0:
        cfg["clips"] = {"frames_per_clip": CLIP_FRAMES, "clips_per_gpu": B // CLIP_FRAMES, "sharding": "whole clips per rank"}
    return cfg


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default=DEFAULT_WORKLOAD, choices=list(WORKLOADS))
    ap.add_argument("--batch", type=int, default=0, help="frames per GPU (default: the workload's)")
    ap.add_argument("--clips", type=int, default=0, help="FCT: whole 16-frame clips per GPU (weak scaling; default 2 = batch 32)")
    ap.add_argument("--strong-clips", type=int, default=0, help="FCT: this many 16-frame clips IN TOTAL, split by frame across the ranks (strong scaling)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extra", action="store_true", help="default workload only (no UNet / Autoencoder32K under 'extra')")
    ap.add_argument("--torch-eager", action="store_true", help="also time the torch-eager (cuDNN/SDPA) library bar on this GPU")
    ap.add_argument("--no-dropout", action="store_true")
    ap.add_argument("--kernel-table", action="store_true", help="print the per-kernel time table to stderr")
    ap.add_argument("--kernel-detail", action="store_true", help="per conv geometry rows in the kernel table")
    ap.add_argument("--infer", action="store_true", help="eval-mode forward only (BASELINE configs[3] inference leg); "
                    "e2e copies the whole output back to pinned host memory")
    ap.add_argument("--no-graph", action="store_true", help="issue every kernel eagerly instead of replaying the captured step")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else max(args.warmup, 1)

    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    key, B0, kind, cfg_name = WORKLOADS[args.workload]
    B = args.batch or B0
    if args.workload == "fct" and args.clips:
        B = CLIP_FRAMES * args.clips
    if args.workload == "fct" and args.strong_clips:
        total = CLIP_FRAMES * args.strong_clips
        if total % world:
            raise SystemExit(f"--strong-clips: {total} frames do not split over {world} ranks")
        B = total // world
    metric, unit = "train frames/sec (fwd+bwd, 256x256)", "frames/s"
    if args.infer:
        metric = "inference frames/sec (eval-mode fwd, 256x256)"

    if args.impl == "reference":
        if rank != 0:
            return
        frames = CPU_FRAMES[args.workload]
        steps, warmup = max(1, args.steps), max(1, args.warmup)
        fps, sec, cores, kimpl = cpu_reference(args.workload, frames, steps, warmup, args.infer)
        sample = cpu_sample_text(kimpl, frames, steps, args.infer, sec)
        print(json.dumps({"impl": "reference", "metric": metric, "value": round(fps, 4), "unit": unit, "n_gpus": args.gpus,
                          "steps": steps, "warmup": warmup, "ms_per_step": round(sec * 1e3, 2), "higher_is_better": True,
                          "scaling": "strong" if args.strong_clips else "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                          "config": make_config(args, args.workload, B, world),
                          "cpu_baseline": {"value": round(fps, 4), "unit": unit, "cores": cores, "kind": kimpl, "sample": sample},
                          "e2e": {"value": round(fps, 4), "unit": unit, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))
        return

    import torch.distributed as dist
    from hybridnet import _lib
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        import datetime
        # keep stdout to the one JSON line: NCCL writes its banner / log lines to fd 1 from C, so fd 1 is pointed
        # at stderr for the whole run and the JSON line goes to the saved descriptor
        sys.stdout.flush()
        json_fd = os.dup(1)
        os.dup2(2, 1)
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
        if "HN_NCCL_DEBUG" in os.environ:
            os.environ["NCCL_DEBUG"] = os.environ["HN_NCCL_DEBUG"]
        dist.init_process_group("nccl", device_id=dev, timeout=datetime.timedelta(seconds=180))
    else:
        json_fd = None
    _lib.call("hn_check_device")
    with_cpu = world == 1 and not args.no_cpu_baseline
    result = run_ours(args, args.workload, B, rank, world, local, dev, args.steps, args.warmup, with_cpu, args.torch_eager)
    is_default = (args.workload == DEFAULT_WORKLOAD and not args.batch and not args.clips and not args.strong_clips
                  and not args.infer and not args.no_graph)
    if world == 1 and is_default and not args.no_extra:
        extra = {}
        for w in EXTRA_WORKLOADS:
            r = run_ours(args, w, WORKLOADS[w][1], rank, world, local, dev, args.steps, args.warmup, with_cpu, args.torch_eager)
            extra[w] = {k: r[k] for k in ("value", "unit", "ms_per_step", "config", "e2e", "gpu_launches", "clocks",
                                           "model_tensor_roofline", "roofline", "cpu_baseline", "torch_eager_b200") if k in r}
        result["extra"] = extra
    if rank == 0:
        if json_fd is None:
            print(json.dumps(result))
        else:
            os.write(json_fd, (json.dumps(result) + "\n").encode())
    if world > 1:
        dist.barrier()
        import gc
        gc.collect()
        torch.cuda.synchronize()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
