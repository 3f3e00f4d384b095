"""Diagnostic (not a test): how far is each arithmetic from the fp32 oracle on UNet / AE32K?
   product (CUDA kernels)  |  oracle with bf16 emulation  |  torch autocast(bf16) of the oracle."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import hybridnet
from oracle import nets
from tests.util import rel_l2, cosine
torch.backends.cudnn.allow_tf32 = False
torch.backends.cuda.matmul.allow_tf32 = False

def run_oracle(name, sd, x, y, mode):
    params = {k: (v.clone().cuda().requires_grad_(True) if v.is_floating_point() and "running" not in k else v.clone().cuda()) for k, v in sd.items()}
    ctx = nets.Ctx(training=True)
    if mode == "emu":
        with nets.bf16_emulation():
            out = nets.FORWARDS[name](params, x, ctx)
            out = out[-1] if isinstance(out, tuple) else out
            loss = nets.dice_loss(out, y) if out.shape[1] == 1 else torch.nn.functional.mse_loss(out, x)
            loss.backward()
    elif mode == "autocast":
        with torch.autocast("cuda", dtype=torch.bfloat16):
            out = nets.FORWARDS[name](params, x, ctx)
        out = (out[-1] if isinstance(out, tuple) else out).float()
        loss = nets.dice_loss(out, y) if out.shape[1] == 1 else torch.nn.functional.mse_loss(out, x)
        loss.backward()
    else:
        out = nets.FORWARDS[name](params, x, ctx)
        out = out[-1] if isinstance(out, tuple) else out
        loss = nets.dice_loss(out, y) if out.shape[1] == 1 else torch.nn.functional.mse_loss(out, x)
        loss.backward()
    return out.detach(), {k: p.grad for k, p in params.items() if p.requires_grad and p.grad is not None}

def run_product(name, sd, x, y):
    m = hybridnet.MODELS[name](); m.load_state_dict(sd); m = m.cuda().train()
    for mod in m.modules():
        if isinstance(mod, (torch.nn.Dropout, torch.nn.Dropout2d)): mod.p = 0.0
    out = m(x); out = out[-1] if isinstance(out, tuple) else out
    loss = nets.dice_loss(out, y) if out.shape[1] == 1 else torch.nn.functional.mse_loss(out, x)
    loss.backward()
    return out.detach(), {k: p.grad for k, p in m.named_parameters() if p.grad is not None}

for name, B, S, jit in [("unet", 1, 256, 0.1), ("unet", 4, 256, 0.1), ("unet", 2, 64, 0.2), ("ae32k_image", 2, 32, 0.2), ("ae32k_image", 2, 64, 0.2), ("ae32k_image", 4, 128, 0.2), ("fct", 1, 256, 0.1)]:
    sd = nets.init_state_dict(name, 7, jit)
    g = torch.Generator().manual_seed(2)
    x = torch.rand(B, 3, S, S, generator=g).cuda(); y = (torch.rand(B, 1, S, S, generator=g) > 0.5).float().cuda()
    ref_o, ref_g = run_oracle(name, sd, x, y, "fp32")
    first = [k for k in ref_g if k.endswith("weight")][0]
    for mode in ("product", "emu", "autocast"):
        o, gr = run_product(name, sd, x, y) if mode == "product" else run_oracle(name, sd, x, y, mode)
        ks = [k for k in ref_g if k in gr]
        ga = torch.cat([gr[k].flatten().float() for k in ks]); gb = torch.cat([ref_g[k].flatten() for k in ks])
        worst = min((cosine(gr[k].float(), ref_g[k]), k) for k in ks if float(ref_g[k].norm()) > 1e-4)
        print(f"{name:12s} B={B} S={S} {mode:9s} out rel={rel_l2(o, ref_o):.4f} maxabs={float((o.float()-ref_o).abs().max()):.4f} "
              f"grad cos={cosine(ga, gb):.5f} rel={rel_l2(ga, gb):.4f} first-layer cos={cosine(gr[first].float(), ref_g[first]):.4f} worst={worst[0]:.4f} {worst[1]}")
    if name != "fct":
        emu_o, emu_g = run_oracle(name, sd, x, y, "emu"); o, gr = run_product(name, sd, x, y)
        ks = [k for k in emu_g if k in gr]
        ga = torch.cat([gr[k].flatten().float() for k in ks]); gb = torch.cat([emu_g[k].flatten() for k in ks])
        print(f"{'':12s} product vs emu: out rel={rel_l2(o, emu_o):.4f} grad cos={cosine(ga, gb):.5f} first cos={cosine(gr[first].float(), emu_g[first]):.4f}")
