"""CPU-side checks of the drop-in boundary: the C-ABI library loads, exports every symbol
include/hybridnet.h declares, and the product modules keep the reference's state_dict layout.
No compute call is made (there is no GPU here)."""
import ctypes
import json
import os
import re

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _header_symbols():
    src = open(os.path.join(ROOT, "include", "hybridnet.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(hn_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    import hybridnet
    from hybridnet import _lib
    assert os.path.exists(_lib.LIB_PATH), "run __graft_entry__.build() first"
    lib = ctypes.CDLL(_lib.LIB_PATH)
    syms = _header_symbols()
    assert len(syms) >= 35
    for s in syms:
        assert hasattr(lib, s), f"{s} declared in include/hybridnet.h but not exported"
    assert sorted(_lib.exported_symbols()) == syms, "ctypes table and header disagree"


def test_state_dict_layout_matches_reference_classes():
    import hybridnet
    keys = json.load(open(os.path.join(ROOT, "tests", "golden", "state_dict_keys.json")))
    for name, ref in keys.items():
        m = hybridnet.MODELS[name]()
        sd = m.state_dict()
        assert list(sd.keys()) == list(ref.keys()), name
        for k, v in sd.items():
            assert list(v.shape) == ref[k], (name, k)


def test_checkpoint_roundtrip_reference_format(tmp_path):
    """FCT.py:366-373 format: {'epoch','model_state_dict','optimizer_state_dict','loss'}; the oracle's
    state_dict (reference key names) must load strictly into the product module and back."""
    import hybridnet
    from oracle import nets
    sd = nets.init_state_dict("fct", 3)
    m = hybridnet.FCT()
    m.load_state_dict(sd, strict=True)
    opt = torch.optim.AdamW(m.parameters(), lr=1e-3)
    path = tmp_path / "FCT_for_cars.tar"
    torch.save({"epoch": 1, "model_state_dict": m.state_dict(), "optimizer_state_dict": opt.state_dict(),
                "loss": 0.5}, path)
    ck = torch.load(path)
    m2 = hybridnet.FCT()
    m2.load_state_dict(ck["model_state_dict"], strict=True)
    for k, v in sd.items():
        assert torch.equal(m2.state_dict()[k], v), k


def test_missing_library_fails_loudly(monkeypatch):
    from hybridnet import _lib
    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setattr(_lib, "LIB_PATH", "/nonexistent/libhybridnet_sm100.so")
    with pytest.raises(RuntimeError, match="no CPU/PyTorch fallback"):
        _lib.load()


def test_product_never_imports_oracle():
    pkg = os.path.join(ROOT, "transformer-cnn-hybrid-network-for-video-processing_b200")
    for fn in os.listdir(pkg):
        if fn.endswith(".py"):
            src = open(os.path.join(pkg, fn)).read()
            assert not re.search(r"^\s*(from|import)\s+oracle", src, flags=re.M), fn
            assert "sys.path" not in src, fn


def test_default_init_matches_reference_when_available():
    ref = os.environ.get("HN_REFERENCE", "/root/reference")
    if not os.path.isdir(ref):
        pytest.skip("reference checkout not present (GPU box)")
    import sys
    sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
    import make_golden
    import hybridnet
    classes, _ = make_golden.load_reference()
    for name in ("fct", "unet", "autoencoder", "i2i2m"):
        torch.manual_seed(0)
        a = classes[name]().state_dict()
        torch.manual_seed(0)
        b = hybridnet.MODELS[name]().state_dict()
        assert all(torch.equal(a[k], b[k]) for k in a), name


def test_fused_adamw_state_dict_layout_is_torchs():
    """hybridnet.optim.AdamW speaks torch.optim.AdamW's state_dict format (the reference stores it in its .tar
    checkpoints, FCT.py:371).  Host-side logic only: no kernel runs here."""
    import hybridnet
    from hybridnet import parallel
    m = hybridnet.FCT()
    dead = parallel.used_parameter_names("fct")
    arena = parallel.GradArena(m, used={n for n, _ in m.named_parameters() if n not in dead})
    before = {n: p.detach().clone() for n, p in m.named_parameters()}
    opt = hybridnet.optim.AdamW(m.parameters(), arena, lr=1e-3)
    for n, p in m.named_parameters():                         # re-pointing into the flat buffer changes nothing
        assert torch.equal(p.detach(), before[n]), n
    ref_params = [torch.nn.Parameter(p.detach().clone()) for p in m.parameters()]
    ref = torch.optim.AdamW(ref_params, lr=1e-3)
    names = [n for n, _ in m.named_parameters()]
    for n, r in zip(names, ref_params):                        # one reference step; never-used tensors keep grad None
        if n not in dead:
            r.grad = torch.full_like(r, 0.01)
    ref.step()
    opt.load_state_dict(ref.state_dict())
    back = opt.state_dict()
    tsd = ref.state_dict()
    assert back["param_groups"] == tsd["param_groups"]
    assert set(back["state"]) == set(tsd["state"]) and len(back["state"]) == len(names) - len(dead)
    for i, st in tsd["state"].items():
        assert float(back["state"][i]["step"]) == float(st["step"]) == 1.0
        assert torch.equal(back["state"][i]["exp_avg"], st["exp_avg"])
        assert torch.equal(back["state"][i]["exp_avg_sq"], st["exp_avg_sq"])
    fresh = torch.optim.AdamW([torch.nn.Parameter(p.detach().clone()) for p in m.parameters()], lr=1e-3)
    fresh.load_state_dict(back)                               # and torch accepts ours


def test_checkpoint_crossload_with_reference_classes(tmp_path):
    """a `.tar` saved from the REFERENCE classes (FCT.py:366-373 format, optimizer state included)
    loads strictly into the product classes and optimizer, and what the product saves loads strictly back into the
    reference classes / torch.optim.AdamW with identical tensors."""
    ref = os.environ.get("HN_REFERENCE", "/root/reference")
    if not os.path.isdir(ref):
        pytest.skip("reference checkout not present (GPU box)")
    import sys
    sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
    import make_golden
    import hybridnet
    from hybridnet import parallel
    classes, _ = make_golden.load_reference()
    for name in ("fct", "unet", "autoencoder", "i2i2m"):
        torch.manual_seed(1)
        rm = classes[name]()
        ropt = torch.optim.AdamW(rm.parameters(), lr=1e-3)
        x = torch.rand(1, 3, 32, 32)
        out = rm(x)
        out = out[-1] if isinstance(out, tuple) else out
        out.mean().backward()
        ropt.step()                                            # real optimizer state, real BN statistics
        path = tmp_path / f"{name}_ref.tar"
        torch.save({"epoch": 3, "model_state_dict": rm.state_dict(), "optimizer_state_dict": ropt.state_dict(),
                    "loss": 0.25}, path)
        ck = torch.load(path)
        pm = hybridnet.MODELS[name]()
        pm.load_state_dict(ck["model_state_dict"], strict=True)
        dead = parallel.used_parameter_names(name)
        arena = parallel.GradArena(pm, used={n for n, _ in pm.named_parameters() if n not in dead})
        popt = hybridnet.optim.AdamW(pm.parameters(), arena, lr=1e-3)
        popt.load_state_dict(ck["optimizer_state_dict"])
        path2 = tmp_path / f"{name}_product.tar"
        torch.save({"epoch": 4, "model_state_dict": pm.state_dict(), "optimizer_state_dict": popt.state_dict(),
                    "loss": 0.2}, path2)
        ck2 = torch.load(path2)
        rm2 = classes[name]()
        rm2.load_state_dict(ck2["model_state_dict"], strict=True)
        for k, v in rm.state_dict().items():
            assert torch.equal(rm2.state_dict()[k], v), (name, k)
        ropt2 = torch.optim.AdamW(rm2.parameters(), lr=1e-3)
        ropt2.load_state_dict(ck2["optimizer_state_dict"])
        a, b = ropt.state_dict(), ropt2.state_dict()
        assert a["param_groups"] == b["param_groups"] and set(a["state"]) == set(b["state"]), name
        for i in a["state"]:
            for f in ("step", "exp_avg", "exp_avg_sq"):
                assert torch.equal(torch.as_tensor(a["state"][i][f]), torch.as_tensor(b["state"][i][f])), (name, i, f)
