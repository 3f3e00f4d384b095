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
