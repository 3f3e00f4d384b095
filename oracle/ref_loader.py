"""Load the reference's unmodified model classes from ``oracle/_ref/`` (see ``oracle/make_ref.py``).

TEST / BASELINE INFRASTRUCTURE ONLY: imported by ``bench.py``'s reference arm / ``cpu_baseline`` leg, the
torch-eager library bar and the tests -- never by the product package.
"""
import contextlib
import importlib
import io
import os
import sys
import types

HERE = os.path.dirname(os.path.abspath(__file__))
REF_ZIP = os.path.join(HERE, "_ref", "ref_models.zip")
_cache = {}


def available():
    return os.path.exists(REF_ZIP)


def _stub_missing_modules():
    """Modules the reference imports but which are neither installed nor on the hot path (FCT.py:4,15,16,21;
    Metrics.py:3)."""
    for name in ("torchsummary", "tensorboardX", "dataset", "metric", "pytorch_msssim"):
        if name not in sys.modules:
            m = types.ModuleType(name)
            m.summary = m.SummaryWriter = m.DataLoader = m.DiceLoss = m.JaccardScore = object
            m.ms_ssim = None
            sys.modules[name] = m


def load():
    """-> ({"fct": FCT, "unet": UNet, "autoencoder": AutoEncoder, "i2i2m": Image2Image2Mask}, DiceLoss)"""
    if _cache:
        return _cache["classes"], _cache["dice"]
    if not available():
        raise RuntimeError("oracle/_ref/ref_models.zip is missing: run `python oracle/make_ref.py` where /root/reference exists")
    _stub_missing_modules()
    sys.path.insert(0, REF_ZIP)                           # zipimport; AE_Unet_Combined.py does `from UNet import UNet`
    try:
        with contextlib.redirect_stdout(io.StringIO()):   # AutoEncoder.py prints a demo forward at import (:138-143)
            unet, ae, comb, metrics, fct = (importlib.import_module(n) for n in
                                            ("UNet", "AutoEncoder", "AE_Unet_Combined", "Metrics", "FCT"))
    finally:
        sys.path.remove(REF_ZIP)
    for m in (unet, ae, comb, metrics, fct):
        if REF_ZIP not in (getattr(m, "__file__", "") or ""):
            raise RuntimeError(f"oracle.ref_loader: module {m.__name__} came from {m.__file__}, not from {REF_ZIP}")
    _cache["classes"] = {"unet": unet.UNet, "autoencoder": ae.AutoEncoder, "i2i2m": comb.Image2Image2Mask, "fct": fct.FCT}
    _cache["dice"] = metrics.DiceLoss
    return _cache["classes"], _cache["dice"]
