"""Recipe for ``oracle/_ref/``: the reference's OWN model sources, staged unmodified so that the timed CPU arm
(``bench.py --impl reference``, ``cpu_baseline.kind = "reference"``) and the checker can run the real classes on
the GPU box, where ``/root/reference`` does not exist.

TEST / BASELINE INFRASTRUCTURE ONLY.  ``oracle/_ref/`` is git-ignored (no reference source enters the history) but
travels with the repo snapshot to the GPU box like the built ``.so``.  Run where the reference checkout exists:

    python oracle/make_ref.py            # also called by __graft_entry__.build()

What is staged -- as ONE archive, ``oracle/_ref/ref_models.zip``, imported through zipimport (byte-for-byte members):
  UNet.py, AutoEncoder.py, AE_Unet_Combined.py, Metrics.py   whole files
  FCT.py                                                      lines 1-255 only: the model classes.  Lines 256-412
      are the trainer, which imports modules that do not exist as source (`dataset`, `metric`) and starts a
      70-epoch training run at import time (FCT.py:410-412).
The Autoencoder32K family ships as py3.8 bytecode only (``__pycache__/AE_32K.cpython-38.pyc``) and cannot be staged:
its CPU arm stays the oracle port (``kind = "port"``).
"""
import os
import sys
import zipfile

HERE = os.path.dirname(os.path.abspath(__file__))
DST = os.path.join(HERE, "_ref")
WHOLE = ("UNet.py", "AutoEncoder.py", "AE_Unet_Combined.py", "Metrics.py")


def main(ref=None):
    ref = ref or os.environ.get("HN_REFERENCE", "/root/reference")
    if not os.path.isdir(ref):
        print(f"oracle/make_ref.py: {ref} not present; keeping whatever oracle/_ref/ already holds")
        return False
    os.makedirs(DST, exist_ok=True)
    with open(os.path.join(ref, "FCT.py")) as src:
        head = "\n".join(src.read().split("\n")[:255]) + "\n"
    with zipfile.ZipFile(os.path.join(DST, "ref_models.zip"), "w", zipfile.ZIP_DEFLATED) as z:
        for f in WHOLE:
            z.write(os.path.join(ref, f), f)
        z.writestr("FCT.py", head)
    print(f"oracle/_ref/ref_models.zip: staged {', '.join(WHOLE)} and FCT.py[1:255] from {ref}")
    return True


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else None)
