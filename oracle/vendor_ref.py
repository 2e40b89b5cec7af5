"""TEST / BASELINE INFRASTRUCTURE ONLY — make the reference's own head modules available on the GPU box.

``/root/reference`` exists only in the build container.  ``python -m oracle.vendor_ref`` copies the four pure-torch files the
reference arm and the oracle pin need — UNMODIFIED, byte for byte — into ``oracle/_ref/`` (git-ignored, so nothing of the
reference enters the history; NOT gpurun-ignored, so the directory travels to the GPU box next to the built ``.so``):

    modules/layers.py  modules/miniViT.py  modules/ObjCAViT.py  losses/SILogLoss.py

``oracle/_ref/MANIFEST.json`` records the sha256 of every file.  ``oracle.ref_shim`` falls back to ``oracle/_ref`` when
``/root/reference`` is absent, so ``bench.py --impl reference`` times the REAL reference modules (``kind: "reference"``) instead
of the restatement.  ``__graft_entry__.build()`` runs this recipe whenever the reference tree is present.
"""
from __future__ import annotations

import hashlib
import json
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
DST = os.path.join(HERE, "_ref")
FILES = ("modules/layers.py", "modules/miniViT.py", "modules/ObjCAViT.py", "losses/SILogLoss.py")


def vendor(src: str = "/root/reference", dst: str = DST) -> str:
    if not os.path.isfile(os.path.join(src, FILES[0])):
        raise RuntimeError(f"reference tree not found at {src}")
    sums = {}
    for rel in FILES:
        out = os.path.join(dst, rel)
        os.makedirs(os.path.dirname(out), exist_ok=True)
        shutil.copyfile(os.path.join(src, rel), out)
        with open(out, "rb") as f:
            sums[rel] = hashlib.sha256(f.read()).hexdigest()
    with open(os.path.join(dst, "MANIFEST.json"), "w") as f:
        json.dump({"source": src, "sha256": sums}, f, indent=1, sort_keys=True)
    return dst


def verify(dst: str = DST) -> bool:
    """True when ``dst`` holds all files with the recorded checksums."""
    try:
        with open(os.path.join(dst, "MANIFEST.json")) as f:
            sums = json.load(f)["sha256"]
        for rel in FILES:
            with open(os.path.join(dst, rel), "rb") as f:
                if hashlib.sha256(f.read()).hexdigest() != sums[rel]:
                    return False
        return True
    except (OSError, KeyError, ValueError):
        return False


if __name__ == "__main__":
    print(vendor(*(sys.argv[1:2] or ["/root/reference"])))
