"""Recipe: place the UNMODIFIED reference hot path under ``oracle/_ref/`` (build container only).

    python oracle/build_ref.py        # also run by __graft_entry__.build() when /root/reference exists

The reference is pure Python (no compiled code on this path), so "building" it means copying, byte for
byte, the files the path needs from where they lie under /root/reference:

    src/models/*.py                  (Encoder / Decoder / BaseVAE / BetaVAE / ConditionalVAE / Disentangled...)
    src/losses/{__init__,vae_losses}.py   (VAELoss; imports lpips / open_clip at module top — stubbed by ref_runner)

``oracle/_ref/`` is git-ignored (nothing of the reference enters the history) but NOT gpurun-ignored, so the
copy travels to the GPU box where /root/reference does not exist.  ``pip install --target baseline/_ref`` of the
reference was tried first and fails offline (build backend ``hatchling`` is not in the wheelhouse; the
dependencies lightning / lpips / open_clip / medmnist / hydra / torchmetrics are absent too).  A MANIFEST with the
sha256 of every copied file is written next to the copy; ``oracle/ref_runner.py`` refuses to run if a file no longer
matches it.  TEST INFRASTRUCTURE: used as the checker (tests) and as the timed baselines of bench.py
(``cpu_baseline.kind = "reference"``, ``gpu_baseline``), never by the product package.
"""
from __future__ import annotations

import hashlib
import json
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("MVAE_REFERENCE", "/root/reference")
DST = os.path.join(HERE, "_ref")
FILES = [
    "src/models/__init__.py", "src/models/base_vae.py", "src/models/beta_vae.py", "src/models/conditional_vae.py",
    "src/models/discriminator.py", "src/models/disentangled_conditional_vae.py", "src/models/encoder_decoder.py",
    "src/losses/__init__.py", "src/losses/vae_losses.py",
]


def sha256(path: str) -> str:
    h = hashlib.sha256()
    with open(path, "rb") as f:
        h.update(f.read())
    return h.hexdigest()


def build_ref(verbose: bool = True) -> str | None:
    """-> path of the copy, or None when the reference tree is not present (GPU box: uses the shipped copy)."""
    if not os.path.isdir(os.path.join(REF, "src", "models")):
        return DST if os.path.exists(os.path.join(DST, "MANIFEST.json")) else None
    manifest = {"source": REF, "files": {}}
    for rel in FILES:
        src, dst = os.path.join(REF, rel), os.path.join(DST, rel)
        os.makedirs(os.path.dirname(dst), exist_ok=True)
        shutil.copyfile(src, dst)
        manifest["files"][rel] = sha256(dst)
    try:
        import subprocess
        manifest["commit"] = subprocess.run(["git", "-C", REF, "rev-parse", "HEAD"], capture_output=True, text=True,
                                            timeout=10).stdout.strip() or None
    except Exception:
        manifest["commit"] = None
    with open(os.path.join(DST, "MANIFEST.json"), "w") as f:
        json.dump(manifest, f, indent=1)
    if verbose:
        print(f"oracle/_ref: {len(FILES)} reference files copied from {REF}")
    return DST


if __name__ == "__main__":
    p = build_ref()
    sys.exit(0 if p else 1)
