"""Run a script of the reference (main.py, generate.py, quick_test.py ...) with `src.models` / `src.losses` served by
the B200 path and everything else (`src.lightning_module`, `src.data`, `src.utils`, configs) by the reference itself:

    cd /path/to/medvae-disentangled-multimodal
    python -m medvae_disentangled_multimodal_b200.launch main.py experiment=conditional_vae_64

The reference's `src` is a namespace package (no src/__init__.py), so putting `shim/` ahead of the script's directory
on sys.path makes `shim/src/{models,losses}` the first portion of `src` — no file of the reference is edited
(INTEGRATION.md 1b').  Hydra `_target_: src.models.ConditionalVAE` (configs/model/conditional_vae.yaml:2) and the
`isinstance` checks of VAELightningModule (src/lightning_module.py:91,116,119) then see the drop-in classes."""
from __future__ import annotations

import os
import runpy
import sys

SHIM = os.path.join(os.path.dirname(os.path.abspath(__file__)), "shim")


def install(script_dir: str | None = None) -> None:
    """put the shim first on sys.path (and the script's directory right behind it, where `python script.py` would
    have put it first)"""
    for p in (script_dir, SHIM):
        if p:
            while p in sys.path:
                sys.path.remove(p)
            sys.path.insert(0, p)
    if "src" in sys.modules or any(m.startswith("src.") for m in sys.modules):
        raise RuntimeError("`src` was imported before the shim was installed; call install() first")


def main(argv=None) -> None:
    argv = list(sys.argv[1:] if argv is None else argv)
    if not argv:
        raise SystemExit("usage: python -m medvae_disentangled_multimodal_b200.launch <reference script> [args...]")
    script = os.path.abspath(argv[0])
    install(os.path.dirname(script))
    sys.argv = [script] + argv[1:]
    runpy.run_path(script, run_name="__main__")


if __name__ == "__main__":
    main()
