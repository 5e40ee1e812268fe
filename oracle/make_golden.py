"""Generate golden vectors by running the UNMODIFIED reference (build container only).

    python oracle/make_golden.py            # writes tests/golden/*.npz

Imports ``src.models`` / ``src.losses.vae_losses`` from /root/reference (read-only), with
``lpips`` / ``open_clip`` stubbed in sys.modules exactly as SURVEY.md 8(c) describes (VAELoss
touches neither).  The reparameterisation noise is captured by patching ``torch.randn_like``
while ``reparameterize`` runs.  Test infrastructure; never imported by the product package.
"""
from __future__ import annotations

import os
import sys
import types

import numpy as np
import torch

REF = os.environ.get("MVAE_REFERENCE", "/root/reference")
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests", "golden")


def load_reference():
    if REF not in sys.path:
        sys.path.insert(0, REF)
    for name in ("lpips", "open_clip"):
        sys.modules.setdefault(name, types.ModuleType(name))
    import src.models as M  # noqa
    import importlib
    L = importlib.import_module("src.losses.vae_losses")
    return M, L


class CaptureEps:
    """Patch torch.randn_like so the reference's eps (base_vae.py:86) is recorded."""

    def __enter__(self):
        self.eps = None
        self._orig = torch.randn_like

        def patched(t, *a, **k):
            e = self._orig(t, *a, **k)
            self.eps = e.detach().clone()
            return e

        torch.randn_like = patched
        return self

    def __exit__(self, *exc):
        torch.randn_like = self._orig


def hook_taps(model, taps):
    from src.models.encoder_decoder import ResnetBlock, AttnBlock, Downsample, Upsample
    hs = []
    for name, mod in model.named_modules():
        # a handful of per-layer activations is enough to localise a mismatch; keeping every
        # block would make the fixtures tens of MB
        keep = isinstance(mod, (AttnBlock, Downsample, Upsample)) or name.endswith(
            ("conv_in", "conv_out", "mid.block_2", "block.0"))
        if keep and (name.startswith("encoder") or name.startswith("decoder")):
            hs.append(mod.register_forward_hook(
                lambda m, i, o, name=name: taps.__setitem__(name, o.detach().clone())))
    return hs


def save(name, **arrs):
    os.makedirs(OUT, exist_ok=True)
    flat = {}
    for k, v in arrs.items():
        if isinstance(v, dict):
            for kk, vv in v.items():
                flat[f"{k}/{kk}"] = vv.detach().cpu().numpy() if torch.is_tensor(vv) else np.asarray(vv)
        else:
            flat[k] = v.detach().cpu().numpy() if torch.is_tensor(v) else np.asarray(v)
    path = os.path.join(OUT, name + ".npz")
    np.savez_compressed(path, **flat)
    print(f"wrote {path}: {os.path.getsize(path)/1024:.1f} KiB, {len(flat)} arrays")


def run_vae(M, L, name, cls, kwargs, batch, loss_kwargs, conditional=False):
    torch.manual_seed(1234)
    model = getattr(M, cls)(**kwargs).float()
    c, r = kwargs.get("input_channels", 1), kwargs["resolution"]
    g = torch.Generator().manual_seed(42)
    x = torch.rand(batch, c, r, r, generator=g) * 2 - 1
    cond = None
    if conditional:
        idx = torch.arange(batch) % 12
        cond = torch.nn.functional.one_hot(idx, 12).float()
    taps = {}
    hooks = hook_taps(model, taps)
    with CaptureEps() as cap:
        out = model(x, cond) if conditional else model(x)
    for h in hooks:
        h.remove()
    crit = L.VAELoss(**loss_kwargs)
    ld = crit(inputs=x, reconstructions=out["reconstruction"], posteriors=out["posterior"],
              priors=out["prior"])
    ld["loss"].backward()
    grads = {n: (p.grad if p.grad is not None else torch.zeros_like(p)) for n, p in model.named_parameters()}
    save(name, x=x, eps=cap.eps, cond=cond if cond is not None else torch.zeros(0),
         sd=dict(model.state_dict()), grads=grads, taps=taps,
         out={k: out[k] for k in ("reconstruction", "mean", "logvar", "z")},
         loss={k: v for k, v in ld.items()})


def run_disentangled(M, name, kwargs, batch, idx, channels):
    torch.manual_seed(4321)
    model = M.DisentangledConditionalVAE(**kwargs).float()
    r = kwargs["resolution"]
    g = torch.Generator().manual_seed(7)
    x = torch.rand(batch, channels, r, r, generator=g) * 2 - 1
    idx = torch.tensor(idx, dtype=torch.long)
    if channels == 3:
        for i, m in enumerate(idx.tolist()):
            if m in (0, 3):
                x[i, 1:] = 0
    with CaptureEps() as cap:
        out = model(x, idx)
    crit = M.DisentangledVAELoss("mse", 1.0, 1.0, 0.1, 0.05)
    ld = crit(out, x)
    ld["loss"].backward()
    grads = {n: (p.grad if p.grad is not None else torch.zeros_like(p)) for n, p in model.named_parameters()}
    save(name, x=x, eps=cap.eps, idx=idx, sd=dict(model.state_dict()), grads=grads,
         out={k: out[k] for k in ("reconstruction", "mean", "logvar", "z", "separation_loss", "contrastive_loss")},
         loss={k: v for k, v in ld.items()})


def main():
    M, L = load_reference()
    tiny = dict(latent_dim=4, hidden_channels=16, num_res_blocks=1, dropout=0.0)
    # BaseVAE, 1x28x28, 3 levels, attention at 14 (and always in mid)  -> C1-shaped, shrunk
    run_vae(M, L, "base_vae_28", "BaseVAE",
            dict(input_channels=1, ch_mult=(1, 2, 4), attn_resolutions=[14], resolution=28,
                 **{**tiny, "hidden_channels": 8}),
            batch=3, loss_kwargs=dict(recon_loss_type="mse", kl_weight=1.0, recon_weight=1.0))
    # BetaVAE beta=4 through VAELoss(kl_weight=4), 3x28x28, L1 loss  -> C2-shaped
    run_vae(M, L, "beta_vae_28", "BetaVAE",
            dict(beta=4.0, input_channels=3, ch_mult=(1, 2), attn_resolutions=[], resolution=28, **tiny),
            batch=2, loss_kwargs=dict(recon_loss_type="l1", kl_weight=4.0, recon_weight=1.0))
    # ConditionalVAE concat, 3x16x16, attention at 8, BCE loss  -> C3/C4-shaped
    run_vae(M, L, "cvae_16", "ConditionalVAE",
            dict(input_channels=3, ch_mult=(1, 2), attn_resolutions=[8], resolution=16,
                 condition_method="concat", **tiny),
            batch=4, loss_kwargs=dict(recon_loss_type="bce", kl_weight=0.5, recon_weight=2.0),
            conditional=True)
    # ConditionalVAE with two res blocks per level and 4 levels at 32x32 (C4 topology, shrunk)
    run_vae(M, L, "cvae_32", "ConditionalVAE",
            dict(input_channels=3, latent_dim=4, hidden_channels=8, num_res_blocks=2, dropout=0.0,
                 ch_mult=(1, 2, 2, 4), attn_resolutions=[8, 4], resolution=32, condition_method="concat"),
            batch=2, loss_kwargs=dict(recon_loss_type="mse", kl_weight=1.0, recon_weight=1.0),
            conditional=True)
    # Disentangled CVAE: the reference's own test shapes (test_modality_channels.py:50-112)
    run_disentangled(M, "disentangled_28_mixed",
                     dict(num_modalities=5, shared_latent_dim=4, modality_latent_dim=4, resolution=28,
                          hidden_channels=16, ch_mult=(1, 2), num_res_blocks=1, attn_resolutions=[]),
                     batch=6, idx=[0, 1, 3, 4, 1, 0], channels=3)
    run_disentangled(M, "disentangled_28_gray",
                     dict(num_modalities=5, shared_latent_dim=4, modality_latent_dim=4, resolution=28,
                          hidden_channels=16, ch_mult=(1, 2), num_res_blocks=1, attn_resolutions=[]),
                     batch=2, idx=[0, 3], channels=1)


if __name__ == "__main__":
    main()
