"""How far does the UNMODIFIED reference itself move when it runs in bf16?  (build container only)

    python oracle/make_bf16_yardstick.py      # writes tests/golden/bf16_yardstick.json

For every VAE fixture in tests/golden/ the reference model is rebuilt from the stored weights and
run under ``torch.autocast("cpu", dtype=torch.bfloat16)`` (what Lightning's ``precision=16/bf16``
does to it) on the stored inputs and reparameterisation noise; the deviation of its outputs and of
the whole gradient vector from the stored fp32 results is recorded.  tests/test_model_gpu.py uses
these numbers as the yardstick for the bf16 mode of the CUDA path on the same (very narrow)
models: it must stay within a small multiple of the reference's own bf16 error.
Test infrastructure; never imported by the product package."""
from __future__ import annotations

import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from make_golden import OUT, load_reference  # noqa: E402

FIXTURES = {
    "base_vae_28": ("BaseVAE", dict(input_channels=1, latent_dim=4, hidden_channels=8, num_res_blocks=1, dropout=0.0,
                                    ch_mult=(1, 2, 4), attn_resolutions=[14], resolution=28),
                    dict(recon_loss_type="mse", kl_weight=1.0, recon_weight=1.0), False),
    "beta_vae_28": ("BetaVAE", dict(beta=4.0, input_channels=3, latent_dim=4, hidden_channels=16, num_res_blocks=1,
                                    dropout=0.0, ch_mult=(1, 2), attn_resolutions=[], resolution=28),
                    dict(recon_loss_type="l1", kl_weight=4.0, recon_weight=1.0), False),
    "cvae_16": ("ConditionalVAE", dict(input_channels=3, latent_dim=4, hidden_channels=16, num_res_blocks=1, dropout=0.0,
                                       ch_mult=(1, 2), attn_resolutions=[8], resolution=16, condition_method="concat"),
                dict(recon_loss_type="bce", kl_weight=0.5, recon_weight=2.0), True),
    "cvae_32": ("ConditionalVAE", dict(input_channels=3, latent_dim=4, hidden_channels=8, num_res_blocks=2, dropout=0.0,
                                       ch_mult=(1, 2, 2, 4), attn_resolutions=[8, 4], resolution=32,
                                       condition_method="concat"),
                dict(recon_loss_type="mse", kl_weight=1.0, recon_weight=1.0), True),
}


def rel(a, b):
    a, b = a.double(), b.double()
    return float((a - b).norm() / (b.norm() + 1e-30))


def main():
    M, L = load_reference()
    res = {}
    for name, (cls, kw, lkw, conditional) in FIXTURES.items():
        z = np.load(os.path.join(OUT, name + ".npz"))
        t = lambda k: torch.from_numpy(np.asarray(z[k]))
        model = getattr(M, cls)(**kw).float()
        model.load_state_dict({k[3:]: t(k) for k in z.files if k.startswith("sd/")})
        eps = t("eps")
        orig = torch.randn_like
        torch.randn_like = lambda s, *a, **k: eps.to(s.dtype)
        try:
            with torch.autocast("cpu", dtype=torch.bfloat16):
                out = model(t("x"), t("cond")) if conditional else model(t("x"))
                ld = L.VAELoss(**lkw)(inputs=t("x"), reconstructions=out["reconstruction"].float(),
                                      posteriors=out["posterior"], priors=out["prior"])
        finally:
            torch.randn_like = orig
        ld["loss"].float().backward()
        num = den = 0.0
        worst = 0.0
        for n, p in model.named_parameters():
            g = t("grads/" + n).double()
            d = (p.grad.double() if p.grad is not None else torch.zeros_like(g)) - g
            num += float(d.pow(2).sum()); den += float(g.pow(2).sum())
            if float(g.norm()) > 1e-3 * (den ** 0.5 + 1e-30):
                worst = max(worst, float(d.norm() / g.norm()))
        res[name] = {
            "reconstruction": rel(out["reconstruction"].float(), t("out/reconstruction")),
            "mean": rel(out["mean"].float(), t("out/mean")),
            "logvar": rel(out["logvar"].float(), t("out/logvar")),
            "z": rel(out["z"].float(), t("out/z")),
            "whole_gradient": (num / den) ** 0.5,
            "worst_tensor": worst,
            "loss_abs": abs(float(ld["loss"]) - float(t("loss/loss"))),
        }
        print(name, res[name])
    with open(os.path.join(OUT, "bf16_yardstick.json"), "w") as f:
        json.dump({"how": "reference under torch.autocast('cpu', bfloat16) vs its own fp32 results (oracle/make_bf16_yardstick.py)",
                   "fixtures": res}, f, indent=1)


if __name__ == "__main__":
    main()
