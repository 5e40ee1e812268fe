"""Golden vectors for the rows either side of the training step, made by
running the UNMODIFIED reference in the build container:

    python oracle/make_golden_widen.py        # writes tests/golden/{linattn_vae_16,dropout_vae_16,film,metrics,feed}.npz

* linattn_vae_16  — BaseVAE(attn_type="linear") forward + VAELoss + backward (LinAttnBlock, encoder_decoder.py:36-65)
* dropout_vae_16  — BaseVAE(dropout=0.25) in training mode; the keep masks nn.Dropout drew are captured with
                    forward hooks and stored, so another implementation can be checked on identical masks
* film            — FiLMLayer forward/backward (conditional_vae.py:191-203)
* metrics         — compute_kl_metrics / compute_latent_metrics of src/utils/metrics.py (imported by file path with
                    ``torchmetrics.functional`` stubbed by the oracle's restatement: the reference's PSNR/SSIM numbers
                    in the fixture therefore come from the restatement and pin nothing — see oracle/metrics_oracle.py)
* feed            — MedMNISTDataset.__getitem__ + Normalize + mixed_modality_collate_fn (src/data/medmnist_data.py)
                    with the absent ``medmnist`` / ``lightning`` packages stubbed: a stub data class hands the
                    dataset PIL images built from stored uint8 arrays, everything after that is reference code

Test infrastructure; never imported by the product package.
"""
from __future__ import annotations

import importlib.util
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, ".."))
from oracle.make_golden import REF, CaptureEps, load_reference, save  # noqa: E402
from oracle import metrics_oracle as MO  # noqa: E402


def load_by_path(name, rel):
    spec = importlib.util.spec_from_file_location(name, os.path.join(REF, rel))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def run_vae(M, L, name, kwargs, batch, train_mode=False):
    torch.manual_seed(77)
    model = M.BaseVAE(**kwargs).float()
    model.train(train_mode)
    c, r = kwargs["input_channels"], kwargs["resolution"]
    x = torch.rand(batch, c, r, r, generator=torch.Generator().manual_seed(5)) * 2 - 1
    masks, hooks = {}, []
    if train_mode:
        for mname, mod in model.named_modules():
            if isinstance(mod, torch.nn.Dropout):
                key = mname[: -len(".dropout")]
                hooks.append(mod.register_forward_hook(
                    lambda m, i, o, key=key: masks.__setitem__(key, (o != 0).to(torch.uint8))))
    with CaptureEps() as cap:
        out = model(x)
    for h in hooks:
        h.remove()
    ld = L.VAELoss("mse", 1.0, 1.0)(inputs=x, reconstructions=out["reconstruction"], posteriors=out["posterior"],
                                   priors=out["prior"])
    ld["loss"].backward()
    grads = {n: (p.grad if p.grad is not None else torch.zeros_like(p)) for n, p in model.named_parameters()}
    save(name, x=x, eps=cap.eps, sd=dict(model.state_dict()), grads=grads, masks=masks,
         out={k: out[k] for k in ("reconstruction", "mean", "logvar", "z")}, loss=dict(ld))


def run_film():
    cv = sys.modules["src.models.conditional_vae"]
    torch.manual_seed(3)
    layer = cv.FiLMLayer(12, 24)
    f = torch.randn(5, 24, 7, 9, requires_grad=True)
    c = torch.randn(5, 12, requires_grad=True)
    out = layer(f, c)
    w = torch.randn(out.shape, generator=torch.Generator().manual_seed(1))
    (out * w).sum().backward()
    save("film", features=f, cond=c, w=w, out=out, sd=dict(layer.state_dict()), d_features=f.grad, d_cond=c.grad,
         grads={n: p.grad for n, p in layer.named_parameters()})


def run_metrics():
    tm = types.ModuleType("torchmetrics")
    tmf = types.ModuleType("torchmetrics.functional")
    tmf.structural_similarity_index_measure = lambda p, t, data_range=1.0: MO.ssim(p, t, data_range)
    tmf.peak_signal_noise_ratio = lambda p, t, data_range=1.0: MO.psnr(p, t, data_range)
    tm.functional = tmf
    sys.modules.setdefault("torchmetrics", tm)
    sys.modules.setdefault("torchmetrics.functional", tmf)
    RM = load_by_path("ref_metrics", "src/utils/metrics.py")
    g = torch.Generator().manual_seed(11)
    x = torch.rand(6, 3, 28, 28, generator=g) * 2 - 1
    r = (x + 0.2 * torch.randn(6, 3, 28, 28, generator=g)).clamp(-1.2, 1.2)
    mean = 0.7 * torch.randn(6, 8, 7, 7, generator=g)
    logvar = 0.5 * torch.randn(6, 8, 7, 7, generator=g) - 0.3
    z = torch.randn(6, 8, 7, 7, generator=g) * 0.3
    mean2, logvar2 = torch.randn(9, 16, generator=g), 0.3 * torch.randn(9, 16, generator=g)
    save("metrics", x=x, r=r, mean=mean, logvar=logvar, z=z, mean2=mean2, logvar2=logvar2,
         recon={k: torch.tensor(v) for k, v in RM.compute_reconstruction_metrics(x, r).items()},
         kl={k: torch.tensor(v) for k, v in RM.compute_kl_metrics(mean, logvar).items()},
         kl2={k: torch.tensor(v) for k, v in RM.compute_kl_metrics(mean2, logvar2).items()},
         latent={k: torch.tensor(v) for k, v in RM.compute_latent_metrics(z).items()})


def run_feed():
    from PIL import Image
    g = np.random.default_rng(2)
    n_per, res = 3, 28
    names = ["chestmnist", "pathmnist", "octmnist", "pneumoniamnist", "dermamnist", "bloodmnist", "tissuemnist",
             "retinamnist", "breastmnist", "organamnist", "organcmnist", "organsmnist"]
    # natural channel counts of the stored arrays (MedMNIST INFO n_channels): colour for path/derma/blood/retina
    natural = {n: (3 if n in ("pathmnist", "dermamnist", "bloodmnist", "retinamnist") else 1) for n in names}
    store = {n: g.integers(0, 256, size=(n_per, res, res, natural[n]) if natural[n] == 3 else (n_per, res, res),
                           dtype=np.uint8) for n in names}
    labels = {n: g.integers(0, 5, size=(n_per, 1)) for n in names}

    class _Data:
        def __init__(self, name, **kw):
            self.name = name

        def __len__(self):
            return n_per

        def __getitem__(self, i):
            return Image.fromarray(store[self.name][i]), labels[self.name][i]

    med = types.ModuleType("medmnist")
    med.INFO = {n: {"n_channels": natural[n], "label": {"0": "a", "1": "b"}, "task": "multi-class",
                    "python_class": "C_" + n} for n in names}
    for n in names:
        setattr(med, "C_" + n, (lambda n: (lambda **kw: _Data(n, **kw)))(n))
    lt = types.ModuleType("lightning")
    lt.LightningDataModule = object
    sys.modules["medmnist"] = med
    sys.modules["lightning"] = lt
    RD = load_by_path("ref_data", "src/data/medmnist_data.py")
    from torchvision import transforms
    samples = []
    for n in names:
        probe = RD.MedMNISTDataset(n, transform=None, download=False, size=res)
        norm = transforms.Normalize([0.5] * probe.target_channels, [0.5] * probe.target_channels)
        ds = RD.MedMNISTDataset(n, transform=transforms.Compose([norm]), download=False, size=res)
        samples += [ds[i] for i in range(n_per)]
    # three batches: mixed (everything), all-gray, all-colour
    order = list(range(len(samples)))
    gray_only = [i for i in order if samples[i][0].shape[0] == 1]
    rgb_only = [i for i in order if samples[i][0].shape[0] == 3]
    out = {}
    for tag, sel in (("mixed", order), ("gray", gray_only), ("rgb", rgb_only)):
        x, lab, onehot, idx = RD.mixed_modality_collate_fn([samples[i] for i in sel])
        out.update({f"{tag}_sel": torch.tensor(sel), f"{tag}_x": x, f"{tag}_labels": lab, f"{tag}_onehot": onehot,
                    f"{tag}_idx": idx})
    # staged uint8 store in sample order, 3 wide (gray in channel 0), + per-sample source channel count
    staged = np.zeros((len(samples), res, res, 3), dtype=np.uint8)
    src_c = []
    k = 0
    for n in names:
        for i in range(n_per):
            a = store[n][i]
            if a.ndim == 2:
                staged[k, :, :, 0] = a
            else:
                staged[k] = a
            src_c.append(natural[n])
            k += 1
    save("feed", staged=torch.from_numpy(staged), src_c=torch.tensor(src_c, dtype=torch.int32),
         modality_idx=torch.tensor([names.index(n) for n in names for _ in range(n_per)]), **out)


def main():
    M, L = load_reference()
    tiny = dict(input_channels=1, latent_dim=4, hidden_channels=8, num_res_blocks=1, ch_mult=(1, 2), resolution=16)
    run_vae(M, L, "linattn_vae_16", dict(attn_resolutions=[8], attn_type="linear", dropout=0.0, **tiny), batch=3)
    run_vae(M, L, "dropout_vae_16", dict(attn_resolutions=[8], dropout=0.25, **tiny), batch=3, train_mode=True)
    run_film()
    run_metrics()
    run_feed()


if __name__ == "__main__":
    main()
