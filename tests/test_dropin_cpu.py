"""CPU-side checks of the drop-in class surface: names, ctor kwargs, state_dict
keys/shapes against the reference's golden state_dicts, seeded-init equality with the reference
when it is importable (build container only)."""
import os
import sys
import types

import pytest
import torch

from conftest import load_golden

CASES = {
    "base_vae_28": ("BaseVAE", dict(input_channels=1, latent_dim=4, hidden_channels=8, num_res_blocks=1,
                                    ch_mult=(1, 2, 4), attn_resolutions=[14], resolution=28)),
    "beta_vae_28": ("BetaVAE", dict(beta=4.0, input_channels=3, latent_dim=4, hidden_channels=16, num_res_blocks=1,
                                    ch_mult=(1, 2), attn_resolutions=[], resolution=28)),
    "cvae_16": ("ConditionalVAE", dict(input_channels=3, latent_dim=4, hidden_channels=16, num_res_blocks=1,
                                       ch_mult=(1, 2), attn_resolutions=[8], resolution=16)),
    "cvae_32": ("ConditionalVAE", dict(input_channels=3, latent_dim=4, hidden_channels=8, num_res_blocks=2,
                                       ch_mult=(1, 2, 2, 4), attn_resolutions=[8, 4], resolution=32)),
}


@pytest.mark.parametrize("name", list(CASES))
def test_state_dict_is_reference_compatible(name):
    import medvae_disentangled_multimodal_b200 as M
    cls, kw = CASES[name]
    g = load_golden(name)
    model = getattr(M, cls)(**kw)
    sd = model.state_dict()
    assert set(sd) == set(g["sd"]), set(sd) ^ set(g["sd"])
    for k, v in g["sd"].items():
        assert tuple(sd[k].shape) == tuple(v.shape), k
    model.load_state_dict(g["sd"])      # strict


def test_public_names_and_kwargs():
    import medvae_disentangled_multimodal_b200 as M
    for n in ("BaseVAE", "BetaVAE", "ConditionalVAE", "Encoder", "Decoder", "VAELoss"):
        assert hasattr(M, n)
    m = M.ConditionalVAE(input_channels=3, latent_dim=4, hidden_channels=8, ch_mult=(1, 2), num_res_blocks=1,
                         attn_resolutions=[], resolution=16, num_modalities=12, some_unknown_kwarg=1)
    assert m.num_modalities == 12 and m.condition_dim == 12 and m.condition_method == "concat"
    assert m.encoder.conv_in.in_channels == 6 and m.latent_dim == 4 and m.encoder_out_res == 8
    assert torch.equal(m.get_modality_condition("oct"), torch.eye(12)[2])
    with pytest.raises(ValueError):
        m.get_modality_condition("nope")
    assert M.BetaVAE(beta=4.0, input_channels=1, latent_dim=4, hidden_channels=8, ch_mult=(1, 2),
                     num_res_blocks=1, attn_resolutions=[], resolution=16).beta == 4.0
    assert M.VAELoss("l1", 2.0, 3.0).kl_weight == 2.0
    with pytest.raises(NotImplementedError):
        M.BaseVAE(attn_type="bogus", resolution=16, hidden_channels=8, ch_mult=(1,), num_res_blocks=1)


@pytest.mark.skipif(not os.path.isdir("/root/reference/src"), reason="reference only exists in the build container")
def test_seeded_init_equals_reference():
    sys.path.insert(0, "/root/reference")
    for n in ("lpips", "open_clip"):
        sys.modules.setdefault(n, types.ModuleType(n))
    import src.models as R
    import medvae_disentangled_multimodal_b200 as M
    kw = dict(input_channels=3, latent_dim=8, hidden_channels=16, ch_mult=(1, 2, 4), num_res_blocks=2,
              attn_resolutions=[14, 7], resolution=28)
    for cls, extra in (("BaseVAE", {}), ("BetaVAE", {"beta": 2.0}), ("ConditionalVAE", {}),
                       ("ConditionalVAE", {"condition_method": "film"})):
        torch.manual_seed(3)
        a = getattr(M, cls)(**kw, **extra).state_dict()
        torch.manual_seed(3)
        b = getattr(R, cls)(**kw, **extra).state_dict()
        assert list(a) == list(b)
        assert all(torch.equal(a[k], b[k]) for k in a), cls


DIS_KW = dict(num_modalities=5, shared_latent_dim=4, modality_latent_dim=4, resolution=28, hidden_channels=16,
              ch_mult=(1, 2), num_res_blocks=1, attn_resolutions=[])


@pytest.mark.parametrize("name", ["disentangled_28_mixed", "disentangled_28_gray"])
def test_disentangled_state_dict_is_reference_compatible(name):
    import medvae_disentangled_multimodal_b200 as M
    g = load_golden(name)
    model = M.DisentangledConditionalVAE(**DIS_KW)
    sd = model.state_dict()
    assert set(sd) == set(g["sd"]), set(sd) ^ set(g["sd"])
    for k, v in g["sd"].items():
        assert tuple(sd[k].shape) == tuple(v.shape), k
    model.load_state_dict(g["sd"])
    assert model.latent_dim == 8 and model.input_channels == 3 and model.num_modalities == 5
    assert model.modality_channels == {0: 1, 1: 3, 2: 3, 3: 1, 4: 3}
    assert sorted(model.modality_input_projectors) == ["0", "3"] == sorted(model.modality_output_projectors)
    crit = M.DisentangledVAELoss("l1", 2.0, 1.0, 0.3, 0.4)
    assert crit.separation_weight == 0.3 and crit.contrastive_weight == 0.4


@pytest.mark.skipif(not os.path.isdir("/root/reference/src"), reason="reference only exists in the build container")
def test_disentangled_seeded_init_equals_reference():
    sys.path.insert(0, "/root/reference")
    for n in ("lpips", "open_clip"):
        sys.modules.setdefault(n, types.ModuleType(n))
    import src.models as R
    import medvae_disentangled_multimodal_b200 as M
    torch.manual_seed(5)
    a = M.DisentangledConditionalVAE(**DIS_KW).state_dict()
    torch.manual_seed(5)
    b = R.DisentangledConditionalVAE(**DIS_KW).state_dict()
    assert list(a) == list(b)
    assert all(torch.equal(a[k], b[k]) for k in a)
