"""The zero-edit drop-in route (INTEGRATION.md 1b'): `python -m medvae_disentangled_multimodal_b200.launch script.py`
makes `src.models` / `src.losses` resolve to the B200 classes while every other `src.*` module still comes from the
reference tree — the reference's `src` is a namespace package (no src/__init__.py).  Run in subprocesses: other tests
import the real `src` from /root/reference in this process."""
import os
import subprocess
import sys
import textwrap

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PROBE = textwrap.dedent("""
    import importlib.util, sys
    from src.models import (BaseVAE, BetaVAE, ConditionalVAE, DisentangledConditionalVAE, DisentangledVAELoss, Encoder,
                            Decoder, NLayerDiscriminator)
    from src.models.disentangled_conditional_vae import DisentangledVAELoss as D2
    from src.models.encoder_decoder import ResnetBlock
    from src.losses import VAELoss, LPIPSLoss, BiomedCLIPLoss, LPIPSWithDiscriminator
    from src.losses.vae_losses import VAELoss as V2
    assert D2 is DisentangledVAELoss and V2 is VAELoss
    for c in (BaseVAE, BetaVAE, ConditionalVAE, DisentangledConditionalVAE, Encoder, VAELoss, ResnetBlock):
        assert c.__module__.startswith("medvae_disentangled_multimodal_b200."), c
    assert issubclass(BetaVAE, BaseVAE) and issubclass(ConditionalVAE, BaseVAE)
    m = ConditionalVAE(input_channels=3, latent_dim=4, hidden_channels=8, ch_mult=(1, 2), num_res_blocks=1,
                       attn_resolutions=[], resolution=16)
    assert m.encoder.conv_in.weight.shape[1] == 6          # conv_in rebuilt for the concatenated condition map
    print("OTHER", importlib.util.find_spec(sys.argv[1]).origin)
""")


def _launch(cwd, script, *args):
    env = dict(os.environ, PYTHONPATH=ROOT + os.pathsep + os.environ.get("PYTHONPATH", ""))
    return subprocess.run([sys.executable, "-m", "medvae_disentangled_multimodal_b200.launch", script, *args], cwd=cwd,
                          env=env, capture_output=True, text=True, timeout=300)


def test_shim_serves_models_and_losses_and_leaves_the_rest(tmp_path):
    ref = tmp_path / "ref"
    (ref / "src" / "models").mkdir(parents=True)
    (ref / "src" / "data").mkdir(parents=True)
    (ref / "src" / "models" / "__init__.py").write_text("raise ImportError('the reference models were imported')\n")
    (ref / "src" / "data" / "__init__.py").write_text("MARK = 'reference data'\n")
    (ref / "probe.py").write_text(PROBE)
    r = _launch(str(ref), "probe.py", "src.data")
    assert r.returncode == 0, r.stderr[-3000:]
    assert f"OTHER {ref / 'src' / 'data' / '__init__.py'}" in r.stdout


@pytest.mark.skipif(not os.path.isdir("/root/reference/src"), reason="the reference tree only exists in the build container")
def test_shim_in_front_of_the_real_reference(tmp_path):
    probe = tmp_path / "probe.py"
    probe.write_text(PROBE)
    # the script sits outside the reference root here, so put that root on the path the way `cd reference` would
    env_probe = "import sys; sys.path.insert(1, '/root/reference')\n" + PROBE
    probe.write_text(env_probe)
    r = _launch("/root/reference", str(probe), "src.lightning_module")
    assert r.returncode == 0, r.stderr[-3000:]
    assert "OTHER /root/reference/src/lightning_module.py" in r.stdout
