"""What the cross-compile produced, checked without a GPU: the hot kernels really are tcgen05 / TMEM / TMA code (the
SASS mnemonics B200_PROFILING.md lists: tcgen05.mma -> UTCHMMA, tcgen05.ld -> LDTM, cp.async.bulk.tensor -> UTMALDG),
the CTA-pair variants exist (.2CTA), everything is built for sm_100a only, and no tcgen05 kernel spills more than a
few words."""
import os
import re
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OBJ = os.path.join(ROOT, "medvae_disentangled_multimodal_b200", "csrc", "_obj")


@pytest.fixture(scope="module")
def built():
    from medvae_disentangled_multimodal_b200.build import build
    build()
    cuobjdump = shutil.which("cuobjdump") or "/usr/local/cuda/bin/cuobjdump"
    if not os.path.exists(cuobjdump):
        pytest.skip("cuobjdump not available")
    return cuobjdump


def _sass(cuobjdump, name):
    return subprocess.run([cuobjdump, "-sass", os.path.join(OBJ, name + ".o")], capture_output=True, text=True,
                          check=True).stdout


def test_conv_kernels_are_tcgen05_tma_code(built):
    sass = _sass(built, "conv_umma")
    assert set(re.findall(r"arch = (sm_\w+)", sass)) == {"sm_100a"}
    count = lambda pat: len(re.findall(pat, sass))
    assert count(r"\bUTCHMMA\b") >= 100            # tcgen05.mma, every gather / reduce instantiation
    assert count(r"\bUTCHMMA\.2CTA\b") >= 8        # cta_group::2 (CTA pairs)
    assert count(r"\bUTMALDG\.4D\b") >= 40         # 4-D TMA boxes over the NHWC views
    assert count(r"\bUTMALDG\.3D\b") >= 10         # packed weight slabs
    assert count(r"\bLDTM\b") >= 50                # tcgen05.ld: accumulators out of TMEM
    assert count(r"\bUTCBAR\.2CTA\.MULTICAST\b") >= 1
    assert not re.search(r"\bHMMA\b|\bHGMMA\b", sass)   # no mma.sync / wgmma on the conv path


def test_attention_core_is_tcgen05_code(built):
    sass = _sass(built, "attn_umma")
    assert re.search(r"\bUTCHMMA\b", sass) and re.search(r"\bUTMALDG\.3D\b", sass) and re.search(r"\bLDTM\b", sass)


def test_tensor_core_kernels_barely_spill(built):
    worst = {}
    for log in ("conv_umma.o.log", "attn_umma.o.log"):
        name = None
        for line in open(os.path.join(OBJ, log)):
            m = re.search(r"Compiling entry function '(\w+)'", line)
            if m:
                name = m.group(1)
            m = re.search(r"(\d+) bytes spill stores, (\d+) bytes spill loads", line)
            if m and name:
                worst[name] = max(int(m.group(1)), int(m.group(2)))
    assert len(worst) >= 20, len(worst)
    bad = {k: v for k, v in worst.items() if v > 64}
    assert not bad, bad
