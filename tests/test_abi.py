"""CPU-side checks of the native boundary: the shared library loads without a GPU, exports every
symbol include/medvae_b200.h declares, its host-only helpers agree with the reference's conv
geometry, and every compute entry point refuses to run without a device (no CPU fallback)."""
import ctypes as C
import os
import re

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _lib():
    from medvae_disentangled_multimodal_b200 import _lib as L
    return L


def header_symbols():
    txt = open(os.path.join(ROOT, "include", "medvae_b200.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(mvae_[a-z0-9_]+)\s*\(", txt)))


def test_library_loads_and_exports_every_declared_symbol():
    L = _lib()
    lib = L.lib()
    syms = header_symbols()
    assert len(syms) >= 25
    for s in syms:
        assert hasattr(lib, s), f"{s} declared in medvae_b200.h but not exported"
    assert set(L.SIGNATURES) == set(syms), set(L.SIGNATURES) ^ set(syms)
    assert lib.mvae_abi_version() == 5


@pytest.mark.parametrize("h,mode,k,expect", [(28, 0, 3, 28), (28, 1, 3, 14), (14, 1, 3, 7), (7, 1, 3, 3),
                                             (64, 1, 3, 32), (7, 2, 3, 14), (16, 0, 1, 16)])
def test_conv_geometry_matches_reference(h, mode, k, expect):
    """Downsample: pad (0,1,0,1) then conv3x3 s2 p0 -> floor((H-2)/2)+1 (encoder_decoder.py:184-188);
    Upsample doubles (:205-209)."""
    L = _lib()
    from medvae_disentangled_multimodal_b200 import ops
    d = ops.conv_desc(2, h, h, 8, 8, k, mode, L.F32)
    assert ops.conv_out_hw(d) == (expect, expect)
    ref = torch.zeros(1, 1, h, h)
    if mode == 1:
        ref = torch.nn.functional.conv2d(torch.nn.functional.pad(ref, (0, 1, 0, 1)), torch.zeros(1, 1, 3, 3), stride=2)
    elif mode == 2:
        ref = torch.nn.functional.interpolate(ref, scale_factor=2.0)
    assert ref.shape[-1] == expect
    assert ops.conv_num_slabs(d) == (16 if mode == 2 else k * k)


def test_engine_selection():
    L = _lib()
    from medvae_disentangled_multimodal_b200 import ops
    lib = L.lib()
    yes = ops.conv_desc(8, 16, 16, 128, 256, 3, 0, L.BF16)
    assert lib.mvae_conv2d_uses_umma(C.byref(yes)) == 1
    for d in (ops.conv_desc(8, 16, 16, 128, 256, 3, 0, L.F32),      # fp32 verification mode
              ops.conv_desc(8, 16, 16, 6, 128, 3, 0, L.BF16),       # encoder conv_in: K = 54
              ops.conv_desc(8, 16, 16, 128, 3, 3, 0, L.BF16),       # decoder conv_out: N = 3
              ops.conv_desc(8, 16, 16, 128, 256, 3, 0, L.BF16, L.ENGINE_SIMT)):
        assert lib.mvae_conv2d_uses_umma(C.byref(d)) == 0
    assert lib.mvae_conv2d_wgrad_workspace_bytes(C.byref(yes)) >= 4 * 9 * 128 * 256


def test_bad_descriptor_is_an_error_not_a_crash():
    L = _lib()
    from medvae_disentangled_multimodal_b200 import ops
    bad = ops.conv_desc(1, 8, 8, 8, 8, 5, 0, L.F32)
    oh, ow = C.c_int32(), C.c_int32()
    assert L.lib().mvae_conv2d_out_hw(C.byref(bad), C.byref(oh), C.byref(ow)) != 0
    assert "bad conv descriptor" in L.last_error()


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-device behaviour")
def test_no_cpu_fallback():
    import medvae_disentangled_multimodal_b200 as M
    L = _lib()
    assert L.lib().mvae_device_ok() == 0
    model = M.BaseVAE(1, 4, 8, [1, 2], 1, [], resolution=16)
    with pytest.raises(L.MvaeError):
        model(torch.randn(1, 1, 16, 16))
    with pytest.raises(L.MvaeError):
        model.encode(torch.randn(1, 1, 16, 16))
    with pytest.raises(L.MvaeError):
        model.decode(torch.randn(1, 4, 8, 8))
    with pytest.raises(L.MvaeError):
        M.VAELoss()(inputs=torch.zeros(1), reconstructions=torch.zeros(1),
                    posteriors=torch.distributions.Normal(torch.zeros(1), torch.ones(1)))


def header_prototypes():
    """-> {name: (return C type, [parameter C types])} parsed from include/medvae_b200.h"""
    txt = open(os.path.join(ROOT, "include", "medvae_b200.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    txt = re.sub(r"//[^\n]*", "", txt)
    out = {}
    for m in re.finditer(r"([A-Za-z_][A-Za-z0-9_ ]*?[\s\*]+)(mvae_[a-z0-9_]+)\s*\(([^)]*)\)\s*;", txt):
        ret, name, args = m.group(1).strip(), m.group(2), m.group(3).strip()
        params = [] if args in ("", "void") else [a.strip() for a in args.split(",")]
        out[name] = (ret, params)
    return out


def _ctype_class(decl: str) -> str:
    """coarse ABI class of a C parameter / return declaration"""
    d = decl.replace("const", " ").strip()
    if "*" in d:
        return "ptr"
    words = d.split()
    ty = " ".join(words[:-1]) if len(words) > 1 and words[-1] not in ("int", "float", "double", "long") and \
        not words[-1].endswith("_t") else " ".join(words)
    return {"int": "i32", "int32_t": "i32", "int64_t": "i64", "long long": "i64", "uint64_t": "u64", "size_t": "u64",
            "float": "f32", "double": "f64", "unsigned long long": "u64"}[ty]


def _ctypes_class(t) -> str:
    if t is None:
        return "void"
    if t in (C.c_void_p, C.c_char_p) or hasattr(t, "contents") or issubclass(t, C._Pointer):
        return "ptr"
    return {C.c_int: "i32", C.c_int32: "i32", C.c_int64: "i64", C.c_longlong: "i64", C.c_uint64: "u64",
            C.c_size_t: "u64", C.c_ulonglong: "u64", C.c_float: "f32", C.c_double: "f64"}[t]


def test_ctypes_signatures_match_the_header_prototypes():
    """argument count and ABI class (pointer / 32-bit int / 64-bit int / float / double) of every entry of
    _lib.SIGNATURES against the prototype in include/medvae_b200.h: a drifted binding would otherwise only show up
    as garbage arguments on the GPU box."""
    L = _lib()
    protos = header_prototypes()
    assert set(protos) == set(L.SIGNATURES), set(protos) ^ set(L.SIGNATURES)
    for name, (ret, params) in sorted(protos.items()):
        restype, argtypes = L.SIGNATURES[name]
        assert len(argtypes) == len(params), f"{name}: header has {len(params)} parameters, ctypes table {len(argtypes)}"
        got = [_ctypes_class(t) for t in argtypes]
        want = [_ctype_class(p) for p in params]
        assert got == want, f"{name}: header {want} vs ctypes {got}"
        assert _ctypes_class(restype) == _ctype_class(ret + " x"), f"{name}: return type"


def test_plain_c99_host_compiles_links_and_runs():
    """include/medvae_b200.h is a C header (no C++ / torch types) and the library links from a C program."""
    import shutil
    import subprocess
    gcc = shutil.which("gcc")
    if gcc is None:
        pytest.skip("gcc not available")
    L = _lib()
    libdir = os.path.dirname(L.LIB_PATH)
    out_dir = os.path.join(ROOT, "tests", "_build")
    os.makedirs(out_dir, exist_ok=True)
    exe = os.path.join(out_dir, "c_host")
    subprocess.run([gcc, "-std=c99", "-Wall", "-Wextra", "-Werror", "-pedantic", "-I", os.path.join(ROOT, "include"),
                    os.path.join(ROOT, "tests", "c_host.c"), "-o", exe, "-L", libdir, "-lmedvae_b200",
                    f"-Wl,-rpath,{libdir}"], check=True, capture_output=True)
    r = subprocess.run([exe], capture_output=True, text=True)
    assert r.returncode == 0, (r.returncode, r.stdout, r.stderr)
    assert "c host ok" in r.stdout


# every distinct conv of the C4 model (H, Cin, Cout, k, mode) and of C1-C3 (28 x 28)
_C4_CONVS = [(8, 1024, 1024, 3, 0), (64, 128, 128, 3, 0), (32, 256, 256, 3, 0), (16, 512, 512, 3, 0),
             (8, 1024, 1024, 3, 2), (16, 512, 512, 3, 2), (32, 256, 256, 3, 2),
             (16, 1024, 512, 3, 0), (32, 512, 256, 3, 0), (64, 256, 128, 3, 0),
             (32, 128, 256, 3, 0), (16, 256, 512, 3, 0), (8, 512, 1024, 3, 0),
             (64, 128, 128, 3, 1), (32, 256, 256, 3, 1), (16, 512, 512, 3, 1),
             (8, 1024, 256, 3, 0), (8, 128, 1024, 3, 0), (8, 1024, 3072, 1, 0), (8, 1024, 1024, 1, 0),
             (16, 512, 1536, 1, 0), (16, 512, 512, 1, 0), (32, 128, 256, 1, 0), (16, 256, 512, 1, 0),
             (8, 512, 1024, 1, 0), (16, 1024, 512, 1, 0), (32, 512, 256, 1, 0), (64, 256, 128, 1, 0)]
_C123_CONVS = [(7, 512, 512, 3, 0), (28, 128, 128, 3, 0), (14, 256, 256, 3, 0), (14, 256, 256, 3, 2), (7, 512, 512, 3, 2),
               (14, 512, 256, 3, 0), (28, 256, 128, 3, 0), (14, 128, 256, 3, 0), (7, 256, 512, 3, 0), (7, 512, 256, 3, 0),
               (28, 128, 128, 3, 1), (14, 256, 256, 3, 1), (7, 128, 512, 3, 0), (7, 512, 1536, 1, 0), (14, 256, 768, 1, 0)]


@pytest.mark.parametrize("h,cin,cout,k,mode", _C4_CONVS + _C123_CONVS)
def test_every_hidden_conv_of_the_bench_models_takes_the_tcgen05_path(h, cin, cout, k, mode):
    """north_star (a): the dense contractions run on tcgen05 — the engine selection (a host-side decision of the
    library) says so for every hidden conv shape of SURVEY Appendix B in bf16, and the GroupNorm either side rides the
    conv's epilogue wherever a GroupNorm consumes / produced the tensor (32 groups)."""
    L = _lib()
    from medvae_disentangled_multimodal_b200 import ops
    lib = L.lib()
    d = ops.conv_desc(256, h, h, cin, cout, k, mode, L.BF16)
    assert lib.mvae_conv2d_uses_umma(C.byref(d)) == 1
    if k == 3 and mode == 0:
        assert lib.mvae_conv2d_fuses_gn(C.byref(d), 32, 0) == 1      # statistics of the consumer GroupNorm (fprop)
        assert lib.mvae_conv2d_fuses_gn(C.byref(d), 32, 1) == 1      # GroupNorm backward behind the dgrad


def test_skinny_edge_convs_are_padded_onto_the_tcgen05_path():
    """encoder conv_in (6 -> 128) and decoder conv_out (128 -> 3) run with the skinny dimension zero-padded to 64
    (blocks.ConvOp.pad_in / pad_out)"""
    import torch.nn as nn
    from medvae_disentangled_multimodal_b200 import blocks
    L = _lib()
    from medvae_disentangled_multimodal_b200 import ops
    cin = blocks.ConvOp([nn.Conv2d(6, 128, 3, padding=1)], 0, pad_in=True)
    cout = blocks.ConvOp([nn.Conv2d(128, 3, 3, padding=1)], 0, pad_out=True)
    assert cin.cin_of(torch.bfloat16) == 64 and cin.cin_of(torch.float32) == 6
    assert cout.cout_of(torch.bfloat16) == 64 and cout.cout_of(torch.float32) == 3
    for op in (cin, cout):
        d = ops.conv_desc(256, 64, 64, op.cin_of(torch.bfloat16), op.cout_of(torch.bfloat16), 3, 0, L.BF16)
        assert L.lib().mvae_conv2d_uses_umma(C.byref(d)) == 1
