"""CPU tests of the C-ABI boundary: the shared library builds, loads and exports every declared symbol."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "kokoro_b200.h")


@pytest.fixture(scope="module")
def lib_path():
    from mlx_audio_b200 import _build
    return _build.build()


def _declared_symbols():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(kk_[a-z0-9_]+)\s*\(", src)))


def test_header_declares_expected_entry_points():
    names = _declared_symbols()
    for must in ["kk_conv1d_bf16_tc", "kk_conv1d_f32", "kk_hnsf_source", "kk_istft_head_n20", "kk_stft_f32",
                 "kk_istft_f32", "kk_logmel_f32", "kk_adain_act", "kk_instnorm_stats", "kk_fold_weight_norm"]:
        assert must in names


def test_library_exports_every_declared_symbol(lib_path):
    lib = ctypes.CDLL(lib_path)
    for name in _declared_symbols():
        assert hasattr(lib, name), f"{name} declared in kokoro_b200.h but not exported"


def test_ctypes_prototypes_cover_the_header(lib_path):
    from mlx_audio_b200 import _lib
    declared = set(_declared_symbols()) - {"kk_last_error"}
    assert declared == set(_lib.PROTOTYPES), declared ^ set(_lib.PROTOTYPES)
    lib = _lib.load()
    assert lib.kk_version() >= 100


def test_conv_desc_layout_matches_c(lib_path):
    # sizeof(kk_conv_desc) as laid out by ctypes must equal the C compiler's (checked via a tiny probe build)
    import subprocess, tempfile, textwrap
    from mlx_audio_b200._lib import ConvDesc
    src = textwrap.dedent(f'''
        #include <stdio.h>
        #include <stddef.h>
        #include "{HEADER}"
        int main(void) {{ printf("%zu %zu %zu %zu\\n", sizeof(kk_conv_desc), offsetof(kk_conv_desc, bias),
                          offsetof(kk_conv_desc, out), offsetof(kk_conv_desc, act_slope)); return 0; }}''')
    with tempfile.TemporaryDirectory() as td:
        c = os.path.join(td, "p.c")
        open(c, "w").write(src)
        exe = os.path.join(td, "p")
        subprocess.check_call(["gcc", c, "-o", exe])
        size, o_bias, o_out, o_slope = map(int, subprocess.check_output([exe]).split())
    assert ctypes.sizeof(ConvDesc) == size
    assert ConvDesc.bias.offset == o_bias and ConvDesc.out.offset == o_out and ConvDesc.act_slope.offset == o_slope


def test_no_cpu_fallback_without_cuda():
    import torch
    if torch.cuda.is_available():
        pytest.skip("CUDA present")
    from mlx_audio_b200 import _lib, utils
    with pytest.raises(_lib.KokoroB200Error):
        utils.stft(torch.zeros(4096))
    from mlx_audio_b200.tts.models.kokoro import Decoder
    from mlx_audio_b200 import synth
    dec = Decoder(**synth.KOKORO_DECODER)
    with pytest.raises(_lib.KokoroB200Error):
        dec.load_weights({})


def test_product_does_not_import_oracle():
    # the product package must never reach into oracle/
    pkg = os.path.join(ROOT, "mlx_audio_b200")
    for dp, _, files in os.walk(pkg):
        for f in files:
            if f.endswith(".py"):
                src = open(os.path.join(dp, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", src, flags=re.M), os.path.join(dp, f)
