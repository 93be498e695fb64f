"""kk_hnsf_source_noalloc (operators' addition, no reference counterpart): the harmonic source with its phase scratch in
a static device array - checked through the C ABI against kk_hnsf_source (caller-provided scratch) on two calls in a
row.  Collected last on purpose."""
import numpy as np
import pytest
import torch

from mlx_audio_b200 import _lib as L
from mlx_audio_b200 import synth

pytestmark = pytest.mark.gpu

MAX_FLOATS = 1 << 20   # KK_HNSF_NOALLOC_MAX_FLOATS (include/kokoro_b200.h)


def _both(lib, B, T, seed, lens=None):
    w = synth.synth_decoder_weights(seed=2000)
    inp = synth.synth_decoder_inputs(B, T, seed=seed)
    f0 = torch.from_numpy(inp["F0_curve"]).cuda()
    noise = torch.from_numpy(inp["noise"]).cuda()
    lw = torch.from_numpy(w["decoder.generator.m_source.l_linear.weight"].reshape(-1).copy()).cuda()
    lb = torch.from_numpy(w["decoder.generator.m_source.l_linear.bias"].reshape(-1).copy()).cuda()
    ld = torch.tensor(lens, dtype=torch.int32, device="cuda") if lens is not None else None
    ws = torch.empty(B, 9, 2 * T, device="cuda")
    ref = torch.full((B, 600 * T), 7.0, device="cuda")
    got = torch.full((B, 600 * T), -7.0, device="cuda")
    st = L.stream()
    L.check(lib.kk_hnsf_source(L.ptr(f0), L.ptr(noise), 0, None, L.ptr(lw), L.ptr(lb), L.ptr(ws), L.ptr(ref), B, 2 * T,
                               300, L.ptr(ld), st), "kk_hnsf_source")
    L.check(lib.kk_hnsf_source_noalloc(L.ptr(f0), L.ptr(noise), 0, None, L.ptr(lw), L.ptr(lb), L.ptr(got), B, 2 * T, 300,
                                       L.ptr(ld), st), "kk_hnsf_source_noalloc")
    torch.cuda.synchronize()
    return ref.cpu().numpy(), got.cpu().numpy()


def test_hnsf_source_noalloc_matches_on_two_calls_in_a_row():
    L.require_cuda()
    lib = L.load()
    # call 1 fills the shared static scratch; call 2 (another shape, ragged) must not see anything of call 1
    for B, T, seed, lens in ((3, 24, 77, None), (2, 40, 78, [40, 13])):
        ref, got = _both(lib, B, T, seed, lens)
        assert np.isfinite(got).all() and np.abs(ref).max() > 1e-3
        np.testing.assert_allclose(got, ref, rtol=0, atol=2e-6)   # the tolerance of test_hnsf_source_and_stft_n20


def test_hnsf_source_noalloc_rejects_sizes_beyond_the_static_maximum():
    L.require_cuda()
    lib = L.load()
    B, T2 = 64, MAX_FLOATS // (64 * 9) + 1           # B * 9 * T2 just over the maximum: rejected before any launch
    dummy = torch.zeros(16, device="cuda")
    rc = lib.kk_hnsf_source_noalloc(L.ptr(dummy), None, 0, None, L.ptr(dummy), L.ptr(dummy), L.ptr(dummy), B, T2, 300,
                                    None, L.stream())
    assert rc == L.KK_ERR_INVALID_ARG and "exceed the static maximum" in L.last_error()
