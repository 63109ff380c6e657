"""Kernel 1 for N = 2048 / hop 512 and N = 1024 / hop 256, 16-warp version with the raw samples carried in tensor memory (csrc/k1_2048tm.cu)
against the ring kernel it replaces by default (stft_fwd_kernel<2048>, CRLOT_K1_NO_TM16=1): the same arithmetic in the
same order, so BIT-IDENTICAL spectra; and against the oracle (kissfft_adapter.cc:83-122 -> kiss_fftr)."""
import os

import numpy as np
import pytest

import pyoracle as po

pytestmark = pytest.mark.gpu


@pytest.fixture()
def cuda():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    torch.cuda.set_device(0)
    return torch


@pytest.fixture()
def dsp():
    import crlot_dsp_b200
    return crlot_dsp_b200


def ring_kernel(fn):
    os.environ["CRLOT_K1_NO_TM16"] = "1"
    try:
        return fn()
    finally:
        del os.environ["CRLOT_K1_NO_TM16"]


@pytest.mark.parametrize("N,H", [(2048, 512), (1024, 256)])
@pytest.mark.parametrize("mode,L,ch", [(po.DROP, (1 << 22) + 12344, 2), (po.ZERO_PAD, 70 * 512 + 76, 3), (po.ZERO_PAD, 70 * 512 + 77, 1),
                                       (po.DROP, 2048, 1), (po.ZERO_PAD, 5, 1), (po.DROP, 64 * 512 + 2048, 1), (po.DROP, 65 * 512 + 2048, 1),
                                       (po.DROP, 127 * 256 + 1024, 1), (po.DROP, 128 * 256 + 1024, 1), (po.ZERO_PAD, 129 * 256 + 3, 1)])
def test_bit_identical_to_the_ring_kernel(cuda, dsp, oracle, N, H, mode, L, ch):
    """Chunk boundaries (64 steps per chunk: 65 and 66 frames at N = 2048, 128 to 130 at N = 1024, where a step is two
    frames and an odd count leaves half a step), ragged ZERO_PAD tails, several channels, and the clean-up
    rule on raw samples that enter the ring once (NaN, a denormal-range value)."""
    w = oracle.window(po.HANN, N)
    plan = dsp.Plan(N, H, analysis_window=w, ola_window=w, boundary_mode=mode)
    x = dsp.synth_uniform(ch, L, seed=L)
    x[0, L // 3] = float("nan")
    x[-1, L // 2] = 1e-35
    a = plan.stft(x).clone()
    b = ring_kernel(lambda: plan.stft(x).clone())
    assert a.shape == b.shape
    assert cuda.equal(cuda.view_as_real(a).contiguous().view(cuda.int32), cuda.view_as_real(b).contiguous().view(cuda.int32))


@pytest.mark.parametrize("N,H", [(2048, 512), (1024, 256)])
def test_against_the_oracle(cuda, dsp, oracle, N, H):
    w = oracle.window(po.HANN, N)
    plan = dsp.Plan(N, H, analysis_window=w, ola_window=w, boundary_mode=po.DROP)
    L = 200 * H + N
    x = dsp.synth_uniform(1, L, seed=77)
    S = plan.stft(x).cpu().numpy()[0]
    ref = oracle.roundtrip(x.cpu().numpy()[0], N, H, w, po.DROP)["spectra"]
    err = np.linalg.norm((S - ref).ravel()) / np.linalg.norm(ref.ravel())
    assert err < 1e-5
    assert not S[:, 0].imag.any() and not S[:, -1].imag.any()      # kiss_fftr.c:93-99: exact zeros
