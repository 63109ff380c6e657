"""Kernel 2 for N = 4096, second version (csrc/k2_4096v2.cu: two 1024-point sub-transforms, bulk-copied spectrum chunks,
o' parked in tensor memory) against the oracle (reference: kissfft_adapter.cc:124-168 -> kiss_fftri) and against the
first version (csrc/k2_istft.cu, CRLOT_K2_4096_V1=1: a different factorisation, so tolerance, not bits)."""
import os

import numpy as np
import pytest

import pyoracle as po

pytestmark = pytest.mark.gpu
TOL = 1e-5
N = 4096
BINS = N // 2 + 1


def rel_l2(a, b):
    a = np.asarray(a, np.float64)
    b = np.asarray(b, np.float64)
    return float(np.linalg.norm((a - b).ravel()) / max(np.linalg.norm(b.ravel()), 1e-300))


def bits(a):
    return np.ascontiguousarray(a).view(np.uint32)


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


def v1(fn):
    os.environ["CRLOT_K2_4096_V1"] = "1"
    try:
        return fn()
    finally:
        del os.environ["CRLOT_K2_4096_V1"]


def spectra(F, seed):
    rng = np.random.default_rng(seed)
    return (rng.standard_normal((F, BINS)) + 1j * rng.standard_normal((F, BINS))).astype(np.complex64)


@pytest.mark.parametrize("F", [1, 2, 3, 11, 12, 13, 77])
@pytest.mark.parametrize("win", [False, True])
def test_rows_vs_oracle_and_first_version(cuda, dsp, oracle, F, win):
    """Odd and even rows start 8 bytes apart modulo 16 (2049 bins per row): both staging offsets; the imaginary parts of
    bins 0 and 2048 are ignored (kiss_fftr.c:131-132)."""
    X = spectra(F, 100 + F)
    w = oracle.window(po.HANN, N)
    plan = dsp.Plan(N, N // 4, analysis_window=w, ola_window=w)
    Xd = cuda.from_numpy(X).cuda()
    y = plan.istft_frames(Xd, fuse_synthesis_window=win).cpu().numpy()
    y1 = v1(lambda: plan.istft_frames(Xd, fuse_synthesis_window=win)).cpu().numpy()
    ref = oracle.irfft_many(X, N)
    if win:
        ref = ref * w
    assert rel_l2(y, ref) < TOL and rel_l2(y1, ref) < TOL
    assert max(rel_l2(y[k], ref[k]) for k in range(F)) < TOL
    assert rel_l2(y, y1) < 1e-6
    X2 = X.copy()
    X2[:, 0].imag = 7.0
    X2[:, -1].imag = -3.0
    y2 = plan.istft_frames(cuda.from_numpy(X2).cuda(), fuse_synthesis_window=win).cpu().numpy()
    assert np.array_equal(bits(y2), bits(y))


def test_many_rows_every_warp_walks_its_ring(cuda, dsp, oracle):
    """More rows than one pass of the grid (148 x 12 warps): every warp carries its staging ring and mbarrier phases
    over several rows; a handful of rows against the oracle, all of them against the first version."""
    F = 5000
    X = spectra(F, 5)
    w = oracle.window(po.HANN, N)
    plan = dsp.Plan(N, N // 4, analysis_window=w, ola_window=w)
    Xd = cuda.from_numpy(X).cuda()
    y = plan.istft_frames(Xd, fuse_synthesis_window=True)
    y1 = v1(lambda: plan.istft_frames(Xd, fuse_synthesis_window=True))
    assert float((y.double() - y1.double()).norm() / y1.double().norm()) < 1e-6
    rows = [0, 1, 1775, 1776, 1777, 3553, 4998, 4999]
    ref = oracle.irfft_many(X[rows], N) * w
    got = y[rows].cpu().numpy()
    assert max(rel_l2(got[i], ref[i]) for i in range(len(rows))) < TOL
    # deterministic: the same bits on a second launch
    y_again = plan.istft_frames(Xd, fuse_synthesis_window=True)
    assert cuda.equal(y.view(cuda.int32), y_again.view(cuda.int32))


def test_cleanup_rule(cuda, dsp, oracle):
    """NaN / Inf / |v| < 1e-30 outputs become 0 (kissfft_adapter.cc:154-165): a row with a NaN bin is all zeros, a row
    of tiny values is all zeros, their neighbours are untouched."""
    X = spectra(6, 9)
    X[1, 100] = np.nan
    X[3] *= np.float32(1e-33)
    X[4, 7] = np.inf
    w = oracle.window(po.HANN, N)
    plan = dsp.Plan(N, N // 4, analysis_window=w, ola_window=w)
    y = plan.istft_frames(cuda.from_numpy(X).cuda()).cpu().numpy()
    ref = oracle.irfft_many(X, N)
    assert np.isfinite(y).all()
    for k in (1, 3, 4):
        assert np.array_equal(ref[k], np.zeros(N, np.float32)) and not y[k].any()
    for k in (0, 2, 5):
        assert rel_l2(y[k], ref[k]) < TOL


def test_spectral_hook_inside_the_quads(cuda, dsp, oracle):
    """Per-bin gains are applied to a lane's own bins and to the partner bins it reads from the staging ring: equal, bit
    for bit, to scaling the spectra first."""
    F = 40
    X = spectra(F, 21)
    rng = np.random.default_rng(22)
    g = rng.uniform(0.0, 2.0, (F, BINS)).astype(np.float32)
    w = oracle.window(po.HANN, N)
    plan = dsp.Plan(N, N // 4, analysis_window=w, ola_window=w)
    Xd = cuda.from_numpy(X).cuda()
    plan.set_spectral_gain(cuda.from_numpy(g).cuda())
    y = plan.istft_frames(Xd, fuse_synthesis_window=True).cpu().numpy()
    plan.set_spectral_gain(None)
    Xg = (X.view(np.float32).reshape(F, BINS, 2) * g[:, :, None]).reshape(F, 2 * BINS).view(np.complex64)
    y_ref = plan.istft_frames(cuda.from_numpy(np.ascontiguousarray(Xg)).cuda(), fuse_synthesis_window=True).cpu().numpy()
    assert np.array_equal(bits(y), bits(y_ref))
    assert rel_l2(y, oracle.irfft_many(Xg, N) * w) < TOL


def test_unaligned_buffers_take_the_first_version(cuda, dsp, oracle):
    """float4 stores and bulk copies want 16-byte aligned bases; anything else runs the first version: same bits as
    asking for it."""
    F = 9
    X = spectra(F, 31)
    w = oracle.window(po.HANN, N)
    plan = dsp.Plan(N, N // 4, analysis_window=w, ola_window=w)
    Xd = cuda.from_numpy(X).cuda()
    want = v1(lambda: plan.istft_frames(Xd)).cpu().numpy()
    buf = cuda.empty(F * N + 2, dtype=cuda.float32, device="cuda")
    out = buf[2:].view(F, N)                       # 8 bytes past a 16-byte boundary
    plan.istft_frames(Xd, out=out)
    assert np.array_equal(bits(out.cpu().numpy()), bits(want))
    sbuf = cuda.empty(F * BINS + 1, dtype=cuda.complex64, device="cuda")
    Xo = sbuf[1:].view(F, BINS)
    Xo.copy_(Xd)
    got = plan.istft_frames(Xo).cpu().numpy()
    assert np.array_equal(bits(got), bits(want))
