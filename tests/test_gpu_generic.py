"""Any-size plans (csrc/k_generic.cu): every even real size and every complex size the reference's
plan accepts (ref: dsp/fft/backends/kissfft_adapter.cc:13-66; kissfft is mixed radix with a generic
butterfly, third_party/kissfft/kiss_fft.c:236-328) runs on the GPU, not only the five register-FFT
sizes; so do STFT plans whose hop is not a multiple of 4.  The reference's own tests use nfft = 128
with strides 2/2 (tests/fft_test.cc:450-520) and a 513-point complex plan (tests/fft_test.cc:229-237).

Oracle: the kissfft restatement, pinned bit-for-bit to the compiled reference for non-power-of-two
sizes too (tests/test_oracle_vs_ref.py::test_fft_random_sizes).  Bar: 1e-5 relative L2 (fp32)."""
import numpy as np
import pyoracle as po
import pytest
from conftest import rel_l2

pytestmark = pytest.mark.gpu
TOL = 1e-5

REAL_SIZES = [2, 4, 6, 10, 30, 64, 128, 250, 480, 1000, 2 * 7 * 73, 2 * 509, 6000, 8192, 16384]
CPX_SIZES = [1, 2, 3, 5, 7, 17, 64, 100, 513, 1000, 1021, 4096, 8192]


def crand(rng, n):
    return (rng.standard_normal(n) + 1j * rng.standard_normal(n)).astype(np.complex64)


@pytest.mark.parametrize("n", REAL_SIZES)
def test_real_plan_any_even_size(cuda, dsp, oracle, n):
    rng = np.random.default_rng(n)
    b = 3
    plan = dsp.FftPlan(n, batch=b)
    x = rng.standard_normal(b * n).astype(np.float32)
    X = plan.forward(x, b)
    want = oracle.rfft(x, n, b)
    assert rel_l2(X, want) < TOL
    nb = n // 2 + 1
    Xr = X.reshape(b, nb)
    assert np.all(Xr[:, 0].imag == 0) and np.all(Xr[:, -1].imag == 0)      # exact zeros (kiss_fftr.c:93-99)
    y = plan.inverse(X, b)
    assert rel_l2(y, oracle.irfft(X, n, b)) < TOL
    assert rel_l2(y, x) < TOL                                             # round trip
    # the inverse uses only the real parts of bins 0 and n/2 (kiss_fftr.c:128-129)
    X2 = X.copy().reshape(b, nb)
    X2[:, 0] += 3j
    X2[:, -1] -= 5j
    assert np.array_equal(plan.inverse(X2.reshape(-1), b), y)


def test_reference_stride_layout_case(cuda, dsp, oracle):
    """tests/fft_test.cc:450-520 StrideLayoutVerification: nfft 128, batch 3, stride_in 2, stride_out 2."""
    n, b, si, so = 128, 3, 2, 2
    plan = dsp.FftPlan(n, batch=b, stride_in=si, stride_out=so)
    x = np.zeros(b * n * si * si, np.float32)
    for bb in range(b):
        for i in range(n):
            x[bb * si * n + i * si] = np.sin(2.0 * np.pi * (bb + 1) * i / n)
    want = oracle.rfft(x, n, b, si, so)
    got = plan.forward(x, b, out_len=want.size)
    assert rel_l2(got, want) < TOL
    nb = n // 2 + 1
    for bb in range(b):                                                    # peak at bin bb+1, as the reference asserts
        row = np.abs(got[bb * so * nb: bb * so * nb + so * nb: so])
        assert int(np.argmax(row)) == bb + 1
    wanti = oracle.irfft(got, n, b, so, si, out_len=x.size)
    assert rel_l2(dsp.FftPlan(n, batch=b, stride_in=so, stride_out=si).inverse(got, b, out_len=x.size), wanti) < TOL


@pytest.mark.parametrize("n", CPX_SIZES)
def test_complex_plan_any_size(cuda, dsp, oracle, n):
    rng = np.random.default_rng(1000 + n)
    b = 2
    plan = dsp.FftPlan(n, batch=b, complex_domain=True)
    x = crand(rng, b * n)
    X = plan.forward_complex(x, b)
    assert rel_l2(X, oracle.cfft(x, n, False, b)) < TOL
    assert rel_l2(X.reshape(b, n), np.fft.fft(x.reshape(b, n).astype(np.complex128), axis=1)) < TOL
    y = plan.inverse_complex(X, b)
    assert rel_l2(y, oracle.cfft(X, n, True, b)) < TOL
    assert rel_l2(y, x) < TOL


def test_complex_device_rows_and_cleanup(cuda, dsp, oracle):
    rng = np.random.default_rng(5)
    n, rows = 513, 700
    plan = dsp.FftPlan(n, complex_domain=True)
    x = crand(rng, rows * n).reshape(rows, n)
    X = plan.complex_device(cuda.from_numpy(x).cuda())
    assert rel_l2(X.cpu().numpy(), np.fft.fft(x.astype(np.complex128), axis=1)) < TOL
    assert rel_l2(plan.complex_device(X, inverse=True).cpu().numpy(), x) < TOL
    tiny = (crand(rng, n) * 1e-33).astype(np.complex64)
    assert np.all(plan.inverse_complex(tiny, 1) == 0)                      # |v| < 1e-30 -> 0 (adapter:227-239)
    bad = crand(rng, n)
    bad[7] = complex(np.nan, np.nan)
    assert np.all(plan.inverse_complex(bad, 1) == 0)


STFT_SHAPES = [(480, 160), (400, 100), (1000, 250), (1024, 250), (2048, 510), (96, 40), (6, 2), (128, 32), (8192, 2048)]


@pytest.mark.parametrize("N,H", STFT_SHAPES)
@pytest.mark.parametrize("mode", [po.DROP, po.ZERO_PAD])
def test_stft_roundtrip_any_size(cuda, dsp, oracle, N, H, mode):
    L = 37 * H + N + 13
    x = oracle.synth_uniform(0xBEEF + N, 0, 0, L)
    w = oracle.window(po.HANN, N, periodic=True)
    ref = oracle.roundtrip(x, N, H, w, mode)
    plan = dsp.Plan(N, H, analysis_window=w, ola_window=w, boundary_mode=mode)
    assert not plan.istft_ola_supported()
    out, spec, frames = plan.roundtrip(cuda.from_numpy(x[None, :]).cuda())
    cuda.cuda.synchronize()
    assert spec.shape[1] == ref["n_frames"]
    assert rel_l2(spec.cpu().numpy()[0], ref["spectra"]) < TOL
    assert rel_l2(out.cpu().numpy()[0], ref["out"]) < TOL
    host = plan.roundtrip_host(x[None, :])[0]
    assert np.array_equal(host.view(np.uint32), out.cpu().numpy()[0].view(np.uint32))


def test_stft_multichannel_unaligned_stride(cuda, dsp, oracle):
    """three channels whose stride is not a multiple of 4 floats: only the any-size kernels take that"""
    N, H, ch = 480, 160, 3
    L = 50 * H + N + 1
    w = oracle.window(po.HAMMING, N)
    xs = np.stack([oracle.synth_uniform(7, c, 0, L) for c in range(ch)])
    plan = dsp.Plan(N, H, analysis_window=w, ola_window=w, boundary_mode=po.DROP)
    out, spec, _ = plan.roundtrip(cuda.from_numpy(xs).cuda())
    cuda.cuda.synchronize()
    for c in range(ch):
        ref = oracle.roundtrip(xs[c], N, H, w, po.DROP)
        assert rel_l2(spec.cpu().numpy()[c], ref["spectra"]) < TOL
        assert rel_l2(out.cpu().numpy()[c], ref["out"]) < TOL


def test_cleanup_rule_any_size(cuda, dsp, oracle):
    """NaN / Inf / denormal-range input samples are zeroed before the transform (adapter:104-108)."""
    N, H = 480, 160
    L = 20 * H + N
    x = oracle.synth_uniform(3, 0, 0, L)
    x[5] = np.nan
    x[700] = np.inf
    x[1500] = 1e-35
    w = np.ones(N, np.float32)
    ref = oracle.roundtrip(x, N, H, w, po.DROP)
    plan = dsp.Plan(N, H, analysis_window=w, ola_window=w, boundary_mode=po.DROP)
    out, spec, _ = plan.roundtrip(cuda.from_numpy(x[None, :]).cuda())
    assert np.isfinite(spec.cpu().numpy()).all()
    assert rel_l2(spec.cpu().numpy()[0], ref["spectra"]) < TOL
    assert rel_l2(out.cpu().numpy()[0], ref["out"]) < TOL


def test_framequeue_and_hook_any_size(cuda, dsp, oracle):
    N, H, L = 480, 120, 5000
    x = oracle.synth_uniform(11, 0, 0, L)
    w = oracle.window(po.HANN, N)
    plan = dsp.Plan(N, H, analysis_window=w, ola_window=w)
    for pad_mode in (dsp.PAD_CONSTANT, dsp.PAD_REFLECT, dsp.PAD_EDGE):
        fr = oracle.framequeue(x, N, H, True, pad_mode)
        spec = plan.stft_framequeue(cuda.from_numpy(x).cuda(), center=True, pad_mode=pad_mode)
        assert spec.shape[-2] == fr.shape[0]
        assert rel_l2(spec.cpu().numpy().reshape(fr.shape[0], -1), oracle.rfft_many(fr * w, N)) < TOL
    # spectral hook in the any-size inverse == stand-alone pass, bit for bit
    rng = np.random.default_rng(2)
    X = crand(rng, 40 * (N // 2 + 1)).reshape(1, 40, N // 2 + 1)
    g = rng.uniform(0, 2, N // 2 + 1).astype(np.float32)
    Xd, gd = cuda.from_numpy(X).cuda(), cuda.from_numpy(g).cuda()
    want = plan.istft_frames(plan.apply_spectral_gain(Xd.clone(), gd), True)
    plan.set_spectral_gain(gd)
    got = plan.istft_frames(Xd, True)
    plan.set_spectral_gain(None)
    assert np.array_equal(got.cpu().numpy().view(np.uint32), want.cpu().numpy().view(np.uint32))


def test_error_behaviour(cuda, dsp):
    with pytest.raises(dsp.CrlotError, match="FFT size must be even for real FFT"):
        dsp.FftPlan(513)                                                   # tests/fft_test.cc:225-227
    dsp.FftPlan(513, complex_domain=True)                                  # tests/fft_test.cc:235-237: no throw
    with pytest.raises(dsp.CrlotUnsupported):
        dsp.FftPlan(16386)
    with pytest.raises(dsp.CrlotUnsupported):
        dsp.FftPlan(8193, complex_domain=True)
    dsp.Stream(dsp.Plan(480, 160, boundary_mode=po.DROP), 4, 160)          # any-size plans stream too (gather + kernel 1)
    with pytest.raises(dsp.CrlotUnsupported):
        dsp.Stream(dsp.Plan(1000, 250, boundary_mode=po.DROP), 4, 250)     # ... unless hop % 4 != 0
