"""Complex-domain FFT plan (ref: dsp/fft/backends/kissfft_adapter.cc:171-246,
tests/fft_test.cc complex cases): forward_complex / inverse_complex against the oracle's kissfft
restatement (pinned bit-for-bit to the compiled reference, tests/test_oracle_vs_ref.py), 1e-5 rel-L2;
batch / stride addressing; clean-up rule on the inverse; round trip; error behaviour."""
import numpy as np
import pytest
from conftest import rel_l2

pytestmark = pytest.mark.gpu
TOL = 1e-5
SIZES = [128, 256, 512, 1024, 2048]


def crand(rng, n):
    return (rng.standard_normal(n) + 1j * rng.standard_normal(n)).astype(np.complex64)


@pytest.mark.parametrize("n", SIZES)
def test_forward_inverse_vs_oracle(cuda, dsp, oracle, n):
    rng = np.random.default_rng(n)
    plan = dsp.FftPlan(n, batch=16, complex_domain=True)
    x = crand(rng, 16 * n)
    X = plan.forward_complex(x, 16)
    assert rel_l2(X, oracle.cfft(x, n, False, 16)) < TOL
    y = plan.inverse_complex(X, 16)
    assert rel_l2(y, oracle.cfft(X, n, True, 16)) < TOL
    assert rel_l2(y, x) < TOL                                    # round trip (tests/fft_test.cc RMS < 1e-5)
    # impulse at 1 -> exp(-2 pi i k/n); constant -> n at DC
    e = np.zeros(n, np.complex64); e[1] = 1
    assert np.allclose(plan.forward_complex(e, 1), np.exp(-2j * np.pi * np.arange(n) / n), atol=2e-6)
    assert abs(plan.forward_complex(np.ones(n, np.complex64), 1)[0] - n) < 1e-3


@pytest.mark.parametrize("n", [128, 1024, 2048])
def test_batch_and_strides(cuda, dsp, oracle, n):
    rng = np.random.default_rng(n + 1)
    for b, si, so in ((3, 1, 1), (2, 2, 3), (4, 3, 1)):
        plan = dsp.FftPlan(n, batch=b, stride_in=si, stride_out=so, complex_domain=True)
        x = crand(rng, b * si * n * si + 4)
        want = oracle.cfft(x, n, False, b, si, so)
        got = plan.forward_complex(x, b, out_len=want.size)
        assert rel_l2(got, want) < TOL
        assert np.array_equal(got == 0, want == 0)               # untouched gaps stay zero
        wanti = oracle.cfft(x, n, True, b, si, so)
        assert rel_l2(plan.inverse_complex(x, b, out_len=wanti.size), wanti) < TOL


@pytest.mark.parametrize("n", [256, 2048])
def test_inverse_cleanup_rule(cuda, dsp, oracle, n):
    rng = np.random.default_rng(3)
    plan = dsp.FftPlan(n, complex_domain=True)
    x = crand(rng, n)
    x[5] = complex(np.nan, np.nan)
    assert np.all(plan.inverse_complex(x, 1) == 0)               # NaN everywhere -> every sample scrubbed
    tiny = (crand(rng, n) * 1e-33).astype(np.complex64)
    assert np.all(plan.inverse_complex(tiny, 1) == 0)            # |v| < 1e-30 -> 0
    assert np.all(oracle.cfft(tiny, n, True) == 0)
    assert np.isnan(plan.forward_complex(x, 1)).all()            # the forward transform does not scrub


@pytest.mark.parametrize("n", SIZES)
def test_device_rows(cuda, dsp, oracle, n):
    rng = np.random.default_rng(n + 9)
    rows = 1000 + n % 7
    x = crand(rng, rows * n).reshape(rows, n)
    plan = dsp.FftPlan(n, complex_domain=True)
    xd = cuda.from_numpy(x).cuda()
    X = plan.complex_device(xd)
    want = np.fft.fft(x.astype(np.complex128), axis=1)
    assert rel_l2(X.cpu().numpy(), want) < TOL
    for r in (0, rows - 1):
        assert rel_l2(X[r].cpu().numpy(), oracle.cfft(x[r], n)) < TOL
    y = plan.complex_device(X, inverse=True)
    assert rel_l2(y.cpu().numpy(), x) < TOL


def test_error_behaviour(cuda, dsp):
    with pytest.raises(dsp.CrlotError, match="Batch size must be between 1 and 16"):
        dsp.FftPlan(256, batch=17, complex_domain=True)
    with pytest.raises(dsp.CrlotError, match="Stride must be at least 1"):
        dsp.FftPlan(256, stride_in=0, complex_domain=True)
    with pytest.raises(dsp.CrlotError, match="In-place"):
        dsp.FftPlan(256, in_place=True, complex_domain=True)
    dsp.FftPlan(100, complex_domain=True)                          # any size <= 8192 (tests/test_gpu_generic.py)
    with pytest.raises(dsp.CrlotUnsupported):
        dsp.FftPlan(8200, complex_domain=True)
    real, cplx = dsp.FftPlan(256), dsp.FftPlan(256, batch=2, complex_domain=True)
    with pytest.raises(dsp.CrlotError, match="Complex FFT not supported for Real domain plan"):
        real.forward_complex(np.zeros(256, np.complex64))
    with pytest.raises(dsp.CrlotError, match="Real FFT not supported for Complex domain plan"):
        cplx.forward(np.zeros(256, np.float32))
    with pytest.raises(dsp.CrlotError, match="Invalid batch size"):
        cplx.forward_complex(np.zeros(3 * 256, np.complex64), 3)
