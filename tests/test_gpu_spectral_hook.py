"""Spectral-domain hook between kernels 1 and 2 (the gap the reference leaves between
IFftPlan::forward and IFftPlan::inverse, ref: bench/e2e_benchmark.cc:158-166).

Definition: X'[row][k] = (fl(Re X * g), fl(Im X * g)), g real, per bin (equaliser) or per row and
bin (time-frequency mask).  Bars:
  * the hook fused into the inverse kernels' loads == the stand-alone pass followed by the plain
    kernels, BIT-IDENTICAL (K2 for every N incl. 4096, fused K2+K3, host-buffer path);
  * against the oracle: oracle inverse + OLA of the numpy-multiplied oracle spectra, 1e-5 rel-L2;
  * the spectra written by kernel 1 are untouched by the fused hook; a gain of ones is the identity.
"""
import numpy as np
import pyoracle as po
import pytest
from conftest import bits, rel_l2

pytestmark = pytest.mark.gpu
TOL = 1e-5


def dev(t, a):
    return t.from_numpy(np.ascontiguousarray(a)).cuda()


def rand_spec(rng, ch, F, N):
    return (rng.standard_normal((ch, F, N // 2 + 1)) + 1j * rng.standard_normal((ch, F, N // 2 + 1))).astype(np.complex64)


def mul_real(X, g):
    """complex64 * float32 with one rounding per component (not numpy's complex product)."""
    out = np.empty_like(X)
    out.real = X.real * g
    out.imag = X.imag * g
    return out


@pytest.mark.parametrize("N,H", [(256, 64), (512, 128), (1024, 256), (2048, 512), (4096, 1024)])
@pytest.mark.parametrize("mask", [False, True])
def test_hook_in_k2_equals_separate_pass(cuda, dsp, oracle, N, H, mask):
    rng = np.random.default_rng(N + int(mask))
    w = oracle.window(po.HANN, N)
    plan = dsp.Plan(N, H, analysis_window=w, ola_window=w)
    for ch, F in ((1, 1), (2, 37), (1, 301)):
        Xh = rand_spec(rng, ch, F, N)
        gh = (rng.uniform(0.0, 2.0, (ch, F, N // 2 + 1)) if mask else rng.uniform(0.0, 2.0, N // 2 + 1)).astype(np.float32)
        X, g = dev(cuda, Xh), dev(cuda, gh)
        # the stand-alone pass is the definition
        X2 = plan.apply_spectral_gain(X.clone(), g)
        cuda.cuda.synchronize()
        assert np.array_equal(bits(X2.cpu().numpy()), bits(mul_real(Xh, gh)))
        want = plan.istft_frames(X2, True)
        plan.set_spectral_gain(g)
        got = plan.istft_frames(X, True)
        plan.set_spectral_gain(None)
        cuda.cuda.synchronize()
        assert np.array_equal(bits(X.cpu().numpy()), bits(Xh)), "fused hook must not modify the spectra"
        assert np.array_equal(bits(got.cpu().numpy()), bits(want.cpu().numpy())), (ch, F)
        # cleared hook: plain inverse again
        plain = plan.istft_frames(X, True)
        ones = dev(cuda, np.ones(N // 2 + 1, np.float32))
        plan.set_spectral_gain(ones)
        ident = plan.istft_frames(X, True)
        plan.set_spectral_gain(None)
        assert np.array_equal(bits(ident.cpu().numpy()), bits(plain.cpu().numpy()))


@pytest.mark.parametrize("N,H", [(256, 64), (1024, 256), (2048, 512), (4096, 1024)])
@pytest.mark.parametrize("mask", [False, True])
def test_hook_in_fused_k23_equals_separate_pass(cuda, dsp, oracle, N, H, mask):
    rng = np.random.default_rng(3 * N + int(mask))
    w = oracle.window(po.HAMMING, N)
    plan = dsp.Plan(N, H, analysis_window=w, ola_window=w, gain=0.9)
    for ch, F in ((1, 2), (2, 45), (3, 700)):
        Xh = rand_spec(rng, ch, F, N)
        gh = (rng.uniform(0.0, 2.0, (ch, F, N // 2 + 1)) if mask else rng.uniform(0.0, 2.0, N // 2 + 1)).astype(np.float32)
        X, g = dev(cuda, Xh), dev(cuda, gh)
        X2 = plan.apply_spectral_gain(X.clone(), g)
        want = plan.ola(plan.istft_frames(X2, True), prewindowed=True)
        plan.set_spectral_gain(g)
        got = plan.istft_ola(X)
        plan.set_spectral_gain(None)
        cuda.cuda.synchronize()
        assert np.array_equal(bits(got.cpu().numpy()), bits(want.cpu().numpy())), (ch, F)


@pytest.mark.parametrize("N,H", [(1024, 256), (2048, 512), (4096, 1024)])
def test_hooked_roundtrip_against_oracle(cuda, dsp, oracle, N, H):
    """x -> K1 -> (gain) -> K2 -> K3 against the oracle: oracle spectra times g, oracle inverse, oracle OLA."""
    L = 40 * H + N
    x = oracle.synth_uniform(0xC0FFEE, 0, 0, L)
    w = oracle.window(po.HANN, N)
    g = (0.25 + 1.5 * np.abs(np.sin(np.arange(N // 2 + 1) * 0.01))).astype(np.float32)
    ref = oracle.roundtrip(x, N, H, w, po.DROP)
    F = ref["n_frames"]
    Y = oracle.irfft_many(mul_real(ref["spectra"], g), N)
    want, _, _ = oracle.ola_stream(Y.reshape(F, 1, N), N, H, 1, w)
    plan = dsp.Plan(N, H, analysis_window=w, ola_window=w, boundary_mode=dsp.DROP)
    gd = dev(cuda, g)
    plan.set_spectral_gain(gd)
    out, spec, frames = plan.roundtrip(dev(cuda, x[None, :]))
    outs = [out.cpu().numpy()[0]]
    if plan.istft_ola_supported():
        outs.append(plan.roundtrip_fused(dev(cuda, x[None, :]))[0].cpu().numpy()[0])
    # host-buffer path (crlot_roundtrip_host) honours the hook as well
    outs.append(plan.roundtrip_host(x[None, :])[0])
    plan.set_spectral_gain(None)
    cuda.cuda.synchronize()
    assert rel_l2(spec.cpu().numpy()[0], ref["spectra"]) < TOL       # spectra are the un-hooked ones
    for o in outs:
        assert o.shape == want[0].shape
        assert rel_l2(o, want[0]) < TOL
    assert np.array_equal(bits(outs[0]), bits(outs[-1]))             # device and host paths agree bit for bit


def test_hook_argument_checks(cuda, dsp, oracle):
    plan = dsp.Plan(1024, 256)
    g = cuda.ones(plan.bins - 1, dtype=cuda.float32, device="cuda")
    with pytest.raises(dsp.CrlotInvalidArgument):
        plan.set_spectral_gain(g)
    import ctypes as C
    rc = dsp.lib().crlot_stft_plan_set_spectral_gain(plan._h, C.c_void_p(g.data_ptr()), C.c_longlong(plan.bins - 1))
    assert rc == 1      # CRLOT_ERR_INVALID_ARGUMENT: stride smaller than a row
