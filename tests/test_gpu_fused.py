"""Kernels 2+3 fused (csrc/k23_istft_ola.cu, C-ABI crlot_istft_ola_execute): the iFFT frames stay in
shared memory.  Bar: BIT-IDENTICAL to kernel 2 -> kernel 3 on the same spectra (same roundings in
the same order), hence the same parity to the oracle as the three-kernel path (1e-5 rel-L2 audio),
for every chunking, offset, window policy, ragged n_out and channel count."""
import numpy as np
import pyoracle as po
import pytest
from conftest import bits, rel_l2

pytestmark = pytest.mark.gpu
TOL = 1e-5
SHAPES = [(256, 64), (256, 16), (256, 128), (256, 256), (512, 128), (512, 32), (512, 96), (1024, 256), (1024, 512),
          (1024, 1024), (2048, 512), (2048, 64), (2048, 1024), (2048, 192), (2048, 2048),
          (4096, 1024), (4096, 512), (4096, 2048), (4096, 4096), (4096, 192), (4096, 64)]


def dev(t, a):
    return t.from_numpy(np.ascontiguousarray(a)).cuda()


def rand_spec(rng, ch, F, N):
    X = (rng.standard_normal((ch, F, N // 2 + 1)) + 1j * rng.standard_normal((ch, F, N // 2 + 1))).astype(np.complex64)
    return X


@pytest.mark.parametrize("N,H", SHAPES)
def test_fused_is_bit_identical_to_k2_k3(cuda, dsp, oracle, N, H):
    rng = np.random.default_rng(N * 7 + H)
    w = oracle.window(po.HANN, N)
    plan = dsp.Plan(N, H, analysis_window=w, ola_window=w, gain=0.8)
    assert plan.istft_ola_supported()
    for ch, F in ((1, 1), (1, 3), (2, 45), (3, 300), (1, 1500)):
        X = dev(cuda, rand_spec(rng, ch, F, N))
        want = plan.ola(plan.istft_frames(X, True), prewindowed=True)
        got = plan.istft_ola(X)
        cuda.cuda.synchronize()
        assert got.shape == want.shape
        assert np.array_equal(bits(got.cpu().numpy()), bits(want.cpu().numpy())), (ch, F)


@pytest.mark.parametrize("N,H", [(256, 64), (1024, 256), (2048, 512), (4096, 1024)])
def test_fused_offsets_and_ragged_output(cuda, dsp, oracle, N, H):
    """t_offset (envelope phase), frame_t0 < 0 (halo frames whose head is cut), frame_t0 > 0 (output
    begins before the first frame), n_out shorter / longer than the covered range, odd n_out."""
    rng = np.random.default_rng(N + 99)
    w = oracle.window(po.HAMMING, N)
    plan = dsp.Plan(N, H, analysis_window=w, ola_window=w)
    F = 130
    X = dev(cuda, rand_spec(rng, 2, F, N))
    fr = plan.istft_frames(X, True)
    full = (F - 1) * H + N
    for t_off, ft0, n_out in ((0, 0, full), (6 * H, -3 * H, full - 3 * H), (123456 * 2, -2 * H, 40 * H + 7),
                              (10, 2 * H, full + 2 * H + 3 * H), (0, -(N - H), 17 * H + 1), (2, 4, full + 64 + 4),
                              (0, 0, 5), (0, -(F + 2) * H, H)):
        # two channels need an even channel stride in the fused kernel: odd n_out runs on one channel
        Xc, frc = (X, fr) if n_out % 2 == 0 else (X[:1], fr[:1])
        want = plan.ola(frc, prewindowed=True, t_offset=t_off, frame_t0=ft0, n_out=n_out)
        got = plan.istft_ola(Xc, t_offset=t_off, frame_t0=ft0, n_out=n_out)
        cuda.cuda.synchronize()
        assert np.array_equal(bits(got.cpu().numpy()), bits(want.cpu().numpy())), (t_off, ft0, n_out)


def test_fused_window_policies_and_scrub(cuda, dsp, oracle):
    N, H, F = 1024, 256, 64
    rng = np.random.default_rng(5)
    w = oracle.window(po.HANN, N)
    Xh = rand_spec(rng, 1, F, N)
    Xh[0, 3, 7] = np.nan
    Xh[0, 9, 0] = np.inf
    Xh[0, 20] *= 1e-34                      # drives iFFT outputs below the 1e-30 clean-up threshold
    X = dev(cuda, Xh)
    for kw in (dict(ola_window=w), dict(ola_window=None), dict(push_window=w, apply_window_inside=False),
               dict(ola_window=w, scrub=False), dict(ola_window=w, gain=-1.25, eps=1e-3)):
        plan = dsp.Plan(N, H, analysis_window=w, **kw)
        want = plan.ola(plan.istft_frames(X, True), prewindowed=True)
        got = plan.istft_ola(X)
        cuda.cuda.synchronize()
        a, b = got.cpu().numpy(), want.cpu().numpy()
        assert np.array_equal(np.isnan(a), np.isnan(b)), kw
        assert np.array_equal(bits(a)[~np.isnan(a)], bits(b)[~np.isnan(b)]), kw


@pytest.mark.parametrize("N,H,wt", [(256, 64, po.HAMMING), (512, 128, po.HANN), (1024, 256, po.HANN), (2048, 512, po.HANN)])
def test_fused_roundtrip_vs_oracle(cuda, dsp, oracle, N, H, wt):
    rng = np.random.default_rng(N + 3)
    L = 411 * H + N + 17
    x = rng.uniform(-1, 1, L).astype(np.float32)
    w = oracle.window(wt, N)
    plan = dsp.Plan(N, H, analysis_window=w, ola_window=w, boundary_mode=dsp.DROP)
    out, spec = plan.roundtrip_fused(dev(cuda, x))
    out3, _, _ = plan.roundtrip(dev(cuda, x))
    cuda.cuda.synchronize()
    ref = oracle.roundtrip(x, N, H, w, po.DROP)
    assert rel_l2(spec.cpu().numpy()[0], ref["spectra"]) < TOL
    assert rel_l2(out.cpu().numpy()[0], ref["out"]) < TOL
    assert np.array_equal(bits(out.cpu().numpy()), bits(out3.cpu().numpy()))


def test_fused_many_chunks_large(cuda, dsp, oracle):
    """Enough frames that every warp of the grid takes several chunks (the warm-up re-accumulation
    at chunk starts is what this exercises), 4 channels."""
    N, H = 2048, 512
    w = oracle.window(po.HANN, N)
    plan = dsp.Plan(N, H, analysis_window=w, ola_window=w)
    x = dsp.synth_uniform(4, 1 << 23, seed=77)
    out, spec = plan.roundtrip_fused(x)
    out3, _, _ = plan.roundtrip(x, spec=spec)
    cuda.cuda.synchronize()
    assert cuda.equal(out.view(cuda.int32), out3.view(cuda.int32))


def test_fused_unsupported_shapes_fail_loudly(cuda, dsp, oracle):
    w = oracle.window(po.HANN, 2048)
    plan = dsp.Plan(2048, 500, analysis_window=w, ola_window=w)       # hop not a multiple of 64
    assert not plan.istft_ola_supported()
    X = dev(cuda, rand_spec(np.random.default_rng(0), 1, 4, 2048))
    with pytest.raises(dsp.CrlotUnsupported):
        plan.istft_ola(X)
    plan = dsp.Plan(2048, 512, analysis_window=w, ola_window=w)
    with pytest.raises(dsp.CrlotUnsupported):
        plan.istft_ola(X, t_offset=1)


def test_roundtrip_host_uses_fused_path_same_bits(cuda, dsp, oracle):
    N, H = 1024, 256
    w = oracle.window(po.HANN, N)
    plan = dsp.Plan(N, H, analysis_window=w, ola_window=w)
    x = np.random.default_rng(8).uniform(-1, 1, 300000).astype(np.float32)
    got = plan.roundtrip_host(x)
    want, _, _ = plan.roundtrip(dev(cuda, x))
    assert np.array_equal(bits(got), bits(want.cpu().numpy()))


def test_table_division_equals_fdiv_rn(cuda, dsp, oracle):
    """The fused kernel divides by tabulated envelope values with the 3-FFMA tail of __fdiv_rn
    (csrc/common.cuh); 2^30 hashed operand pairs per plan must agree bit for bit with __fdiv_rn."""
    for N, H, wt, eps in ((2048, 512, po.HANN, 1e-8), (1024, 256, po.HAMMING, 1e-3), (256, 64, po.BLACKMAN, 1e-8)):
        w = oracle.window(wt, N)
        plan = dsp.Plan(N, H, analysis_window=w, ola_window=w, eps=eps)
        assert plan.selftest_division(1 << 30, seed=N + 1) == 0


def test_fused_extreme_amplitudes_take_exact_division(cuda, dsp, oracle):
    """Accumulator values outside the table division's operand range (huge, tiny, zero) must still be
    bit-identical to kernel 3's __fdiv_rn."""
    N, H, F = 2048, 512, 40
    rng = np.random.default_rng(11)
    w = oracle.window(po.HANN, N)
    plan = dsp.Plan(N, H, analysis_window=w, ola_window=w, scrub=False)
    Xh = rand_spec(rng, 1, F, N)
    Xh[0, 5:9] *= 1e25            # |acc| ~ 1e24 > 2^60
    Xh[0, 15:19] *= 1e-27         # |acc| ~ 1e-26 < 2^-60 (kept: scrub is off)
    Xh[0, 25:30] = 0
    X = dev(cuda, Xh)
    want = plan.ola(plan.istft_frames(X, True), prewindowed=True)
    got = plan.istft_ola(X)
    cuda.cuda.synchronize()
    assert np.array_equal(bits(got.cpu().numpy()), bits(want.cpu().numpy()))
