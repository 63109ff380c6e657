"""FrameQueue centre / pad front end (ref: dsp/frame/FrameQueue.cc:9-115,
dsp/frame/Indexing.h:18-70, tests/frame_queue_test.cc).

Bars: frame extraction and frame counts BIT-EXACT against the oracle (which is pinned bit-for-bit
to the compiled reference, tests/test_oracle_vs_ref.py); kernel 1 in FrameQueue addressing mode ==
kernel 1 run over the explicitly padded signal, BIT-IDENTICAL spectra, and within 1e-5 rel-L2 of the
oracle's rfft of the oracle's frames."""
import numpy as np
import pyoracle as po
import pytest
from conftest import bits, rel_l2

pytestmark = pytest.mark.gpu
TOL = 1e-5
MODES = [po.PAD_CONSTANT, po.PAD_REFLECT, po.PAD_EDGE]


def dev(t, a):
    return t.from_numpy(np.ascontiguousarray(a)).cuda()


@pytest.mark.parametrize("mode", MODES)
@pytest.mark.parametrize("center", [True, False])
def test_extract_is_bit_exact(cuda, dsp, oracle, mode, center):
    rng = np.random.default_rng(17 + mode)
    for L in (0, 1, 2, 3, 5, 31, 32, 33, 100, 1000, 4099):
        for N, H in ((8, 4), (32, 16), (32, 32), (16, 4), (256, 64), (9, 3), (31, 15), (1024, 256)):
            x = rng.standard_normal(L).astype(np.float32)
            want = oracle.framequeue(x, N, H, center, mode)
            assert dsp.framequeue_count(L, N, H, center) == want.shape[0]
            got = dsp.framequeue_extract(dev(cuda, x) if L else cuda.empty(0, device="cuda"), N, H, center, mode)
            assert got.shape == want.shape
            assert np.array_equal(bits(got.cpu().numpy()), bits(want)), (L, N, H)


def test_reference_known_answers(cuda, dsp):
    """tests/frame_queue_test.cc:321-364 pattern and the asymmetric reflect101 (Indexing.h:23-29)."""
    x = dev(cuda, np.array([1, 2, 3, 4], np.float32))
    got = dsp.framequeue_extract(x, 8, 4, True, dsp.PAD_REFLECT).cpu().numpy()
    assert got.tolist() == [[4, 3, 2, 1, 1, 2, 3, 4], [1, 2, 3, 4, 3, 2, 1, 1]]
    got = dsp.framequeue_extract(x, 8, 4, True, dsp.PAD_EDGE).cpu().numpy()
    assert got.tolist() == [[1, 1, 1, 1, 1, 2, 3, 4], [1, 2, 3, 4, 4, 4, 4, 4]]
    got = dsp.framequeue_extract(x, 8, 4, True, dsp.PAD_CONSTANT).cpu().numpy()
    assert got.tolist() == [[0, 0, 0, 0, 1, 2, 3, 4], [1, 2, 3, 4, 0, 0, 0, 0]]
    with pytest.raises(dsp.CrlotInvalidArgument):
        dsp.framequeue_count(10, 0, 4)
    with pytest.raises(dsp.CrlotInvalidArgument):
        dsp.framequeue_count(10, 8, 0)


@pytest.mark.parametrize("N,H", [(256, 64), (512, 128), (1024, 256), (2048, 512), (4096, 1024), (1024, 512), (2048, 2048)])
@pytest.mark.parametrize("mode", MODES)
def test_kernel1_framequeue_mode(cuda, dsp, oracle, N, H, mode):
    w = oracle.window(po.HANN, N)
    plan = dsp.Plan(N, H, analysis_window=w, ola_window=w, boundary_mode=dsp.DROP)
    rng = np.random.default_rng(N + H + mode)
    for ch, L in ((1, 3), (1, N // 2 - 4), (2, 5 * N + 12), (3, 70 * H + 40), (1, 300 * H + 4)):
        x = rng.standard_normal((ch, L)).astype(np.float32)
        Lp = (L + 3) // 4 * 4                                   # row stride must be a multiple of 4 floats
        xd = cuda.zeros((ch, Lp), dtype=cuda.float32, device="cuda")
        xd[:, :L] = dev(cuda, x)
        got = plan.stft_framequeue(xd, True, mode, length=L)
        F = dsp.framequeue_count(L, N, H, True)
        assert got.shape == (ch, F, N // 2 + 1)
        # (a) against the oracle: rfft of the oracle's FrameQueue frames times the window
        for c in range(ch):
            fr = oracle.framequeue(x[c], N, H, True, mode)
            want = oracle.rfft_many(fr * w[None, :], N)
            assert rel_l2(got[c].cpu().numpy(), want) < TOL, (ch, L, c)
        # (b) bit-identical to the plain kernel over the explicitly padded signal (DROP framing)
        pad = N // 2
        xp = np.stack([np.concatenate([[oracle_pad(x[c], i, mode) for i in range(-pad, 0)], x[c],
                                       [oracle_pad(x[c], i, mode) for i in range(L, L + pad)]]).astype(np.float32)
                       for c in range(ch)])
        P = L + 2 * pad
        Pp = (P + 3) // 4 * 4
        xpd = cuda.zeros((ch, Pp), dtype=cuda.float32, device="cuda")
        xpd[:, :P] = dev(cuda, xp)
        ref = plan.stft(xpd, length=P, n_frames=F)
        cuda.cuda.synchronize()
        assert np.array_equal(bits(got.cpu().numpy()), bits(ref.cpu().numpy())), (ch, L)


def oracle_pad(x, i, mode):
    n = len(x)
    if 0 <= i < n:
        return x[i]
    if mode == po.PAD_CONSTANT:
        return 0.0
    if mode == po.PAD_EDGE:
        return x[0] if i < 0 else x[n - 1]
    if n <= 1:
        return x[0]
    while i < 0 or i >= n:
        i = -i - 1 if i < 0 else 2 * n - 2 - i
    return x[i]


def test_uncentered_equals_plain_drop_framing(cuda, dsp, oracle):
    N, H, L = 1024, 256, 50 * 256 + 1024 + 8
    w = oracle.window(po.HAMMING, N)
    plan = dsp.Plan(N, H, analysis_window=w, boundary_mode=dsp.DROP)
    x = dsp.synth_uniform(2, L, seed=5)
    a = plan.stft_framequeue(x, False, dsp.PAD_REFLECT)
    b = plan.stft(x)
    cuda.cuda.synchronize()
    assert a.shape == b.shape and np.array_equal(bits(a.cpu().numpy()), bits(b.cpu().numpy()))


@pytest.mark.parametrize("N,H", [(1024, 256), (2048, 512), (4096, 1024), (2048, 2048), (480, 160)])
@pytest.mark.parametrize("mode", MODES)
def test_host_entry_equals_device_entry(cuda, dsp, oracle, N, H, mode):
    """crlot_stft_host_framequeue (slabs through the host-path slots) == crlot_stft_execute_framequeue on the whole
    signal, bit for bit: short signals (one slab with both ends), both centre settings, several channels."""
    w = oracle.window(po.HANN, N)
    plan = dsp.Plan(N, H, analysis_window=w, ola_window=w, boundary_mode=dsp.DROP)
    rng = np.random.default_rng(3 * N + H + mode)
    for ch, L, center in ((1, 5, True), (1, N // 2 + 3, True), (2, 5 * N + 12, True), (3, 70 * H + 41, True), (2, 9 * N + 7, False)):
        if not center and L < N:
            continue
        x = rng.standard_normal((ch, L)).astype(np.float32)
        got = plan.stft_host_framequeue(x, center, mode)
        Lp = (L + 3) // 4 * 4
        xd = cuda.zeros((ch, Lp), dtype=cuda.float32, device="cuda")
        xd[:, :L] = dev(cuda, x)
        want = plan.stft_framequeue(xd, center, mode, length=L).cpu().numpy()
        assert got.shape == want.shape
        assert np.array_equal(bits(got.view(np.float32)), bits(want.view(np.float32))), (ch, L, center)


@pytest.mark.parametrize("N,H,mode", [(2048, 512, po.PAD_REFLECT), (4096, 1024, po.PAD_EDGE), (1024, 256, po.PAD_CONSTANT)])
def test_host_entry_across_slabs(cuda, dsp, oracle, N, H, mode):
    """A signal of several slabs (2^24 samples each): the first slab holds the left pad, interior slabs run the plain
    path with a zero pad offset, the last one the right pad -- and a ragged length leaves a short last slab whose
    reflection reaches back before its own first frame.  Bit-identical to the one-launch device entry."""
    w = oracle.window(po.HANN, N)
    plan = dsp.Plan(N, H, analysis_window=w, ola_window=w, boundary_mode=dsp.DROP)
    for L in ((1 << 25) + 3 * H + 5, 2 * (1 << 24) + N // 2 - 100):
        x = np.asarray(dsp.synth_uniform(1, L, seed=L & 0xffff).cpu().numpy())
        got = plan.stft_host_framequeue(x, True, mode)
        Lp = (L + 3) // 4 * 4
        xd = cuda.zeros((1, Lp), dtype=cuda.float32, device="cuda")
        xd[:, :L] = dev(cuda, x)
        want = plan.stft_framequeue(xd, True, mode, length=L)
        gd = dev(cuda, got.view(np.float32))
        assert cuda.equal(gd.view(cuda.int32).reshape(-1), cuda.view_as_real(want).contiguous().view(cuda.int32).reshape(-1)), L
        del xd, want, gd
