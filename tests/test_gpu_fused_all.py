"""The one-kernel round trip (crlot_roundtrip_fused_execute, csrc/k23_istft_ola.cu FW variant): frame + window +
rFFT + [spectral hook] + irFFT + synthesis window + overlap-add + normalise in one launch, the spectra living only in
registers and the overlapping part of consecutive frames in the lanes' tensor memory.
Bar: BIT-IDENTICAL to K1 -> K2 -> K3 (same roundings in the same order), hence the same 1e-5 rel-L2 to the oracle:
channels, ZERO_PAD tails, ragged lengths, chunk boundaries, time-shard offsets, the spectral hook, the clean-up rule."""
import numpy as np
import pyoracle as po
import pytest
from conftest import bits, rel_l2

pytestmark = pytest.mark.gpu
TOL = 1e-5
N, H = 2048, 512


def test_supported_shapes(cuda, dsp, oracle):
    w = oracle.window(po.HANN, N)
    assert dsp.Plan(N, H, analysis_window=w, ola_window=w).roundtrip_fused_all_supported()
    assert not dsp.Plan(N, 256, analysis_window=w, ola_window=w).roundtrip_fused_all_supported()
    w1 = oracle.window(po.HANN, 1024)
    p = dsp.Plan(1024, 256, analysis_window=w1, ola_window=w1)
    assert not p.roundtrip_fused_all_supported()
    with pytest.raises(dsp.CrlotUnsupported):
        p.roundtrip_fused_all(dsp.synth_uniform(1, 1 << 14))


@pytest.mark.parametrize("mode", [po.DROP, po.ZERO_PAD])
@pytest.mark.parametrize("ch,L", [(1, N), (1, N + 3 * H), (2, 100 * H + N), (3, 400 * H + N + 37 * 4), (1, 5000 * H + 100)])
def test_one_kernel_equals_three_kernels_and_oracle(cuda, dsp, oracle, mode, ch, L):
    w = oracle.window(po.HANN, N)
    plan = dsp.Plan(N, H, analysis_window=w, ola_window=w, boundary_mode=mode, gain=0.9)
    L4 = (L + 3) // 4 * 4
    x = dsp.synth_uniform(ch, L4, seed=L)[:, :L] if L4 != L else dsp.synth_uniform(ch, L, seed=L)
    if x.stride(0) % 4 or not x.is_contiguous():
        buf = cuda.zeros((ch, L4), dtype=cuda.float32, device="cuda")
        buf[:, :L] = x
        x = buf
    want, _, _ = plan.roundtrip(x, length=L)
    got = plan.roundtrip_fused_all(x, length=L)
    cuda.cuda.synchronize()
    assert got.shape == want.shape
    assert np.array_equal(bits(got.cpu().numpy()), bits(want.cpu().numpy()))
    ref = oracle.roundtrip(x[ch - 1, :L].cpu().numpy(), N, H, w, mode, gain=0.9)
    assert rel_l2(got[ch - 1].cpu().numpy(), ref["out"]) < TOL


def test_one_kernel_with_hook_offsets_and_cleanup(cuda, dsp, oracle):
    rng = np.random.default_rng(4)
    w = oracle.window(po.HAMMING, N)
    L = 700 * H + N
    xh = oracle.synth_uniform(9, 0, 0, L)
    xh[12345] = np.nan                       # the clean-up rule on the analysis side (kissfft_adapter.cc:104-108)
    xh[200000] = np.inf
    xh[300001] = 1e-36
    x = cuda.from_numpy(xh).cuda().unsqueeze(0)
    plan = dsp.Plan(N, H, analysis_window=w, ola_window=w, boundary_mode=po.DROP)
    F = plan.n_frames(L)
    # time-frequency mask, then an equaliser table
    for gain in (cuda.from_numpy(rng.uniform(0, 2, (1, F, N // 2 + 1)).astype(np.float32)).cuda(),
                 cuda.from_numpy(rng.uniform(0, 2, N // 2 + 1).astype(np.float32)).cuda()):
        plan.set_spectral_gain(gain)
        want, _, _ = plan.roundtrip(x)
        got = plan.roundtrip_fused_all(x)
        cuda.cuda.synchronize()
        assert np.array_equal(bits(got.cpu().numpy()), bits(want.cpu().numpy()))
    plan.set_spectral_gain(None)
    # a shard-like call: left-halo frames (frame_t0 < 0), envelope phase, ragged n_out
    spec = plan.stft(x)
    for t_off, ft0, n_out in ((6 * H, -3 * H, 300 * H), (123456 * 2, -2 * H, 40 * H + 6), (0, 0, (F - 1) * H + N), (10, 2 * H, 50 * H)):
        want = plan.istft_ola(spec, t_offset=t_off, frame_t0=ft0, n_out=n_out)
        got = plan.roundtrip_fused_all(x, t_offset=t_off, frame_t0=ft0, n_out=n_out)
        cuda.cuda.synchronize()
        assert np.array_equal(bits(got.cpu().numpy()), bits(want.cpu().numpy())), (t_off, ft0, n_out)


def test_one_kernel_full_size_properties(cuda, dsp, oracle):
    """2^28 samples: bit-identical to K1 + fused K2+K3, and a slice against the oracle deep inside the signal."""
    w = oracle.window(po.HANN, N)
    plan = dsp.Plan(N, H, analysis_window=w, ola_window=w, boundary_mode=po.DROP)
    L = 1 << 28
    x = dsp.synth_uniform(1, L, seed=0xC0FFEE)
    want, _ = plan.roundtrip_fused(x)
    got = plan.roundtrip_fused_all(x)
    cuda.cuda.synchronize()
    assert cuda.equal(got, want)
    a, ref = oracle.slice_roundtrip(0xC0FFEE, 0, 200_000_000, 1 << 15, N, H, w, plan.ring_len)
    assert rel_l2(got[0, a:a + (1 << 15)].cpu().numpy(), ref) < TOL
