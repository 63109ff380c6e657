"""Interleaved (AoS) multichannel I/O (ref: dsp/ola/aos_to_soa.cc:5-19,
dsp/frame/framer.cc:133-141): device de/interleave copies are bit-exact, and the interleaved
host-buffer round trip equals the planar one bit for bit (hence the same parity to the oracle)."""
import numpy as np
import pyoracle as po
import pytest
from conftest import bits, rel_l2

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("ch", [1, 2, 3, 4, 7])
def test_device_transposes_are_exact(cuda, dsp, ch):
    rng = np.random.default_rng(ch)
    for n in (1, 5, 1000, 65537):
        a = rng.standard_normal((n, ch)).astype(np.float32)
        ad = cuda.from_numpy(a).cuda()
        cs = (n + 3) // 4 * 4
        pl = dsp.deinterleave(ad, cs)
        assert np.array_equal(bits(pl.cpu().numpy()[:, :n]), bits(np.ascontiguousarray(a.T)))
        back = dsp.interleave(pl, n)
        assert np.array_equal(bits(back.cpu().numpy()), bits(a))
        # unaligned interleaved base (odd element offset): the scalar path
        if n > 3:
            ad2 = cuda.from_numpy(np.concatenate([[0.0] * 1, a.ravel()]).astype(np.float32)).cuda()[1:].view(n, ch)
            assert np.array_equal(bits(dsp.deinterleave(ad2, cs).cpu().numpy()[:, :n]), bits(np.ascontiguousarray(a.T)))


@pytest.mark.parametrize("N,H,ch", [(1024, 256, 2), (2048, 512, 2), (256, 64, 4), (4096, 1024, 2), (1024, 256, 3), (512, 96, 2)])
def test_interleaved_host_roundtrip_equals_planar(cuda, dsp, oracle, N, H, ch):
    L = 77 * H + N + 5
    rng = np.random.default_rng(N + ch)
    x = rng.uniform(-1, 1, (L, ch)).astype(np.float32)
    w = oracle.window(po.HANN, N)
    plan = dsp.Plan(N, H, analysis_window=w, ola_window=w, boundary_mode=dsp.DROP)
    y = plan.roundtrip_host_interleaved(x)
    yp = plan.roundtrip_host(np.ascontiguousarray(x.T))
    assert y.shape == (yp.shape[1], ch)
    assert np.array_equal(bits(np.ascontiguousarray(y.T)), bits(yp))
    ref = oracle.roundtrip(np.ascontiguousarray(x[:, ch - 1]), N, H, w, po.DROP)
    assert rel_l2(y[:, ch - 1], ref["out"]) < 1e-5


def test_interleaved_host_roundtrip_many_slabs(cuda, dsp, oracle):
    """Longer than one slab (2^24 / hop / channels hops per slab): slab seams must be invisible."""
    N, H, ch = 1024, 256, 2
    L = (1 << 24) + 3 * H + 17
    x = np.stack([oracle.synth_uniform(7, c, 0, L) for c in range(ch)], axis=1)
    w = oracle.window(po.HANN, N)
    plan = dsp.Plan(N, H, analysis_window=w, ola_window=w, boundary_mode=dsp.DROP)
    y = plan.roundtrip_host_interleaved(np.ascontiguousarray(x))
    yp = plan.roundtrip_host(np.ascontiguousarray(x.T))
    assert np.array_equal(bits(np.ascontiguousarray(y.T)), bits(yp))
