"""crlot_roundtrip_execute_static: the round trip on the library's static device scratch (include/crlot_cuda.h, last
section; no counterpart in the reference -- added for callers that cannot call cudaMalloc).  Called through the C-ABI
and compared with the existing entry point crlot_roundtrip_execute on the same inputs: same kernels, same roundings,
so the comparison is bit for bit (and within the 1e-5 rel-L2 bar of the oracle, like every round-trip test here).

Two calls in a row on different inputs share the scratch -- the second must not see the first's spectra or counter --
then a plan on the unfused route (iFFT frames in the scratch as well), the size limit, and one replay from a CUDA graph.
This file sorts last on purpose."""
import ctypes as C

import numpy as np
import pyoracle as po
import pytest
from conftest import bits, rel_l2

pytestmark = pytest.mark.gpu
TOL = 1e-5
SCRATCH = 32 << 20          # CRLOT_STATIC_SCRATCH_BYTES


def _existing(plan, x):
    out, _, _ = plan.roundtrip(x)
    return out


def _static(cuda, plan, x):
    ch = 1 if x.dim() == 1 else x.shape[0]
    out = cuda.full((ch, plan.out_len(x.shape[-1])), float("nan"), dtype=cuda.float32, device="cuda")
    return plan.roundtrip_static(x, out)


@pytest.mark.parametrize("N,H", [(2048, 512), (1024, 256), (4096, 1024)])
def test_two_calls_in_a_row_match_existing_entry_point(cuda, dsp, oracle, N, H):
    w = oracle.window(po.HANN, N)
    plan = dsp.Plan(N, H, analysis_window=w, ola_window=w, boundary_mode=dsp.DROP)
    first = dsp.synth_uniform(2, 150 * H + N, seed=101)
    second = dsp.synth_uniform(1, 37 * H + N + 12, seed=202)[0]      # smaller and ragged: stale rows lie behind it
    for x in (first, second, first):
        want = _existing(plan, x)
        got = _static(cuda, plan, x)
        cuda.cuda.synchronize()
        want, got = want.cpu().numpy(), got.cpu().numpy()
        assert not np.isnan(got).any()
        assert rel_l2(got, want) < TOL
        assert np.array_equal(bits(got), bits(want))
    ref = oracle.roundtrip(second.cpu().numpy(), N, H, w, po.DROP)
    got = _static(cuda, plan, second)[0].cpu().numpy()
    assert rel_l2(got, ref["out"]) < TOL


def test_unfused_route_keeps_frames_in_the_scratch(cuda, dsp, oracle):
    """nfft 480 / hop 160 runs the any-size kernels: K1 -> K2 -> K3 with spectra AND iFFT frames in the static array."""
    N, H = 480, 160
    w = oracle.window(po.HAMMING, N)
    plan = dsp.Plan(N, H, analysis_window=w, ola_window=w, boundary_mode=dsp.ZERO_PAD)
    for seed, L in ((5, 200 * H + 77), (6, 31 * H)):
        x = dsp.synth_uniform(2, L, seed=seed)
        want = _existing(plan, x).cpu().numpy()
        got = _static(cuda, plan, x).cpu().numpy()
        assert np.array_equal(bits(got), bits(want))
        ref = oracle.roundtrip(x[0].cpu().numpy(), N, H, w, po.ZERO_PAD)
        assert rel_l2(got[0], ref["out"]) < TOL


def test_no_device_memory_is_taken_and_the_limit_is_enforced(cuda, dsp, oracle):
    N, H = 2048, 512
    w = oracle.window(po.HANN, N)
    plan = dsp.Plan(N, H, analysis_window=w, ola_window=w, boundary_mode=dsp.DROP)
    x = dsp.synth_uniform(1, 64 * H + N, seed=9)
    out = cuda.empty((1, plan.out_len(x.shape[-1])), dtype=cuda.float32, device="cuda")
    plan.roundtrip_static(x, out)                     # first call on the device resolves the scratch address
    cuda.cuda.synchronize()
    free0, _ = cuda.cuda.mem_get_info()
    for _ in range(3):
        plan.roundtrip_static(x, out)
    cuda.cuda.synchronize()
    free1, _ = cuda.cuda.mem_get_info()
    assert free1 >= free0                             # the calls took no device memory

    bins = N // 2 + 1
    max_rows = (SCRATCH - 256) // (bins * 8)
    big = cuda.zeros(((max_rows + 1 - 1) * H + N,), dtype=cuda.float32, device="cuda")      # max_rows + 1 frames
    big_out = cuda.empty((1, plan.out_len(big.numel())), dtype=cuda.float32, device="cuda")
    assert plan.n_frames(big.numel()) == max_rows + 1
    with pytest.raises(dsp.CrlotInvalidArgument) as err:
        plan.roundtrip_static(big, big_out)
    assert "CRLOT_STATIC_SCRATCH_BYTES" in str(err.value)
    fits = big[:(max_rows - 1) * H + N]               # exactly max_rows frames: accepted
    want = _existing(plan, fits).cpu().numpy()
    got = plan.roundtrip_static(fits, big_out[:, :plan.out_len(fits.numel())].contiguous()).cpu().numpy()
    assert np.array_equal(bits(got), bits(want))


def test_replays_from_a_cuda_graph(cuda, dsp, oracle):
    """What the entry point is for: a capture in the strictest mode fails on any allocation inside it."""
    N, H = 2048, 512
    w = oracle.window(po.HANN, N)
    plan = dsp.Plan(N, H, analysis_window=w, ola_window=w, boundary_mode=dsp.DROP)
    x = dsp.synth_uniform(2, 90 * H + N, seed=77)
    want = _existing(plan, x).cpu().numpy()
    out = cuda.zeros((2, plan.out_len(x.shape[-1])), dtype=cuda.float32, device="cuda")
    plan.roundtrip_static(x, out)                     # warm-up outside the capture
    cuda.cuda.synchronize()
    out.zero_()
    g = cuda.cuda.CUDAGraph()
    side = cuda.cuda.Stream()
    with cuda.cuda.graph(g, stream=side, capture_error_mode="global"):
        plan.roundtrip_static(x, out)
    cuda.cuda.synchronize()
    assert float(out.abs().sum()) == 0.0              # captured, not run
    for _ in range(2):
        out.zero_()
        g.replay()
        cuda.cuda.synchronize()
        assert np.array_equal(bits(out.cpu().numpy()), bits(want))
