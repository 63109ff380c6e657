"""Drop-in surfaces over the C-ABI: IFftPlan (batch/stride layout, argument errors) and
OLAAccumulator (bookkeeping, window policies, AoS == SoA), mirroring the reference's own tests
(tests/fft_test.cc, tests/ola_accumulator_test.cc)."""
import numpy as np
import pyoracle as po
import pytest
from conftest import bits, rel_l2

pytestmark = pytest.mark.gpu
TOL = 1e-5


def test_fft_plan_known_answers(cuda, dsp, oracle):
    for n in (512, 1024, 2048):
        f = dsp.FftPlan(n)
        X = f.forward(np.ones(n, np.float32))
        assert abs(X[0].real - n) < 1e-3 and X[0].imag == 0 and np.abs(X[1:]).max() < 1e-3      # fft_test.cc:131-149
        t = np.arange(n)
        X = f.forward((2 * np.cos(2 * np.pi * 10 * t / n)).astype(np.float32))
        assert abs(abs(X[10]) - n) < 1e-2 * n / 512 and abs(np.angle(X[10])) < 1e-4             # :151-173
        X = f.forward(((-1.0) ** t).astype(np.float32))
        assert abs(X[n // 2].real - n) < 1e-3 and X[n // 2].imag == 0                           # :364-378
        x = np.random.default_rng(n).uniform(-1, 1, n).astype(np.float32)
        y = f.inverse(f.forward(x))
        assert np.sqrt(np.mean((y - x) ** 2)) < 1e-5                                            # :57-129
        assert f.size() == n and f.max_batch_size() == 16


def test_fft_plan_batch_and_stride_layout(cuda, dsp, oracle, gold):
    g = gold("fft_kat")
    f = dsp.FftPlan(512, batch=4, stride_in=2, stride_out=2)
    Xb = f.forward(g["xb"], batch=4, out_len=4 * 2 * 257)
    mask = np.abs(g["Xb"]) > 0
    assert rel_l2(Xb[mask], g["Xb"][mask]) < TOL and not Xb[~mask].any()     # untouched slots stay untouched
    f1 = dsp.FftPlan(512, batch=4)
    x = np.random.default_rng(4).uniform(-1, 1, 4 * 512).astype(np.float32)
    X = f1.forward(x, batch=4)
    assert rel_l2(X, oracle.rfft(x, 512, batch=4)) < TOL
    assert rel_l2(f1.inverse(X, batch=4), oracle.irfft(oracle.rfft(x, 512, batch=4), 512, batch=4)) < TOL


def test_fft_plan_argument_errors(cuda, dsp):
    # kissfft_adapter.cc:13-39 -> std::runtime_error
    for kw in (dict(batch=17), dict(batch=0), dict(stride_in=0), dict(in_place=True)):
        with pytest.raises(dsp.CrlotError):
            dsp.FftPlan(512, **kw)
    with pytest.raises(dsp.CrlotError):
        dsp.FftPlan(511)
    f = dsp.FftPlan(512, batch=2)
    with pytest.raises(dsp.CrlotError):
        f.forward(np.zeros(3 * 512, np.float32), batch=3)       # batch > plan batch: "Invalid batch size"


@pytest.mark.parametrize("N,H,Cn", [(256, 64, 1), (1024, 256, 2), (512, 512, 1)])
def test_ola_accumulator_streaming(cuda, dsp, oracle, N, H, Cn):
    rng = np.random.default_rng(N + Cn)
    F = 30
    fr = rng.uniform(-1, 1, (F, Cn, N)).astype(np.float32)
    w = oracle.window(po.HANN, N)
    want, n_want, peak = oracle.ola_stream(fr, N, H, Cn, w, True, False, True, gain=0.9)
    ola = dsp.OlaAccumulator(N, H, Cn)
    assert ola.ring_size() == oracle.ring_len(N, H) and not ola.has_window()
    ola.set_window(w)
    assert ola.has_window()
    got = [[] for _ in range(Cn)]
    for k in range(F):
        ola.add_frame_SoA([fr[k, c] for c in range(Cn)], None, k * H, 0, N, 0.9)
        assert ola.produced_samples() == k * H + N                               # ola_accumulator_test.cc:194-219
        outs, n = ola.produce(H)
        assert n == H
        for c in range(Cn):
            got[c].append(outs[c])
    if N > H:
        outs, n = ola.produce(N - H)
        for c in range(Cn):
            got[c].append(outs[c])
    got = np.stack([np.concatenate(g) for g in got])
    assert np.array_equal(bits(got), bits(want))
    assert abs(ola.meter_peak() - peak) == 0
    ola.reset()
    assert ola.produced_samples() == 0 and ola.read_pos() == 0 and not ola.has_window()


def test_ola_accumulator_aos_equals_soa_and_policies(cuda, dsp, oracle):
    rng = np.random.default_rng(8)
    N, H, Cn, F = 256, 64, 3, 20
    fr = rng.uniform(-1, 1, (F, Cn, N)).astype(np.float32)
    w = oracle.window(po.HAMMING, N)
    a = dsp.OlaAccumulator(N, H, Cn)
    b = dsp.OlaAccumulator(N, H, Cn)
    a.set_window(w)
    b.set_window(w)
    for k in range(F):
        a.add_frame_SoA([fr[k, c] for c in range(Cn)], None, k * H, 0, N, 1.3)
        b.push_frame_AoS(np.ascontiguousarray(fr[k].T), None, k * H, 0, N, 1.3)
        oa, _ = a.produce(H)
        ob, _ = b.produce(H)
        for c in range(Cn):
            assert np.array_equal(bits(oa[c]), bits(ob[c]))
    # external-window policy: apply_window_inside=False honours the per-call window, envelope = 1
    want, _, _ = oracle.ola_stream(fr, N, H, Cn, w, True, True, False, gain=1.0)
    c_ = dsp.OlaAccumulator(N, H, Cn, apply_window_inside=False)
    c_.set_window(w)
    outs = []
    for k in range(F):
        c_.add_frame_SoA([fr[k, c] for c in range(Cn)], w, k * H, 0, N, 1.0)
        outs.append(np.stack(c_.produce(H)[0]))
    assert np.array_equal(bits(np.concatenate(outs, axis=1)), bits(want[:, : F * H]))
    # argument errors -> std::invalid_argument
    with pytest.raises(ValueError):
        dsp.OlaAccumulator(0, 64)
    with pytest.raises(ValueError):
        a.set_window(w[:100])
    # produce with nothing accumulated returns 0 samples
    e = dsp.OlaAccumulator(N, H, 1)
    assert e.produce(64)[1] == 0
    e.flush()
    assert e.produce(64)[1] == 64
