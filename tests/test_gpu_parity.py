"""GPU parity: the CUDA path, called through the C-ABI, against the CPU oracle and the golden
vectors.  Bars (BASELINE.json north_star): frame indexing bit-exact; spectra and audio within
1e-5 relative L2 (fp32); kernel 3 bit-identical to the reference given identical frames."""
import numpy as np
import pyoracle as po
import pytest
from conftest import bits, rel_l2

pytestmark = pytest.mark.gpu
TOL = 1e-5            # north_star: "within 1e-5 relative L2 (fp32)"
SIZES = [(256, 64, po.HAMMING), (512, 128, po.HANN), (1024, 256, po.HANN), (2048, 512, po.HANN), (4096, 1024, po.HANN)]


def dev(t, a):
    return t.from_numpy(np.ascontiguousarray(a)).cuda()


@pytest.mark.parametrize("N,H,wt", SIZES)
def test_k1_framing_is_bit_exact(cuda, dsp, oracle, N, H, wt):
    """Frame indexing: a one-hot analysis window w = e_i turns bin 0 of frame k into x[k*H+i]
    exactly (a sum of one non-zero term), for every frame including the zero-padded tail."""
    rng = np.random.default_rng(N)
    L = 37 * H + N + 3 * H // 4          # ragged end: ZERO_PAD adds partly-empty frames
    x = rng.uniform(-1, 1, L).astype(np.float32)
    for mode in (dsp.DROP, dsp.ZERO_PAD):
        frames = oracle.framer(x, N, H, 1, mode)
        for i in (0, 1, N // 2 - 1, N // 2, N - 2, N - 1, 2 * (N // 7) + 1):
            w = np.zeros(N, np.float32)
            w[i] = 1.0
            plan = dsp.Plan(N, H, analysis_window=w, boundary_mode=mode)
            S = plan.stft(dev(cuda, x)).cpu().numpy()[0]
            assert S.shape[0] == frames.shape[0] == oracle.frame_count(L, N, H, mode)
            assert np.array_equal(bits(S[:, 0].real.copy()), bits(frames[:, i].copy())), (mode, i)
            assert not S[:, 0].imag.any()


@pytest.mark.parametrize("N,H,wt", SIZES)
def test_k1_spectra_vs_oracle(cuda, dsp, oracle, N, H, wt):
    rng = np.random.default_rng(N + 1)
    L = 203 * H + N + 5
    x = rng.uniform(-1, 1, L).astype(np.float32)
    w = oracle.window(wt, N)
    for mode in (dsp.DROP, dsp.ZERO_PAD):
        plan = dsp.Plan(N, H, analysis_window=w, ola_window=w, boundary_mode=mode)
        S = plan.stft(dev(cuda, x)).cpu().numpy()[0]
        ref = oracle.rfft_many(oracle.framer(x, N, H, 1, mode) * w, N)
        assert S.shape == ref.shape
        assert rel_l2(S, ref) < TOL
        worst = max(rel_l2(S[k], ref[k]) for k in range(S.shape[0]))
        assert worst < TOL, worst
        assert not S[:, 0].imag.any() and not S[:, -1].imag.any()      # kiss_fftr.c:93-99: exact zeros


@pytest.mark.parametrize("N,H,wt", SIZES)
def test_k2_frames_vs_oracle(cuda, dsp, oracle, N, H, wt):
    rng = np.random.default_rng(N + 2)
    F = 77
    X = (rng.standard_normal((F, N // 2 + 1)) + 1j * rng.standard_normal((F, N // 2 + 1))).astype(np.complex64)
    w = oracle.window(wt, N)
    plan = dsp.Plan(N, H, analysis_window=w, ola_window=w)
    y = plan.istft_frames(dev(cuda, X)).cpu().numpy()
    ref = oracle.irfft_many(X, N)
    assert rel_l2(y, ref) < TOL
    assert max(rel_l2(y[k], ref[k]) for k in range(F)) < TOL
    # fused synthesis window == one extra fp32 rounding of y*w
    yw = plan.istft_frames(dev(cuda, X), fuse_synthesis_window=True).cpu().numpy()
    assert np.array_equal(bits(yw), bits(y * w))


@pytest.mark.parametrize("N,H,wt", SIZES)
@pytest.mark.parametrize("channels", [1, 3])
def test_k3_is_bit_identical_given_identical_frames(cuda, dsp, oracle, N, H, wt, channels):
    rng = np.random.default_rng(N + 3)
    F = 61
    fr = rng.uniform(-1, 1, (F, channels, N)).astype(np.float32)
    w = oracle.window(wt, N)
    for gain in (1.0, 0.7):
        want, n, _ = oracle.ola_stream(fr, N, H, channels, w, True, False, True, gain=gain)
        plan = dsp.Plan(N, H, ola_window=w, gain=gain)
        assert np.array_equal(bits(plan.envelope()), bits(oracle.ola_envelope(w, N, H)))
        dfr = dev(cuda, fr.transpose(1, 0, 2))
        got = plan.ola(dfr).cpu().numpy()
        assert got.shape == want.shape
        assert np.array_equal(bits(got), bits(want)), (gain,)
        # pre-windowed frames (what kernel 2 hands over) give the same bits
        pre = (fr * w).astype(np.float32)
        got2 = plan.ola(dev(cuda, pre.transpose(1, 0, 2)), prewindowed=True).cpu().numpy()
        assert np.array_equal(bits(got2), bits(want))
    # window policies of add_frame_SoA (OLAAccumulator.cc:82-83) and envelope special cases
    for inside, setw, passw in [(False, True, True), (False, False, False), (True, False, False)]:
        want, _, _ = oracle.ola_stream(fr, N, H, channels, w, setw, passw, inside, gain=1.0)
        plan = dsp.Plan(N, H, ola_window=w if setw else None, push_window=w if passw else None, apply_window_inside=inside)
        got = plan.ola(dev(cuda, fr.transpose(1, 0, 2))).cpu().numpy()
        assert np.array_equal(bits(got), bits(want)), (inside, setw, passw)


def test_k3_no_overlap_identity(cuda, dsp, oracle):
    # tests/ola_accumulator_test.cc:729-735: H == N -> output == frame (w/max(w,eps) = 1 where w > eps)
    N = H = 256
    rng = np.random.default_rng(9)
    fr = rng.uniform(-1, 1, (9, 1, N)).astype(np.float32)
    w = oracle.window(po.HAMMING, N)
    want, _, _ = oracle.ola_stream(fr, N, H, 1, w, True, False, True)
    plan = dsp.Plan(N, H, ola_window=w)
    got = plan.ola(dev(cuda, fr.transpose(1, 0, 2))).cpu().numpy()
    assert np.array_equal(bits(got), bits(want))
    assert np.abs(got[0] - fr[:, 0, :].ravel()).max() < 1e-6


@pytest.mark.parametrize("tag", ["rt256", "rt256zp", "rt1024", "rt2048", "rt4096"])
def test_roundtrip_vs_golden(cuda, dsp, oracle, gold, tag):
    g = gold("roundtrip_small")
    N, H, L, t, mode = [int(v) for v in g[f"{tag}_cfg"]]
    w = oracle.window(t, N)
    plan = dsp.Plan(N, H, analysis_window=w, ola_window=w, boundary_mode=mode)
    out, spec, frames = plan.roundtrip(dev(cuda, g[f"{tag}_x"]))
    spec, frames, out = spec.cpu().numpy()[0], frames.cpu().numpy()[0], out.cpu().numpy()[0]
    assert spec.shape == g[f"{tag}_spectra"].shape and out.shape == g[f"{tag}_out"].shape
    assert rel_l2(spec, g[f"{tag}_spectra"]) < TOL
    assert rel_l2(frames, g[f"{tag}_iframes"] * w) < TOL
    assert rel_l2(out, g[f"{tag}_out"]) < TOL
    # same answer through the host-buffer entry point
    out_h = plan.roundtrip_host(g[f"{tag}_x"])
    assert np.array_equal(bits(out_h[0]), bits(out))


def test_config1_oboe(cuda, dsp, oracle, gold):
    g = gold("oboe_c1")
    x = g["lr_sum"].astype(np.float32) / 65536.0
    w = oracle.window(po.HANN, 1024)
    plan = dsp.Plan(1024, 256, analysis_window=w, ola_window=w, boundary_mode=dsp.DROP)
    out, spec, _ = plan.roundtrip(dev(cuda, x))
    out, spec = out.cpu().numpy()[0], spec.cpu().numpy()[0]
    assert spec.shape[0] == 1111 and out.size == int(g["out_len"])
    ref = oracle.roundtrip(x, 1024, 256, w, po.DROP)
    assert rel_l2(spec, ref["spectra"]) < TOL and rel_l2(out, ref["out"]) < TOL
    assert rel_l2(out[:16384], g["out_head"]) < TOL and rel_l2(out[-4096:], g["out_tail"]) < TOL
    assert rel_l2(spec[g["spec_pick_idx"]], g["spec_pick"]) < TOL
    nb = g["out_block_energy"].size
    e = (out[: nb * 4096].astype(np.float64) ** 2).reshape(nb, 4096).sum(1)
    assert np.allclose(e, g["out_block_energy"], rtol=1e-4, atol=1e-9)
    n = out.size
    snr = 10 * np.log10(np.sum(x[1024:n - 1024].astype(np.float64) ** 2) / np.sum((x[1024:n - 1024].astype(np.float64) - out[1024:n - 1024]) ** 2))
    assert abs(snr - float(g["snr_db_vs_input"])) < 0.01        # the reference's own ~7.3 dB


def test_multichannel_and_ragged(cuda, dsp, oracle):
    """Planar [channels][L] batch with a row stride larger than L; every channel equals its own
    mono run; last chunk per channel is ragged."""
    rng = np.random.default_rng(21)
    N, H, Cn, L = 1024, 256, 5, 70 * 256 + 1024 + 8
    w = oracle.window(po.HANN, N)
    x = rng.uniform(-1, 1, (Cn, L)).astype(np.float32)
    plan = dsp.Plan(N, H, analysis_window=w, ola_window=w)
    big = cuda.zeros((Cn, L + 64), dtype=cuda.float32, device="cuda")
    big[:, :L] = dev(cuda, x)
    out, spec, _ = plan.roundtrip(big, length=L)
    out, spec = out.cpu().numpy(), spec.cpu().numpy()
    for c in range(Cn):
        ref = oracle.roundtrip(x[c], N, H, w, po.DROP)
        assert rel_l2(spec[c], ref["spectra"]) < TOL and rel_l2(out[c], ref["out"]) < TOL
    out1, _, _ = plan.roundtrip(dev(cuda, x[2]))
    assert np.array_equal(bits(out1.cpu().numpy()[0]), bits(out[2]))     # deterministic, channel-independent


def test_edge_cases(cuda, dsp, oracle):
    N, H = 1024, 256
    w = oracle.window(po.HANN, N)
    plan = dsp.Plan(N, H, analysis_window=w, ola_window=w, boundary_mode=dsp.DROP)
    # shorter than one frame: DROP yields nothing, ZERO_PAD one padded frame per started hop
    short = np.linspace(-1, 1, 600).astype(np.float32)
    assert plan.n_frames(600) == 0 and plan.roundtrip_host(short).shape == (1, 0)
    pz = dsp.Plan(N, H, analysis_window=w, ola_window=w, boundary_mode=dsp.ZERO_PAD)
    r = oracle.roundtrip(short, N, H, w, po.ZERO_PAD)
    out, spec, _ = pz.roundtrip(dev(cuda, np.pad(short, (0, 8))), length=600)
    assert spec.shape[1] == r["n_frames"] == 3
    assert rel_l2(out.cpu().numpy()[0], r["out"]) < TOL
    # exactly one frame
    one = np.random.default_rng(1).uniform(-1, 1, N).astype(np.float32)
    r = oracle.roundtrip(one, N, H, w, po.DROP)
    assert rel_l2(plan.roundtrip(dev(cuda, one))[0].cpu().numpy()[0], r["out"]) < TOL
    # silence stays exactly silent
    z, _, _ = plan.roundtrip(cuda.zeros(8192, device="cuda"))
    assert not z.cpu().numpy().any()
    # argument errors mirror the reference's exception types
    with pytest.raises(ValueError):
        dsp.Plan(0, 256)                                    # framer.cc:17-19
    with pytest.raises(ValueError):
        dsp.Plan(1024, 0)                                   # framer.cc:20-22
    with pytest.raises(ValueError):
        dsp.Plan(1024, 256, eps=0.0)                        # OLAConfig::isValid
    with pytest.raises(dsp.CrlotError):
        dsp.Plan(16386, 4096)                               # beyond the GPU path's sizes: error, never a CPU fallback
    with pytest.raises(dsp.CrlotError):
        dsp.Plan(1001, 250)                                 # odd real size (kissfft_adapter.cc:36-39)


def test_scrub_rules(cuda, dsp, oracle, gold):
    """NaN/Inf/|v|<1e-30 clean-up on the forward input and the inverse output, and the ignored
    imaginary parts of DC/Nyquist (kissfft_adapter.cc:104-108,159-163; kiss_fftr.c:131-132)."""
    g = gold("fft_kat")
    f = dsp.FftPlan(512)
    X = f.forward(g["x_scrub"])
    assert np.isfinite(X.view(np.float32)).all() and rel_l2(X, g["X_scrub"]) < TOL
    assert rel_l2(f.inverse(g["X_dcny"]), g["y_dcny"]) < TOL
    y = f.inverse(g["X_nanbin"])
    assert np.array_equal(y, g["y_nanbin"]) and not y.any()       # a NaN bin poisons every sample -> all scrubbed to 0
    tiny = np.full(512, 1e-35, np.float32)
    assert not f.forward(tiny).view(np.float32).any()


def test_large_signal_properties(cuda, dsp, oracle):
    """Sizes the oracle cannot finish in seconds: size-independent properties plus slice oracles
    (a fresh accumulator started at a multiple of ring_len reproduces the full run
    after the first N samples)."""
    N, H = 2048, 512
    L = 1 << 24
    w = oracle.window(po.HANN, N)
    plan = dsp.Plan(N, H, analysis_window=w, ola_window=w)
    x = dsp.synth_uniform(1, L, seed=0xC0FFEE)
    # generator parity (bit-exact) on scattered slices
    for i0 in (0, 12345, L - 4096):
        assert np.array_equal(bits(x[0, i0:i0 + 4096].cpu().numpy()), bits(oracle.synth_uniform(0xC0FFEE, 0, i0, 4096)))
    out, spec, frames = plan.roundtrip(x)
    F = plan.n_frames(L)
    assert spec.shape == (1, F, N // 2 + 1) and out.shape == (1, (F - 1) * H + N)
    # linearity of the whole path: T(a*x) = a*T(x) for a power of two (exact in fp32)
    out2, _, _ = plan.roundtrip(x * 0.5)
    assert cuda.equal(out2, out * 0.5)
    # determinism run to run
    out3, _, _ = plan.roundtrip(x)
    assert cuda.equal(out3, out)
    # slice oracles at ring-aligned offsets
    ring = plan.ring_len
    for blk in (0, 7, (L // ring) - 3):
        s0 = blk * ring
        seg = x[0, s0:s0 + 40 * H + N].cpu().numpy()
        r = oracle.roundtrip(seg, N, H, w, po.DROP)
        lo, hi = (N if blk else 0), 40 * H
        assert rel_l2(out[0, s0 + lo:s0 + hi].cpu().numpy(), r["out"][lo:hi]) < TOL
        k0 = s0 // H
        assert rel_l2(spec[0, k0:k0 + 40].cpu().numpy(), r["spectra"][:40]) < TOL
    # Parseval per frame: sum|x w|^2 = (|X0|^2 + |XN/2|^2 + 2 sum|Xk|^2)/N
    k = 1000
    fr = (x[0, k * H:k * H + N].cpu().numpy() * w).astype(np.float64)
    Xk = spec[0, k].cpu().numpy().astype(np.complex128)
    e = (abs(Xk[0]) ** 2 + abs(Xk[-1]) ** 2 + 2 * (abs(Xk[1:-1]) ** 2).sum()) / N
    assert abs(e - (fr ** 2).sum()) / (fr ** 2).sum() < 1e-5
    # host-buffer path (slabbed, pipelined) == device path
    xh = x[0, : (1 << 22) + 777].cpu().numpy()
    o_dev = plan.roundtrip(dev(cuda, xh))[0].cpu().numpy()
    assert np.array_equal(bits(plan.roundtrip_host(xh)), bits(o_dev))


@pytest.mark.parametrize("N,H", [(2048, 512), (1024, 256), (512, 128), (256, 64), (1024, 512)])
def test_k1_cleanup_inside_a_stream(cuda, dsp, oracle, N, H):
    """NaN / +-Inf / sub-1e-30 samples scattered through a long stream: every frame that contains one
    must come out exactly as the reference's per-sample clean-up makes it (kernel 1 screens samples
    once, on entry, when nfft == 4*hop; the (1024,512) case runs the per-sample rule)."""
    rng = np.random.default_rng(N + H)
    L = 300 * H + N
    x = rng.uniform(-1, 1, L).astype(np.float32)
    w = oracle.window(po.HANN, N)
    bad = {5: np.nan, 3 * H + 17: np.inf, 3 * H + 18: -np.inf, 40 * H: 1e-33, 40 * H + 1: -3e-36, 41 * H - 1: 1e-39,
           100 * H + N // 2: np.nan, 200 * H + 1: 1e-27, 250 * H: 1e30, L - 2: np.inf, 64 * H + 5: 2e-31}
    for i, v in bad.items():
        x[i] = v
    x[150 * H:150 * H + N] = 0.0
    x[150 * H + 7] = 1e-32
    plan = dsp.Plan(N, H, analysis_window=w, ola_window=w)
    S = plan.stft(dev(cuda, x)).cpu().numpy()[0]
    with np.errstate(invalid="ignore", over="ignore"):
        ref = oracle.rfft_many(oracle.framer(x, N, H, 1, po.DROP) * w, N)
    assert np.isfinite(S.view(np.float32)).all()
    for k in range(S.shape[0]):
        den = np.linalg.norm(ref[k])
        err = np.linalg.norm(S[k] - ref[k])
        assert (err / den < TOL) if den > 0 else (err == 0), k
    # the all-zero-but-one-tiny frame comes out exactly zero, as the reference's does
    assert not ref[150].any() and not S[150].any()


@pytest.mark.parametrize("N,H,fused_ok", [(2048, 512, True), (4096, 1024, True)])
def test_full_size_2p30_properties(cuda, dsp, oracle, N, H, fused_ok):
    """BASELINE configs[1] (c2) and one c4 shard at their FULL size (2^30 samples, 43 GB of device
    arrays): size-independent properties and slice oracles at ring-aligned offsets spread over the
    whole signal, including the very end."""
    if cuda.cuda.get_device_properties(0).total_memory < 60e9:
        pytest.skip("needs 60 GB of device memory")
    L = 1 << 30
    w = oracle.window(po.HANN, N)
    plan = dsp.Plan(N, H, analysis_window=w, ola_window=w, boundary_mode=dsp.DROP)
    x = dsp.synth_uniform(1, L, seed=0xC0FFEE)
    F = plan.n_frames(L)
    assert F == (L - N) // H + 1
    out, spec, frames = plan.roundtrip(x)
    assert out.shape == (1, (F - 1) * H + N)
    ring = plan.ring_len
    nblk = L // ring
    for blk in (0, 1, nblk // 3, nblk // 2 + 1, nblk - 2):
        s0 = blk * ring
        n = min(24 * H + N, L - s0)
        seg = oracle.synth_uniform(0xC0FFEE, 0, s0, n)
        assert np.array_equal(bits(x[0, s0:s0 + n].cpu().numpy()), bits(seg))           # generator, bit-exact
        r = oracle.roundtrip(seg, N, H, w, po.DROP)
        lo, hi = (N if blk else 0), (r["n_frames"] - 1) * H
        assert rel_l2(out[0, s0 + lo:s0 + hi].cpu().numpy(), r["out"][lo:hi]) < TOL, blk
        k0 = s0 // H
        assert rel_l2(spec[0, k0:k0 + r["n_frames"]].cpu().numpy(), r["spectra"]) < TOL, blk
    # the tail of the signal: last frames and the final N-H samples of the output
    s0 = (nblk - 1) * ring
    seg = oracle.synth_uniform(0xC0FFEE, 0, s0, L - s0)
    r = oracle.roundtrip(seg, N, H, w, po.DROP)
    assert s0 // H + r["n_frames"] == F
    assert rel_l2(out[0, s0 + N:].cpu().numpy(), r["out"][N:]) < TOL
    del frames
    # determinism and (K2+K3 fused) == (K2 -> K3) over all 2^30 samples, compared on the device
    if fused_ok:
        out_f = plan.istft_ola(spec)
        assert cuda.equal(out_f.view(cuda.int32), out.view(cuda.int32))
        del out_f
    # linearity with a power-of-two factor is exact in fp32
    x.mul_(0.25)
    out2 = plan.roundtrip_fused(x, spec=spec)[0] if fused_ok else plan.roundtrip(x, spec=spec)[0]
    assert cuda.equal(out2, out * 0.25)
