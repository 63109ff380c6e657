"""GPU parity of the streaming path (config 5: many concurrent streams, chunked ring-buffer
ingest) through the C-ABI crlot_stream_* entry points.

Per stream the reference is: Framer(DROP) fed in chunks -> window -> forward -> inverse ->
push_frame_AoS(k*hop) -> produce(hop), final produce(nfft-hop)  (framer.cc:37-182,
bench/ola_benchmark.cc:150-157).  The frames a chunk-fed DROP Framer emits are the frames of the
one-shot Framer (tests/framer_test.cc:299-321 formula), so the concatenated tick outputs must equal
the oracle's round trip of the concatenated input: within 1e-5 rel-L2 (FFT rounding), and
BIT-IDENTICAL to the batch CUDA path (same arithmetic, same order)."""
import numpy as np
import pyoracle as po
import pytest
from conftest import bits, rel_l2

pytestmark = pytest.mark.gpu
TOL = 1e-5


def dev(t, a):
    return t.from_numpy(np.ascontiguousarray(a)).cuda()


def run_ticks(cuda, st, x, chunk):
    """feed [S, L] in chunks of `chunk` with crlot_stream_tick; returns concatenated output"""
    xs = dev(cuda, x)
    outs = []
    for t0 in range(0, x.shape[1] - chunk + 1, chunk):
        o = st.tick(xs[:, t0:t0 + chunk])
        assert o.shape[1] % st.plan.hop == 0
        outs.append(o.cpu().numpy().copy())
    outs.append(st.drain().cpu().numpy().copy())
    return np.concatenate(outs, axis=1)


@pytest.mark.parametrize("N,H,wt,chunk", [(256, 64, po.HAMMING, 64), (256, 64, po.HAMMING, 256), (256, 64, po.HAMMING, 100),
                                          (1024, 256, po.HANN, 256), (512, 128, po.HANN, 36), (2048, 512, po.HANN, 512),
                                          # beyond the ring-reading kernels: nfft = 4096 and any-size plans (gather + kernel 1)
                                          (4096, 1024, po.HANN, 1024), (4096, 1024, po.HANN, 600), (480, 160, po.HANN, 160),
                                          (6000, 1500, po.HAMMING, 1000), (8192, 2048, po.HANN, 2048)])
def test_stream_ticks_equal_batch_and_oracle(cuda, dsp, oracle, N, H, wt, chunk):
    rng = np.random.default_rng(N + chunk)
    S, T = (13, 41) if N <= 2048 else (5, 23)
    L = T * chunk
    x = rng.uniform(-1, 1, (S, L)).astype(np.float32)
    w = oracle.window(wt, N)
    plan = dsp.Plan(N, H, analysis_window=w, ola_window=w, boundary_mode=dsp.DROP)
    st = dsp.Stream(plan, S, max_chunk=chunk)
    assert st.ring_capacity() >= N + chunk
    got = run_ticks(cuda, st, x, chunk)
    F = oracle.frame_count(L, N, H, po.DROP)
    assert st.samples_ingested() == L and st.frames_processed() == F
    assert got.shape[1] == (F - 1) * H + N
    batch, _, _ = plan.roundtrip(dev(cuda, x))
    assert np.array_equal(bits(got), bits(batch.cpu().numpy()))          # same kernels' arithmetic, same order
    for s in (0, S - 1):
        ref = oracle.roundtrip(x[s], N, H, w, po.DROP)
        assert rel_l2(got[s], ref["out"]) < TOL


def test_stream_run_graph_replay_equals_ticks(cuda, dsp, oracle):
    """crlot_stream_run (tick loop in the library, CUDA-graph replay) == tick-by-tick, and the state
    carries across calls (run, then more ticks, then drain)."""
    rng = np.random.default_rng(5)
    N, H, S, chunk, T = 256, 64, 300, 64, 70
    w = oracle.window(po.HAMMING, N)
    x = rng.uniform(-1, 1, (S, T * chunk)).astype(np.float32)
    plan = dsp.Plan(N, H, analysis_window=w, ola_window=w, boundary_mode=dsp.DROP)
    a = dsp.Stream(plan, S, max_chunk=chunk)
    want = run_ticks(cuda, a, x, chunk)
    b = dsp.Stream(plan, S, max_chunk=chunk)
    xs = dev(cuda, x)
    o1 = b.run(xs[:, :50 * chunk], chunk).cpu().numpy().copy()           # 50 ticks: graph path (>= 16)
    o2 = b.run(xs[:, 50 * chunk:], chunk, n_ticks=5).cpu().numpy().copy()  # 5 ticks: direct launches
    rest = [b.tick(xs[:, t * chunk:(t + 1) * chunk]).cpu().numpy().copy() for t in range(55, T)]
    got = np.concatenate([o1, o2] + rest + [b.drain().cpu().numpy()], axis=1)
    assert np.array_equal(bits(got), bits(want))
    # reset brings the state back to the beginning
    b.reset()
    assert b.samples_ingested() == 0 and b.frames_processed() == 0
    again = run_ticks(cuda, b, x, chunk)
    assert np.array_equal(bits(again), bits(want))


def test_stream_warmup_ticks_and_spectra_tap(cuda, dsp, oracle):
    """No output until the first frame is complete (DROP); the spectra of the last tick are exposed."""
    N, H, S, chunk = 256, 64, 4, 64
    w = oracle.window(po.HAMMING, N)
    x = np.random.default_rng(9).uniform(-1, 1, (S, 6 * chunk)).astype(np.float32)
    plan = dsp.Plan(N, H, analysis_window=w, ola_window=w, boundary_mode=dsp.DROP)
    st = dsp.Stream(plan, S, max_chunk=chunk)
    xs = dev(cuda, x)
    for t in range(3):
        assert st.frames_ready(chunk) == 0
        assert st.tick(xs[:, t * chunk:(t + 1) * chunk]).shape[1] == 0
    assert st.frames_ready(chunk) == 1
    o = st.tick(xs[:, 3 * chunk:4 * chunk])
    assert o.shape[1] == H
    ptr, nf = st.last_spectra()
    assert nf == 1 and ptr
    # read the tap back through the C-ABI buffer calls (borrowed pointer -> download)
    import ctypes as C
    cuda.cuda.synchronize()
    L = dsp.lib()
    nbytes = S * (N // 2 + 1) * 8
    h, host = C.c_void_p(), np.empty(S * (N // 2 + 1), np.complex64)
    assert L.crlot_buf_wrap(C.c_void_p(ptr), C.c_size_t(nbytes), C.c_int(0), C.byref(h)) == 0
    assert L.crlot_buf_download(h, C.c_size_t(0), host.ctypes.data_as(C.c_void_p), C.c_size_t(nbytes), None) == 0
    assert L.crlot_cuda_stream_sync(C.c_int(0), None) == 0
    assert L.crlot_buf_destroy(h) == 0
    tap = host.reshape(S, N // 2 + 1)
    ref = oracle.rfft_many((x[:, :N] * w), N)
    assert rel_l2(tap, ref) < TOL
    S1 = plan.stft(dev(cuda, x[:, :N].copy())).cpu().numpy()[:, 0]
    assert np.array_equal(bits(tap), bits(S1))                           # identical to kernel 1's arithmetic


def test_stream_argument_errors(cuda, dsp, oracle):
    w = oracle.window(po.HANN, 256)
    zp = dsp.Plan(256, 64, analysis_window=w, ola_window=w, boundary_mode=dsp.ZERO_PAD)
    with pytest.raises(dsp.CrlotUnsupported):
        dsp.Stream(zp, 4, 64)                       # chunked ZERO_PAD is not a sane drive (framer.cc:98-103)
    odd = dsp.Plan(1000, 250, boundary_mode=dsp.DROP)          # hop % 4 != 0: the ring kernels move 16-byte groups
    with pytest.raises(dsp.CrlotUnsupported):
        dsp.Stream(odd, 4, 250)
    plan = dsp.Plan(256, 64, analysis_window=w, ola_window=w, boundary_mode=dsp.DROP)
    with pytest.raises(dsp.CrlotInvalidArgument):
        dsp.Stream(plan, 0, 64)
    st = dsp.Stream(plan, 2, 64)
    with pytest.raises(dsp.CrlotInvalidArgument):
        st.tick(cuda.zeros((2, 128), dtype=cuda.float32, device="cuda"))     # chunk_len > max_chunk
    assert st.drain().shape[1] == 0                  # nothing pushed yet: produce() has nothing to give


def test_stream_run_called_repeatedly_repoints_its_graphs(cuda, dsp, oracle):
    """crlot_stream_run keeps its instantiated graphs on the state and re-points them per call (different input /
    output pointers and offsets each time): the concatenation of several runs equals one run, and the call does not
    block the host stream (results are read only after an explicit synchronize)."""
    rng = np.random.default_rng(11)
    N, H, S, chunk, T = 256, 64, 64, 64, 192
    w = oracle.window(po.HAMMING, N)
    x = rng.uniform(-1, 1, (S, T * chunk)).astype(np.float32)
    plan = dsp.Plan(N, H, analysis_window=w, ola_window=w, boundary_mode=dsp.DROP)
    a = dsp.Stream(plan, S, max_chunk=chunk)
    xs = dev(cuda, x)
    want = a.run(xs, chunk).cpu().numpy().copy()
    b = dsp.Stream(plan, S, max_chunk=chunk)
    parts = []
    for i in range(3):                                           # three calls of 64 ticks, three different buffers
        seg = xs[:, i * 64 * chunk:(i + 1) * 64 * chunk].clone()
        parts.append(b.run(seg, chunk))
    cuda.cuda.synchronize()
    got = np.concatenate([p.cpu().numpy() for p in parts], axis=1)
    assert np.array_equal(bits(got), bits(want))


@pytest.mark.parametrize("N,H,chunk", [(256, 64, 64), (1024, 256, 256), (256, 64, 160), (4096, 1024, 1024)])
def test_stream_host_ticks_zero_copy_staged_and_pipelined(cuda, dsp, oracle, N, H, chunk):
    """crlot_stream_tick_host on page-locked buffers (zero-copy: the tick kernel reads / writes host memory itself),
    on pageable buffers (staged copies), and crlot_stream_submit_host / _wait_host with several ticks in flight:
    all bit-identical to device-resident ticks."""
    rng = np.random.default_rng(N + chunk)
    S, T = 9, 30
    w = oracle.window(po.HANN, N)
    x = rng.uniform(-1, 1, (S, T * chunk)).astype(np.float32)
    plan = dsp.Plan(N, H, analysis_window=w, ola_window=w, boundary_mode=dsp.DROP)
    ref = dsp.Stream(plan, S, max_chunk=chunk)
    want = run_ticks(cuda, ref, x, chunk)[:, :-(N - H)]                  # without the final drain
    max_out = (chunk // H + 1) * H
    # (a) pinned, synchronous
    a = dsp.Stream(plan, S, max_chunk=chunk)
    hin = cuda.from_numpy(x).pin_memory()
    hout = cuda.empty((S, max_out), dtype=cuda.float32).pin_memory()
    hin_np, hout_np = hin.numpy(), hout.numpy()
    outs = []
    for t in range(T):
        n = a.tick_host(hin_np[:, t * chunk:(t + 1) * chunk], hout_np)
        outs.append(hout_np[:, :n].copy())
    assert np.array_equal(bits(np.concatenate(outs, axis=1)), bits(want))
    # (b) pageable, synchronous (staged through device buffers)
    b = dsp.Stream(plan, S, max_chunk=chunk)
    pout = np.empty((S, max_out), np.float32)
    outs = []
    for t in range(T):
        n = b.tick_host(np.ascontiguousarray(x[:, t * chunk:(t + 1) * chunk]), pout)
        outs.append(pout[:, :n].copy())
    assert np.array_equal(bits(np.concatenate(outs, axis=1)), bits(want))
    # (c) pinned, three ticks in flight
    c = dsp.Stream(plan, S, max_chunk=chunk)
    depth = 3
    ring = [cuda.empty((S, max_out), dtype=cuda.float32).pin_memory() for _ in range(depth)]
    ring_np = [r.numpy() for r in ring]
    pend, outs = [], []
    for t in range(T):
        if len(pend) == depth:
            tk, n, slot = pend.pop(0)
            c.wait_host(tk)
            outs.append(ring_np[slot][:, :n].copy())
        n, tk = c.submit_host(hin_np[:, t * chunk:(t + 1) * chunk], ring_np[t % depth])
        pend.append((tk, n, t % depth))
    for tk, n, slot in pend:
        c.wait_host(tk)
        outs.append(ring_np[slot][:, :n].copy())
    assert np.array_equal(bits(np.concatenate(outs, axis=1)), bits(want))
