"""Row a11 (dsp::ring::RingBuffer) and the kernel-dispatch seam on device pointers.

RingBuffer: the reference's split / write / shadow cases (ref: tests/ring_buffer_test.cc:25-234,
dsp/ring/ring_buffer.cc:44-205) against a pure-integer model of the reference arithmetic, over random
(capacity, start, len).  Seam: crlot_axpy / crlot_axpy_windowed / crlot_normalize_and_clear bit-exact against the
oracle's restatement of the reference's scalar kernels (ref: dsp/ola/kernels.cc:18-36), including the eps guard
and the closed forms of tests/kernels_test.cc:67-208."""
import numpy as np
import pytest
from conftest import bits

pytestmark = pytest.mark.gpu


def model_split(cap, start, length):
    """ring_buffer.cc:44-85, integers only"""
    if length == 0:
        return (0, 0), (0, 0)
    length = min(length, (1 << 64) - 1 - start, cap)
    start %= cap
    if start + length <= cap:
        return (start, length), (0, 0)
    return (start, cap - start), (0, length - (cap - start))


def test_ring_split_reference_cases_and_random(cuda, dsp):
    rb = dsp.RingBuffer(10)
    assert rb.capacity() == 10 and rb.write_pos() == 0 and rb.physical_capacity() == 10
    assert rb.split(2, 5) == ((2, 5), (0, 0))                    # SplitNoWrapAround
    assert rb.split(7, 5) == ((7, 3), (0, 2))                    # SplitWithWrapAround
    assert rb.split(5, 0) == ((0, 0), (0, 0))                    # SplitEmpty
    a, b = rb.split(0, (1 << 64) - 1)                            # OverflowProtection
    assert a[1] + b[1] == 10
    rb5 = dsp.RingBuffer(5)
    assert rb5.split(0, 5) == ((0, 5), (0, 0))                   # SplitFullBuffer
    assert rb5.split(3, 5) == ((3, 2), (0, 3))                   # SplitWrapAroundFullBuffer
    assert rb5.split(7, 3) == ((2, 3), (0, 0))                   # SplitStartBeyondCapacity
    assert rb5.split(0, 10) == ((0, 5), (0, 0))                  # LargeLenClamping
    with pytest.raises(ValueError, match="capacity must be > 0"):
        dsp.RingBuffer(0)                                        # ZeroCapacityThrows
    rng = np.random.default_rng(3)
    for _ in range(300):
        cap = int(rng.integers(1, 5000))
        r = dsp.RingBuffer(cap, shadow=bool(rng.integers(0, 2)))
        for _ in range(10):
            start, length = int(rng.integers(0, 4 * cap)), int(rng.integers(0, 3 * cap))
            assert r.split(start, length) == model_split(cap, start, length)


def test_ring_write_wrap_shadow_and_device_access(cuda, dsp):
    rb = dsp.RingBuffer(8, shadow=True)
    assert rb.physical_capacity() == 16
    assert rb.write(np.arange(1, 7, dtype=np.float32)) == 6 and rb.write_pos() == 6
    assert rb.write(np.arange(7, 12, dtype=np.float32)) == 5 and rb.write_pos() == 3     # 2 in the tail, 3 wrap
    v = rb.host_view()
    assert list(v[:8]) == [9, 10, 11, 4, 5, 6, 7, 8]
    assert list(v[8:11]) == [9, 10, 11]                          # the mirrored head (ring_buffer.cc:197-200)
    assert rb.contiguous_offset(6) == 6 and dsp.RingBuffer(8).contiguous_offset(13) == 5
    assert rb.write(np.zeros(0, np.float32)) == 0
    # exactly to the end: write_pos returns to 0 (ring_buffer.cc:177-179)
    r2 = dsp.RingBuffer(4)
    assert r2.write(np.ones(4, np.float32)) == 4 and r2.write_pos() == 0
    # the cells are device memory: a device-side write lands in the same cells the host reads
    r3 = dsp.RingBuffer(6)
    src = cuda.arange(10, 14, dtype=cuda.float32, device="cuda")
    assert r3.write_device(src) == 4
    assert r3.write_device(src) == 4 and r3.write_pos() == 2
    cuda.cuda.synchronize()
    assert list(r3.host_view()) == [12, 13, 12, 13, 10, 11]
    # shadow_sync(n) mirrors the head explicitly
    r4 = dsp.RingBuffer(4, shadow=True)
    r4.write(np.array([5, 6, 7], np.float32))
    r4.shadow_sync(2)
    assert list(r4.host_view()[4:6]) == [5, 6]


def test_seam_kernels_bit_exact_vs_oracle(cuda, dsp, oracle):
    rng = np.random.default_rng(8)
    for n in (1, 7, 256, 4097, 100003):
        dst = rng.standard_normal(n).astype(np.float32)
        src = rng.standard_normal(n).astype(np.float32)
        win = rng.uniform(0, 1, n).astype(np.float32)
        g = np.float32(0.37)
        d = cuda.from_numpy(dst.copy()).cuda()
        dsp.axpy(d, cuda.from_numpy(src).cuda(), float(g))
        assert np.array_equal(bits(d.cpu().numpy()), bits(oracle.axpy(dst, src, g)))
        d = cuda.from_numpy(dst.copy()).cuda()
        dsp.axpy(d, cuda.from_numpy(src).cuda(), float(g), window=cuda.from_numpy(win).cuda())
        assert np.array_equal(bits(d.cpu().numpy()), bits(oracle.axpy_windowed(dst, src, win, g)))
        norm = rng.uniform(-0.1, 2, n).astype(np.float32)
        norm[::5] = 0.0                                          # the eps guard (kernels_test.cc:431-458)
        acc = cuda.from_numpy(dst.copy()).cuda()
        out = dsp.normalize_and_clear(acc, cuda.from_numpy(norm).cuda(), 1e-3)
        want_out, want_acc = oracle.normalize_and_clear(dst, norm, np.float32(1e-3))
        assert np.array_equal(bits(out.cpu().numpy()), bits(want_out))
        assert not acc.any().item()
    # closed forms (kernels_test.cc:67-120): dst = 1 + 2*0.5, and windowed 1 + 2*0.25*0.5
    one, two = cuda.ones(64, device="cuda"), cuda.full((64,), 2.0, device="cuda")
    assert (dsp.axpy(one.clone(), two, 0.5) == 2.0).all().item()
    assert (dsp.axpy(one.clone(), two, 0.5, window=cuda.full((64,), 0.25, device="cuda")) == 1.25).all().item()
