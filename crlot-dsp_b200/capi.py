"""ctypes binding of include/crlot_cuda.h.  Host-side mirror of the reference's interface for
the hot path: same names / argument meaning / error behaviour (status codes come back as the
Python analogue of the reference's exception types)."""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
HANN, HAMMING, BLACKMAN, RECT, BLACKMAN_HARRIS = 0, 1, 2, 3, 4
NORM_NONE, NORM_SUM_TO_ONE, NORM_L2, NORM_OLA_UNITY_GAIN, NORM_OLA_SUM_WSQ = 0, 1, 2, 3, 4
ZERO_PAD, DROP = 0, 1


class CrlotError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"crlot_cuda status {code}: {msg}")
        self.code = code


class CrlotInvalidArgument(CrlotError, ValueError):     # std::invalid_argument
    pass


class CrlotUnsupported(CrlotError):
    pass


def lib_path() -> str:
    return os.path.join(HERE, "libcrlot_cuda.so")


_lib = None


class _Desc(C.Structure):
    _fields_ = [("nfft", C.c_int), ("hop", C.c_int), ("boundary_mode", C.c_int),
                ("analysis_window", C.c_void_p), ("ola_window", C.c_void_p), ("push_window", C.c_void_p),
                ("apply_window_inside", C.c_int), ("gain", C.c_float), ("eps", C.c_float), ("scrub", C.c_int),
                ("envelope", C.c_void_p), ("envelope_len", C.c_size_t), ("device", C.c_int)]


class _Info(C.Structure):
    _fields_ = [("nfft", C.c_int), ("hop", C.c_int), ("bins", C.c_int), ("ring_len", C.c_size_t),
                ("synthesis_window_active", C.c_int), ("device", C.c_int)]


def lib():
    """Load the CUDA C-ABI library.  There is no fallback: a missing library is an error."""
    global _lib
    if _lib is None:
        p = lib_path()
        if not os.path.exists(p):
            raise ImportError(f"{p} not built: run `python crlot-dsp_b200/build.py` (the product has no CPU path)")
        L = C.CDLL(p)
        L.crlot_cuda_last_error.restype = C.c_char_p
        L.crlot_frame_count.restype = C.c_longlong
        L.crlot_framequeue_count.restype = C.c_longlong
        L.crlot_ola_ring_len.restype = C.c_size_t
        L.crlot_buf_device_ptr.restype = C.c_void_p
        L.crlot_buf_bytes.restype = C.c_size_t
        L.crlot_ola_get.restype = C.c_size_t
        L.crlot_ola_meter_peak.restype = C.c_float
        L.crlot_stream_get.restype = C.c_longlong
        _lib = L
    return _lib


def _check(rc):
    if rc != 0:
        msg = lib().crlot_cuda_last_error().decode()
        if rc == 1:
            raise CrlotInvalidArgument(rc, msg)
        if rc == 4:
            raise CrlotUnsupported(rc, msg)
        if rc == 3:
            raise MemoryError(msg)
        raise CrlotError(rc, msg)


def _f32(a):
    return None if a is None else np.ascontiguousarray(a, np.float32)


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


# ---- host-side table builders -----------------------------------------------------------------
def window_make(wtype, n, periodic=False, norm=NORM_NONE):
    out = np.empty(max(int(n), 0), np.float32)
    _check(lib().crlot_window_make(C.c_int(wtype), C.c_size_t(n), C.c_int(int(periodic)), C.c_int(norm), _p(out)))
    return out


def ola_ring_len(frame, hop):
    return int(lib().crlot_ola_ring_len(C.c_size_t(frame), C.c_size_t(hop)))


def build_norm_linear(window, ring_len, frame, hop):
    w = _f32(window)
    norm = np.empty(ring_len, np.float32)
    _check(lib().crlot_build_norm_linear(_p(norm), _p(w), C.c_size_t(ring_len), C.c_size_t(frame), C.c_size_t(hop)))
    return norm


def ola_envelope(window, frame, hop, apply_window_inside=True, eps=1e-8):
    ring = ola_ring_len(frame, hop)
    norm = np.empty(ring, np.float32)
    w = _f32(window)
    _check(lib().crlot_ola_envelope(_p(norm), _p(w), C.c_size_t(ring), C.c_size_t(frame), C.c_size_t(hop),
                                    C.c_int(int(apply_window_inside)), C.c_float(eps)))
    return norm


def frame_count(length, frame, hop, mode):
    return int(lib().crlot_frame_count(C.c_longlong(length), C.c_int(frame), C.c_int(hop), C.c_int(mode)))


PAD_CONSTANT, PAD_REFLECT, PAD_EDGE = 0, 1, 2      # dsp::PadMode (ref: dsp/frame/FrameQueue.h:8-12)


def framequeue_count(length, frame, hop, center=True):
    """FrameQueue::getNumFrames() (ref: dsp/frame/FrameQueue.cc:100-115)."""
    n = int(lib().crlot_framequeue_count(C.c_longlong(length), C.c_int(frame), C.c_int(hop), C.c_int(int(center))))
    if n < 0:
        raise CrlotInvalidArgument(1, "Frame size and hop size must be greater than 0")
    return n


def framequeue_extract(signal, frame, hop, center=True, pad_mode=PAD_CONSTANT, device=0):
    """FrameQueue(...).getAllFrames() on the device: cuda float32 [L] -> [F, frame], bit-exact."""
    torch = _torch()
    L = signal.numel()
    F = framequeue_count(L, frame, hop, center)
    out = torch.empty((F, frame), dtype=torch.float32, device=signal.device)
    _check(lib().crlot_framequeue_extract(C.c_void_p(signal.data_ptr() if L else 0), C.c_longlong(L), C.c_int(frame), C.c_int(hop),
                                          C.c_int(int(center)), C.c_int(pad_mode), C.c_longlong(F), C.c_void_p(out.data_ptr() if F else 0),
                                          C.c_int(device), _stream_ptr(device)))
    return out


def bytes_per_sample(nfft, hop):
    """Algorithmic HBM bytes per input sample of the three-kernel contract."""
    return 8.0 + 16.0 * (nfft / 2 + 1) / hop + 8.0 * nfft / hop


def _torch():
    import torch
    return torch


def _stream_ptr(device):
    torch = _torch()
    return C.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def synth_uniform(channels, n, seed=0xC0FFEE, channel0=0, i0=0, device=0, out=None):
    torch = _torch()
    if out is None:
        out = torch.empty((channels, n), dtype=torch.float32, device=f"cuda:{device}")
    _check(lib().crlot_synth_uniform(C.c_void_p(out.data_ptr()), C.c_longlong(out.stride(0) if out.dim() == 2 else n),
                                     C.c_int(channels), C.c_longlong(n), C.c_ulonglong(seed), C.c_ulonglong(channel0),
                                     C.c_ulonglong(i0), C.c_int(device), _stream_ptr(device)))
    return out


# ---- STFT plan ----------------------------------------------------------------------------------
class Plan:
    """crlot_stft_plan_*: tables for one (nfft, hop, windows, OLA policy) on one device.

    analysis_window : what the caller multiplies frames by before forward()  (e2e_benchmark.cc:154-156)
    ola_window      : what OLAAccumulator::set_window() received
    push_window     : per-call window argument of add_frame_SoA (used when not apply_window_inside)
    """

    def __init__(self, nfft, hop, analysis_window=None, ola_window=None, push_window=None,
                 apply_window_inside=True, gain=1.0, eps=1e-8, boundary_mode=DROP, scrub=True,
                 envelope=None, device=0):
        self._h = C.c_void_p()
        self._keep = [_f32(analysis_window), _f32(ola_window), _f32(push_window), _f32(envelope)]
        for w in self._keep[:3]:
            if w is not None and w.size != nfft:
                raise CrlotInvalidArgument(1, "Window size must match frame size")
        d = _Desc(nfft, hop, boundary_mode, _p(self._keep[0]), _p(self._keep[1]), _p(self._keep[2]),
                  int(apply_window_inside), gain, eps, int(scrub), _p(self._keep[3]),
                  0 if envelope is None else self._keep[3].size, device)
        _check(lib().crlot_stft_plan_create(C.byref(d), C.byref(self._h)))
        info = _Info()
        _check(lib().crlot_stft_plan_info(self._h, C.byref(info)))
        self.nfft, self.hop, self.bins, self.ring_len = info.nfft, info.hop, info.bins, info.ring_len
        self.synthesis_window_active = bool(info.synthesis_window_active)
        self.device = device
        self.boundary_mode = boundary_mode

    def close(self):
        if getattr(self, "_h", None) is not None and self._h.value:
            lib().crlot_stft_plan_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def envelope(self):
        out = np.empty(self.ring_len, np.float32)
        _check(lib().crlot_stft_plan_envelope(self._h, _p(out)))
        return out

    # -- spectral-domain hook -------------------------------------------------
    def set_spectral_gain(self, gain):
        """Attach a per-bin real gain applied as the inverse kernels load each spectrum row.
        gain: cuda float32 [bins] (same gains for every frame) or [..., bins] with one row per spectrum
        row (time-frequency mask, laid out like the spectra); None clears the hook."""
        if gain is None:
            self._sgain = None
            _check(lib().crlot_stft_plan_set_spectral_gain(self._h, C.c_void_p(0), C.c_longlong(0)))
            return
        if gain.shape[-1] != self.bins or not gain.is_contiguous():
            raise CrlotInvalidArgument(1, "gain must be contiguous with bins as the last dimension")
        self._sgain = gain                      # keep the tensor alive while the plan points at it
        _check(lib().crlot_stft_plan_set_spectral_gain(self._h, C.c_void_p(gain.data_ptr()),
                                                       C.c_longlong(0 if gain.dim() == 1 else self.bins)))

    def apply_spectral_gain(self, spec, gain):
        """The hook as its own in-place pass over spec (complex64 [..., bins])."""
        rows = spec.numel() // self.bins
        _check(lib().crlot_spectral_gain_apply(self._h, C.c_void_p(spec.data_ptr()), C.c_longlong(rows),
                                               C.c_void_p(gain.data_ptr()), C.c_longlong(0 if gain.dim() == 1 else self.bins),
                                               _stream_ptr(self.device)))
        return spec

    def n_frames(self, length):
        return frame_count(length, self.nfft, self.hop, self.boundary_mode)

    def out_len(self, length):
        f = self.n_frames(length)
        return (f - 1) * self.hop + self.nfft if f > 0 else 0

    # -- kernels, device tensors -------------------------------------------------------------
    def stft(self, signal, length=None, n_frames=None, out=None):
        """signal: cuda float32 [channels, L] (row stride multiple of 4) -> complex64 [channels, F, bins]."""
        torch = _torch()
        if signal.dim() == 1:
            signal = signal.unsqueeze(0)
        ch = signal.shape[0]
        length = signal.shape[1] if length is None else length
        F = self.n_frames(length) if n_frames is None else n_frames
        if out is None:
            out = torch.empty((ch, F, self.bins), dtype=torch.complex64, device=signal.device)
        _check(lib().crlot_stft_execute(self._h, C.c_void_p(signal.data_ptr()), C.c_longlong(signal.stride(0) if ch > 1 else 0),
                                        C.c_longlong(length), C.c_int(ch), C.c_longlong(F), C.c_void_p(out.data_ptr()),
                                        _stream_ptr(self.device)))
        return out

    def stft_framequeue(self, signal, center=True, pad_mode=PAD_CONSTANT, length=None, out=None):
        """Kernel 1 over FrameQueue's frames (centre padding nfft/2, CONSTANT / REFLECT / EDGE) without
        materialising the padded signal.  signal: cuda float32 [channels, L] -> complex64 [channels, F, bins]."""
        torch = _torch()
        if signal.dim() == 1:
            signal = signal.unsqueeze(0)
        ch = signal.shape[0]
        length = signal.shape[1] if length is None else length
        F = framequeue_count(length, self.nfft, self.hop, center)
        if out is None:
            out = torch.empty((ch, F, self.bins), dtype=torch.complex64, device=signal.device)
        _check(lib().crlot_stft_execute_framequeue(self._h, C.c_void_p(signal.data_ptr()),
                                                   C.c_longlong(signal.stride(0) if ch > 1 else 0), C.c_longlong(length),
                                                   C.c_int(ch), C.c_longlong(F), C.c_int(int(center)), C.c_int(pad_mode),
                                                   C.c_void_p(out.data_ptr()), _stream_ptr(self.device)))
        return out

    def stft_host_framequeue(self, x, center=True, pad_mode=PAD_CONSTANT, out=None):
        """x: numpy float32 [channels, L] or [L] (ideally page-locked) -> numpy complex64 [channels, F, bins]
        (crlot_stft_host_framequeue: FrameQueue's frames, windowed and transformed, slab by slab)."""
        x = np.asarray(x)
        assert x.dtype == np.float32 and x.flags.c_contiguous
        x2 = x.reshape(1, -1) if x.ndim == 1 else x
        ch, L = x2.shape
        F = max(0, framequeue_count(L, self.nfft, self.hop, center))
        if out is None:
            out = np.empty((ch, F, self.bins), np.complex64)
        got = C.c_longlong(0)
        _check(lib().crlot_stft_host_framequeue(self._h, _p(x2), C.c_longlong(L), C.c_longlong(L), C.c_int(ch),
                                                C.c_int(int(center)), C.c_int(pad_mode), _p(out), C.byref(got)))
        assert got.value == F
        return out

    def istft_frames(self, spec, fuse_synthesis_window=False, out=None):
        """spec: complex64 [..., bins] -> float32 [..., nfft] frames (1/N scaled, cleaned up)."""
        torch = _torch()
        rows = spec.numel() // self.bins
        if out is None:
            out = torch.empty(spec.shape[:-1] + (self.nfft,), dtype=torch.float32, device=spec.device)
        _check(lib().crlot_istft_execute(self._h, C.c_void_p(spec.data_ptr()), C.c_longlong(rows), C.c_void_p(out.data_ptr()),
                                         C.c_int(int(fuse_synthesis_window)), _stream_ptr(self.device)))
        return out

    def ola(self, frames, prewindowed=False, t_offset=0, frame_t0=0, n_out=None, out=None):
        """frames: float32 [channels, F, nfft] -> float32 [channels, n_out]."""
        torch = _torch()
        if frames.dim() == 2:
            frames = frames.unsqueeze(0)
        ch, F = frames.shape[0], frames.shape[1]
        if n_out is None:
            n_out = (F - 1) * self.hop + self.nfft if F > 0 else 0
        if out is None:
            out = torch.empty((ch, n_out), dtype=torch.float32, device=frames.device)
        _check(lib().crlot_ola_accumulate(self._h, C.c_void_p(frames.data_ptr()), C.c_longlong(F), C.c_longlong(F), C.c_int(ch),
                                          C.c_int(int(prewindowed)), C.c_longlong(t_offset), C.c_longlong(frame_t0),
                                          C.c_void_p(out.data_ptr()), C.c_longlong(out.stride(0) if ch > 1 else 0),
                                          C.c_longlong(n_out), _stream_ptr(self.device)))
        return out

    def selftest_division(self, n=1 << 28, seed=1):
        """Mismatches between the fused kernel's table-driven division and __fdiv_rn over n operand pairs."""
        bad = C.c_ulonglong(0)
        _check(lib().crlot_selftest_division(self._h, C.c_ulonglong(n), C.c_ulonglong(seed), C.byref(bad)))
        return bad.value

    def istft_ola_supported(self, t_offset=0, frame_t0=0, out=None):
        ptr = C.c_void_p(out.data_ptr() if out is not None else 0)
        stride = out.stride(0) if out is not None and out.dim() == 2 and out.shape[0] > 1 else 0
        return bool(lib().crlot_istft_ola_supported(self._h, C.c_longlong(t_offset), C.c_longlong(frame_t0), ptr,
                                                    C.c_longlong(stride)))

    def istft_ola(self, spec, t_offset=0, frame_t0=0, n_out=None, out=None, frames_per_chan=None):
        """Kernels 2+3 fused.  spec: complex64 [channels, F, bins] -> float32 [channels, n_out];
        bit-identical to ola(istft_frames(spec, True), prewindowed=True, ...)."""
        torch = _torch()
        if spec.dim() == 2:
            spec = spec.unsqueeze(0)
        ch, rows = spec.shape[0], spec.shape[1]
        F = rows if frames_per_chan is None else frames_per_chan
        if n_out is None:
            n_out = (F - 1) * self.hop + self.nfft if F > 0 else 0
        if out is None:
            out = torch.empty((ch, n_out), dtype=torch.float32, device=spec.device)
        _check(lib().crlot_istft_ola_execute(self._h, C.c_void_p(spec.data_ptr()), C.c_longlong(rows), C.c_longlong(F),
                                             C.c_int(ch), C.c_longlong(t_offset), C.c_longlong(frame_t0),
                                             C.c_void_p(out.data_ptr()), C.c_longlong(out.stride(0) if ch > 1 else 0),
                                             C.c_longlong(n_out), _stream_ptr(self.device)))
        return out

    def roundtrip_fused(self, signal, spec=None, out=None, length=None):
        """K1 -> fused K2+K3 (iFFT frames never materialised); returns (out, spec)."""
        torch = _torch()
        if signal.dim() == 1:
            signal = signal.unsqueeze(0)
        ch = signal.shape[0]
        length = signal.shape[1] if length is None else length
        F = self.n_frames(length)
        if spec is None:
            spec = torch.empty((ch, F, self.bins), dtype=torch.complex64, device=signal.device)
        if out is None:
            out = torch.empty((ch, self.out_len(length)), dtype=torch.float32, device=signal.device)
        _check(lib().crlot_roundtrip_execute(self._h, C.c_void_p(signal.data_ptr()), C.c_longlong(signal.stride(0) if ch > 1 else 0),
                                             C.c_longlong(length), C.c_int(ch), C.c_void_p(spec.data_ptr()),
                                             C.c_void_p(0), C.c_void_p(out.data_ptr()),
                                             C.c_longlong(out.stride(0) if ch > 1 else 0), _stream_ptr(self.device)))
        return out, spec

    def roundtrip_fused_all_supported(self, t_offset=0, frame_t0=0):
        return bool(lib().crlot_roundtrip_fused_supported(self._h, C.c_void_p(0), C.c_longlong(0), C.c_longlong(t_offset),
                                                          C.c_longlong(frame_t0), C.c_void_p(0), C.c_longlong(0)))

    def roundtrip_fused_all(self, signal, length=None, n_frames=None, t_offset=0, frame_t0=0, n_out=None, out=None):
        """The whole round trip in one kernel (crlot_roundtrip_fused_execute): spectra never materialised."""
        torch = _torch()
        if signal.dim() == 1:
            signal = signal.unsqueeze(0)
        ch = signal.shape[0]
        length = signal.shape[1] if length is None else length
        F = self.n_frames(length) if n_frames is None else n_frames
        if n_out is None:
            n_out = (F - 1) * self.hop + self.nfft if F > 0 else 0
        if out is None:
            out = torch.empty((ch, n_out), dtype=torch.float32, device=signal.device)
        _check(lib().crlot_roundtrip_fused_execute(self._h, C.c_void_p(signal.data_ptr()), C.c_longlong(signal.stride(0) if ch > 1 else 0),
                                                   C.c_longlong(length), C.c_int(ch), C.c_longlong(F), C.c_longlong(t_offset),
                                                   C.c_longlong(frame_t0), C.c_void_p(out.data_ptr()),
                                                   C.c_longlong(out.stride(0) if ch > 1 else 0), C.c_longlong(n_out),
                                                   _stream_ptr(self.device)))
        return out

    def roundtrip(self, signal, spec=None, frames=None, out=None, length=None):
        """K1 -> K2 (synthesis window fused) -> K3 on device tensors; returns (out, spec, frames)."""
        torch = _torch()
        if signal.dim() == 1:
            signal = signal.unsqueeze(0)
        ch = signal.shape[0]
        length = signal.shape[1] if length is None else length
        F = self.n_frames(length)
        if spec is None:
            spec = torch.empty((ch, F, self.bins), dtype=torch.complex64, device=signal.device)
        if frames is None:
            frames = torch.empty((ch, F, self.nfft), dtype=torch.float32, device=signal.device)
        if out is None:
            out = torch.empty((ch, self.out_len(length)), dtype=torch.float32, device=signal.device)
        _check(lib().crlot_roundtrip_execute(self._h, C.c_void_p(signal.data_ptr()), C.c_longlong(signal.stride(0) if ch > 1 else 0),
                                             C.c_longlong(length), C.c_int(ch), C.c_void_p(spec.data_ptr()),
                                             C.c_void_p(frames.data_ptr()), C.c_void_p(out.data_ptr()),
                                             C.c_longlong(out.stride(0) if ch > 1 else 0), _stream_ptr(self.device)))
        return out, spec, frames

    def roundtrip_static(self, signal, out, length=None):
        """crlot_roundtrip_execute_static: the round trip with every intermediate in the library's static device
        scratch; nothing is allocated here or in the call -- the caller passes signal and output."""
        ch = 1 if signal.dim() == 1 else signal.shape[0]
        length = signal.shape[-1] if length is None else length
        _check(lib().crlot_roundtrip_execute_static(
            self._h, C.c_void_p(signal.data_ptr()), C.c_longlong(signal.stride(0) if ch > 1 else 0), C.c_longlong(length),
            C.c_int(ch), C.c_void_p(out.data_ptr()), C.c_longlong(out.stride(0) if ch > 1 else 0), _stream_ptr(self.device)))
        return out

    # -- host buffers (H2D / D2H inside the call) ----------------------------------------------
    def roundtrip_host(self, x, out=None):
        """x: numpy float32 [channels, L] or [L] (ideally page-locked) -> numpy [channels, n_out]."""
        x = np.asarray(x)
        assert x.dtype == np.float32 and x.flags.c_contiguous
        x2 = x.reshape(1, -1) if x.ndim == 1 else x
        ch, L = x2.shape
        n_out = self.out_len(L)
        if out is None:
            out = np.empty((ch, n_out), np.float32)
        got = C.c_longlong(0)
        _check(lib().crlot_roundtrip_host(self._h, _p(x2), C.c_longlong(L), C.c_longlong(L), C.c_int(ch), _p(out),
                                          C.c_longlong(out.shape[-1] if out.ndim == 2 else n_out), C.byref(got)))
        assert got.value == n_out
        return out


    def roundtrip_host_interleaved(self, x, out=None):
        """x: numpy float32 [L, channels] interleaved -> numpy [n_out, channels] (crlot_roundtrip_host_interleaved)."""
        x = np.asarray(x)
        assert x.dtype == np.float32 and x.flags.c_contiguous and x.ndim == 2
        L, ch = x.shape
        n_out = self.out_len(L)
        if out is None:
            out = np.empty((n_out, ch), np.float32)
        got = C.c_longlong(0)
        _check(lib().crlot_roundtrip_host_interleaved(self._h, _p(x), C.c_longlong(L), C.c_int(ch), _p(out), C.byref(got)))
        assert got.value == n_out
        return out


def deinterleave(aos, chan_stride=None, device=0):
    """cuda float32 [n, channels] -> [channels, chan_stride] planar (first n columns written)."""
    torch = _torch()
    n, ch = aos.shape
    cs = n if chan_stride is None else chan_stride
    out = torch.zeros((ch, cs), dtype=torch.float32, device=aos.device)
    _check(lib().crlot_deinterleave(C.c_void_p(aos.data_ptr()), C.c_longlong(n), C.c_int(ch), C.c_void_p(out.data_ptr()),
                                    C.c_longlong(cs), C.c_int(device), _stream_ptr(device)))
    return out


def interleave(planar, n=None, device=0):
    """cuda float32 [channels, stride] planar -> [n, channels] interleaved."""
    torch = _torch()
    ch = planar.shape[0]
    n = planar.shape[1] if n is None else n
    out = torch.empty((n, ch), dtype=torch.float32, device=planar.device)
    _check(lib().crlot_interleave(C.c_void_p(planar.data_ptr()), C.c_longlong(planar.stride(0)), C.c_longlong(n), C.c_int(ch),
                                  C.c_void_p(out.data_ptr()), C.c_int(device), _stream_ptr(device)))
    return out


# ---- drop-in FFT plan (dsp::fft::IFftPlan over host arrays) ---------------------------------------
class FftPlan:
    def __init__(self, nfft, batch=1, stride_in=1, stride_out=1, in_place=False, device=0, complex_domain=False):
        self._h = C.c_void_p()
        self.nfft, self.batch, self.si, self.so = nfft, batch, stride_in, stride_out
        self.complex_domain, self.device = complex_domain, device
        create = lib().crlot_fft_plan_create_complex if complex_domain else lib().crlot_fft_plan_create
        _check(create(C.c_int(nfft), C.c_int(int(in_place)), C.c_int(batch), C.c_int(stride_in),
                      C.c_int(stride_out), C.c_int(device), C.byref(self._h)))

    def _complex(self, fn, x, batch, out_len):
        x = np.ascontiguousarray(x, np.complex64)
        out = np.zeros(out_len if out_len is not None else batch * self.so * self.nfft, np.complex64)
        _check(fn(self._h, _p(x.view(np.float32)), _p(out.view(np.float32)), C.c_int(batch)))
        return out

    def forward_complex(self, x, batch=1, out_len=None):
        return self._complex(lib().crlot_fft_forward_complex, x, batch, out_len)

    def inverse_complex(self, x, batch=1, out_len=None):
        return self._complex(lib().crlot_fft_inverse_complex, x, batch, out_len)

    def complex_device(self, x, inverse=False):
        """cuda complex64 [rows, nfft] -> [rows, nfft] (crlot_fft_complex_device, any number of rows)."""
        torch = _torch()
        out = torch.empty_like(x)
        _check(lib().crlot_fft_complex_device(self._h, C.c_void_p(x.data_ptr()), C.c_void_p(out.data_ptr()),
                                              C.c_longlong(x.numel() // self.nfft), C.c_int(int(inverse)), _stream_ptr(self.device)))
        return out

    def __del__(self):
        try:
            if self._h.value:
                lib().crlot_fft_plan_destroy(self._h)
        except Exception:
            pass

    def size(self):
        return self.nfft

    def max_batch_size(self):
        return 16

    def forward(self, x, batch=1, out_len=None):
        x = _f32(x)
        nb = self.nfft // 2 + 1
        out = np.zeros(2 * (out_len if out_len is not None else batch * self.so * nb), np.float32)
        _check(lib().crlot_fft_forward(self._h, _p(x), _p(out), C.c_int(batch)))
        return out.view(np.complex64)

    def inverse(self, X, batch=1, out_len=None):
        X = np.ascontiguousarray(X, np.complex64)
        out = np.zeros(out_len if out_len is not None else batch * self.so * self.nfft, np.float32)
        _check(lib().crlot_fft_inverse(self._h, _p(X.view(np.float32)), _p(out), C.c_int(batch)))
        return out


# ---- drop-in OLAAccumulator -----------------------------------------------------------------------
class OlaAccumulator:
    """dsp::OLAAccumulator with device-resident rings (crlot_ola_*)."""

    def __init__(self, frame_size, hop_size, channels=1, sample_rate=48000, eps=1e-8, apply_window_inside=True,
                 shadow_ring=False, device=0):
        self._h = C.c_void_p()
        self.N, self.H, self.C = frame_size, hop_size, channels
        _check(lib().crlot_ola_create(C.c_int(sample_rate), C.c_size_t(frame_size), C.c_size_t(hop_size),
                                      C.c_size_t(channels), C.c_float(eps), C.c_int(int(apply_window_inside)),
                                      C.c_int(int(shadow_ring)), C.c_int(device), C.byref(self._h)))

    def __del__(self):
        try:
            if self._h.value:
                lib().crlot_ola_destroy(self._h)
        except Exception:
            pass

    def set_window(self, w, wlen=None):
        w = _f32(w)
        _check(lib().crlot_ola_set_window(self._h, _p(w), C.c_int(w.size if wlen is None else wlen)))

    def add_frame_SoA(self, ch_frames, window, start_sample, start_off, size, gain):
        arrs = [_f32(a) for a in ch_frames]
        ptrs = (C.c_void_p * len(arrs))(*[a.ctypes.data for a in arrs])
        w = _f32(window)
        _check(lib().crlot_ola_add_frame_soa(self._h, ptrs, _p(w), C.c_size_t(start_sample), C.c_size_t(start_off),
                                             C.c_size_t(size), C.c_float(gain)))

    def push_frame_AoS(self, interleaved, window, start_sample, start_off, size, gain):
        x = _f32(interleaved)
        w = _f32(window)
        _check(lib().crlot_ola_push_frame_aos(self._h, _p(x), _p(w), C.c_size_t(start_sample), C.c_size_t(start_off),
                                              C.c_size_t(size), C.c_float(gain)))

    def produce(self, n):
        outs = [np.zeros(n, np.float32) for _ in range(self.C)]
        ptrs = (C.c_void_p * self.C)(*[a.ctypes.data for a in outs])
        got = C.c_size_t(0)
        _check(lib().crlot_ola_produce(self._h, ptrs, C.c_size_t(n), C.byref(got)))
        return [a[: got.value] for a in outs], got.value

    def flush(self):
        _check(lib().crlot_ola_flush(self._h))

    def reset(self):
        _check(lib().crlot_ola_reset(self._h))

    def produced_samples(self):
        return int(lib().crlot_ola_get(self._h, 0))

    def read_pos(self):
        return int(lib().crlot_ola_get(self._h, 1))

    def ring_size(self):
        return int(lib().crlot_ola_get(self._h, 2))

    def has_window(self):
        return bool(lib().crlot_ola_get(self._h, 3))

    def meter_peak(self):
        return float(lib().crlot_ola_meter_peak(self._h))


# ---- streaming state (config 5) --------------------------------------------------------------------
class Stream:
    """crlot_stream_*: `n_streams` concurrent streams fed in chunks (device tensors in / out).

    Per stream this is the reference's chunk-fed DROP Framer -> window -> forward -> inverse ->
    OLAAccumulator.push_frame_AoS(k*hop) -> produce(hop) drive with device-resident rings."""

    def __init__(self, plan: Plan, n_streams: int, max_chunk: int):
        self._h = C.c_void_p()
        self.plan, self.n_streams, self.max_chunk = plan, n_streams, max_chunk
        _check(lib().crlot_stream_create(plan._h, C.c_int(n_streams), C.c_int(max_chunk), C.byref(self._h)))

    def close(self):
        if getattr(self, "_h", None) is not None and self._h.value:
            lib().crlot_stream_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def reset(self):
        _check(lib().crlot_stream_reset(self._h, _stream_ptr(self.plan.device)))

    def frames_ready(self, chunk_len):
        return int(lib().crlot_stream_frames_ready(self._h, C.c_int(chunk_len)))

    def tick(self, chunk, out=None):
        """chunk: cuda float32 [n_streams, chunk_len] (row stride free) -> view of out [n_streams, n_out]."""
        torch = _torch()
        chunk_len = chunk.shape[1]
        if out is None:
            out = torch.empty((self.n_streams, max(self.frames_ready(chunk_len), 1) * self.plan.hop),
                              dtype=torch.float32, device=chunk.device)
        n = C.c_int(0)
        _check(lib().crlot_stream_tick(self._h, C.c_void_p(chunk.data_ptr()), C.c_longlong(chunk.stride(0)),
                                       C.c_int(chunk_len), C.c_void_p(out.data_ptr()), C.c_longlong(out.stride(0)),
                                       C.byref(n), _stream_ptr(self.plan.device)))
        return out[:, :n.value]

    def run(self, signal, chunk_len, n_ticks=None, out=None):
        """Feed signal [n_streams, >= n_ticks*chunk_len] in n_ticks chunks; returns out[:, :n_out]."""
        torch = _torch()
        if n_ticks is None:
            n_ticks = signal.shape[1] // chunk_len
        if out is None:
            n_max = (n_ticks * chunk_len // self.plan.hop + 1) * self.plan.hop
            out = torch.empty((self.n_streams, n_max), dtype=torch.float32, device=signal.device)
        n = C.c_longlong(0)
        _check(lib().crlot_stream_run(self._h, C.c_void_p(signal.data_ptr()), C.c_longlong(signal.stride(0)),
                                      C.c_int(chunk_len), C.c_longlong(n_ticks), C.c_void_p(out.data_ptr()),
                                      C.c_longlong(out.stride(0)), C.byref(n), _stream_ptr(self.plan.device)))
        return out[:, :n.value]

    def tick_host(self, chunk, out):
        """chunk/out: numpy float32 [n_streams, chunk_len] / [n_streams, >= frames*hop] (ideally pinned)."""
        n = C.c_int(0)
        _check(lib().crlot_stream_tick_host(self._h, _p(chunk), C.c_longlong(chunk.strides[0] // 4), C.c_int(chunk.shape[1]),
                                            _p(out), C.c_longlong(out.strides[0] // 4), C.byref(n)))
        return n.value

    def submit_host(self, chunk, out):
        """tick_host without the wait: returns (n_out, ticket); wait_host(ticket) completes it."""
        n, t = C.c_int(0), C.c_ulonglong(0)
        _check(lib().crlot_stream_submit_host(self._h, _p(chunk), C.c_longlong(chunk.strides[0] // 4), C.c_int(chunk.shape[1]),
                                              _p(out), C.c_longlong(out.strides[0] // 4), C.byref(n), C.byref(t)))
        return n.value, t.value

    def wait_host(self, ticket):
        _check(lib().crlot_stream_wait_host(self._h, C.c_ulonglong(ticket)))

    def drain(self, out=None):
        torch = _torch()
        if out is None:
            out = torch.empty((self.n_streams, max(self.plan.nfft - self.plan.hop, 4)), dtype=torch.float32,
                              device=f"cuda:{self.plan.device}")
        n = C.c_int(0)
        _check(lib().crlot_stream_drain(self._h, C.c_void_p(out.data_ptr()), C.c_longlong(out.stride(0)), C.byref(n),
                                        _stream_ptr(self.plan.device)))
        return out[:, :n.value]

    def last_spectra(self):
        torch = _torch()
        ptr, nf = C.c_void_p(), C.c_int(0)
        _check(lib().crlot_stream_last_spectra(self._h, C.byref(ptr), C.byref(nf)))
        return ptr.value, nf.value

    def samples_ingested(self):
        return int(lib().crlot_stream_get(self._h, 0))

    def frames_processed(self):
        return int(lib().crlot_stream_get(self._h, 1))

    def ring_capacity(self):
        return int(lib().crlot_stream_get(self._h, 2))


# ---- kernel-dispatch seam on device pointers (dsp::ola::axpy / axpy_windowed / normalize_and_clear) -------
def axpy(dst, src, g, window=None, device=0):
    """dst[i] = fma(src[i] (* window[i], rounded), g, dst[i]) on cuda float32 tensors, in place."""
    n = dst.numel()
    if window is None:
        _check(lib().crlot_axpy(C.c_void_p(dst.data_ptr()), C.c_void_p(src.data_ptr()), C.c_float(g), C.c_size_t(n),
                                C.c_int(device), _stream_ptr(device)))
    else:
        _check(lib().crlot_axpy_windowed(C.c_void_p(dst.data_ptr()), C.c_void_p(src.data_ptr()), C.c_void_p(window.data_ptr()),
                                         C.c_float(g), C.c_size_t(n), C.c_int(device), _stream_ptr(device)))
    return dst


def normalize_and_clear(acc, norm, eps, device=0):
    """-> out = acc / max(norm, eps) (true division); acc is zeroed in place."""
    torch = _torch()
    out = torch.empty_like(acc)
    _check(lib().crlot_normalize_and_clear(C.c_void_p(out.data_ptr()), C.c_void_p(acc.data_ptr()), C.c_void_p(norm.data_ptr()),
                                           C.c_float(eps), C.c_size_t(acc.numel()), C.c_int(device), _stream_ptr(device)))
    return out


# ---- drop-in RingBuffer<float> ------------------------------------------------------------------------------
class _Span(C.Structure):
    _fields_ = [("off1", C.c_size_t), ("len1", C.c_size_t), ("off2", C.c_size_t), ("len2", C.c_size_t)]


class RingBuffer:
    """dsp::ring::RingBuffer<float> over crlot_ring_* (unified-memory cells; offsets instead of pointers)."""

    def __init__(self, capacity, shadow=False, device=0):
        self._h = C.c_void_p()
        L = lib()
        L.crlot_ring_data.restype = C.c_void_p
        for f in (L.crlot_ring_capacity, L.crlot_ring_physical_capacity, L.crlot_ring_write_pos, L.crlot_ring_contiguous_offset):
            f.restype = C.c_size_t
        _check(L.crlot_ring_create(C.c_size_t(capacity), C.c_size_t(4), C.c_int(int(shadow)), C.c_int(device), C.byref(self._h)))
        self.device = device

    def __del__(self):
        try:
            if self._h.value:
                lib().crlot_ring_destroy(self._h)
        except Exception:
            pass

    def capacity(self):
        return lib().crlot_ring_capacity(self._h)

    def physical_capacity(self):
        return lib().crlot_ring_physical_capacity(self._h)

    def write_pos(self):
        return lib().crlot_ring_write_pos(self._h)

    def data_ptr(self):
        return lib().crlot_ring_data(self._h)

    def host_view(self):
        """numpy view of all physical cells (unified memory: synchronise the device before touching it)."""
        n = self.physical_capacity()
        return np.ctypeslib.as_array((C.c_float * n).from_address(self.data_ptr()))

    def split(self, start, length):
        sp = _Span()
        _check(lib().crlot_ring_split(self._h, C.c_size_t(start), C.c_size_t(length), C.byref(sp)))
        return (sp.off1, sp.len1), (sp.off2, sp.len2)

    def shadow_sync(self, n):
        _check(lib().crlot_ring_shadow_sync(self._h, C.c_size_t(n)))

    def contiguous_offset(self, read_pos):
        return lib().crlot_ring_contiguous_offset(self._h, C.c_size_t(read_pos))

    def write(self, src):
        src = _f32(src)
        n = C.c_size_t(0)
        _check(lib().crlot_ring_write(self._h, _p(src), C.c_size_t(0 if src is None else src.size), C.byref(n)))
        return n.value

    def write_device(self, src):
        n = C.c_size_t(0)
        _check(lib().crlot_ring_write_device(self._h, C.c_void_p(src.data_ptr()), C.c_size_t(src.numel()), C.byref(n),
                                             _stream_ptr(self.device)))
        return n.value

    def prefetch(self, to_device=True):
        _check(lib().crlot_ring_prefetch(self._h, C.c_int(int(to_device)), _stream_ptr(self.device)))


# ---- drop-in Framer state (device-resident sample store) -------------------------------------------------------
class Framer:
    """dsp::Framer over crlot_framer_* (set_params at construction)."""

    def __init__(self, frame_size, hop_size, channels=1, boundary_mode=ZERO_PAD, device=0):
        self._h = C.c_void_p()
        L = lib()
        L.crlot_framer_available.restype = C.c_size_t
        L.crlot_framer_get.restype = C.c_size_t
        _check(L.crlot_framer_create(C.c_size_t(frame_size), C.c_size_t(hop_size), C.c_size_t(channels), C.c_int(boundary_mode),
                                     C.c_int(device), C.byref(self._h)))
        self.N, self.H, self.C, self.device = frame_size, hop_size, channels, device

    def __del__(self):
        try:
            if self._h.value:
                lib().crlot_framer_destroy(self._h)
        except Exception:
            pass

    def push(self, interleaved):
        x = _f32(interleaved)
        _check(lib().crlot_framer_push(self._h, _p(x), C.c_size_t(x.size // self.C)))

    def push_device(self, planar):
        """planar: cuda float32 [channels, n] (or [n] for one channel), on the current stream."""
        if planar.dim() == 1:
            planar = planar.unsqueeze(0)
        _check(lib().crlot_framer_push_device(self._h, C.c_void_p(planar.data_ptr()), C.c_longlong(planar.stride(0)),
                                              C.c_size_t(planar.shape[1]), _stream_ptr(self.device)))

    def available_frames(self):
        return lib().crlot_framer_available(self._h)

    def drainable_frames(self):
        return lib().crlot_framer_get(self._h, 8)

    def pop(self):
        out = np.empty(self.N * self.C, np.float32)
        ok = C.c_int(0)
        _check(lib().crlot_framer_pop(self._h, _p(out), C.byref(ok)))
        return out if ok.value else None

    def stft(self, plan, max_frames=-1, out=None):
        """every poppable frame through kernel 1 of `plan` on the current stream -> complex64 [channels, F, bins]"""
        torch = _torch()
        F = self.drainable_frames() if max_frames < 0 else min(self.drainable_frames(), max_frames)
        if out is None:
            out = torch.empty((self.C, F, plan.bins), dtype=torch.complex64, device=f"cuda:{self.device}")
        n = C.c_longlong(0)
        _check(lib().crlot_framer_stft(self._h, plan._h, C.c_longlong(max_frames), C.c_void_p(out.data_ptr()), C.byref(n),
                                       _stream_ptr(self.device)))
        return out[:, :n.value]

    def reset(self):
        _check(lib().crlot_framer_reset(self._h))
