"""ctypes doorway to the two CPU checkers.  TEST INFRASTRUCTURE ONLY.

* ``Oracle``  -> oracle/libcrlot_oracle.so, our plain-C restatement (oracle/crlot_oracle.c);
  rebuilt from source with gcc when missing (works on the GPU box too).
* ``Ref``     -> oracle/_ref/libcrlot_ref.so (or the ``_strict`` variant), the UNMODIFIED
  reference compiled from /root/reference by oracle/Makefile.  Prebuilt in the build
  container; on the GPU box only the prebuilt file exists (``Ref.available()``).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module.  Both classes expose the same method names so a test can run one body
against either library.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_f32p = np.ctypeslib.ndpointer(dtype=np.float32, flags="C_CONTIGUOUS")
_vp = C.c_void_p

HANN, HAMMING, BLACKMAN, RECT = 0, 1, 2, 3
NORM_NONE, NORM_SUM_TO_ONE, NORM_L2, NORM_OLA_UNITY, NORM_OLA_SUM_WSQ = 0, 1, 2, 3, 4
ZERO_PAD, DROP = 0, 1
PAD_CONSTANT, PAD_REFLECT, PAD_EDGE = 0, 1, 2


def build_oracle(force: bool = False) -> str:
    so = os.path.join(_HERE, "libcrlot_oracle.so")
    src = os.path.join(_HERE, "crlot_oracle.c")
    if force or not os.path.exists(so) or os.path.getmtime(so) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "oracle"], stdout=subprocess.DEVNULL)
    return so


def build_ref() -> str | None:
    """Build oracle/_ref from /root/reference when that tree is present; else keep prebuilt."""
    so = os.path.join(_HERE, "_ref", "libcrlot_ref.so")
    if os.path.isdir("/root/reference"):
        subprocess.check_call(["make", "-C", _HERE, "ref"], stdout=subprocess.DEVNULL)
    return so if os.path.exists(so) else None


def _ptr(a):
    return None if a is None else a.ctypes.data_as(_vp)


class _Common:
    """Shared call surface; ``self._p`` is the symbol prefix ('orc_' or 'ref_')."""

    def _fn(self, name):
        return getattr(self.lib, self._p + name)

    # -- window ------------------------------------------------------------------------
    def window(self, wtype, n, periodic=False, norm=NORM_NONE):
        out = np.empty(n, np.float32)
        f = self._fn("window")
        f.restype = C.c_int
        rc = f(C.c_int(wtype), C.c_long(n), C.c_int(int(periodic)), C.c_int(norm), _ptr(out))
        if rc != 0:
            raise ValueError("window: bad arguments")
        return out

    # -- framer ------------------------------------------------------------------------
    def framer(self, x, frame, hop, channels=1, mode=ZERO_PAD):
        x = np.ascontiguousarray(x, np.float32)
        frames_in = x.size // channels
        f = self._fn("framer")
        f.restype = C.c_long
        n = f(_ptr(x), C.c_long(frames_in), C.c_long(frame), C.c_long(hop), C.c_long(channels),
              C.c_int(mode), None, C.c_long(0))
        out = np.empty((n, frame * channels), np.float32)
        if n:
            f(_ptr(x), C.c_long(frames_in), C.c_long(frame), C.c_long(hop), C.c_long(channels),
              C.c_int(mode), _ptr(out), C.c_long(n))
        return out

    def framequeue(self, x, frame, hop, center=True, pad_mode=0):
        """FrameQueue(...).getAllFrames() -> [F][frame] (pad_mode: 0 CONSTANT, 1 REFLECT, 2 EDGE)."""
        x = np.ascontiguousarray(x, np.float32)
        f = self._fn("framequeue")
        f.restype = C.c_long
        args = [_ptr(x) if x.size else None, C.c_long(x.size), C.c_long(frame), C.c_long(hop), C.c_int(int(center)), C.c_int(pad_mode)]
        n = f(*args, None, C.c_long(0))
        if n < 0:
            raise ValueError("framequeue: bad arguments")
        out = np.empty((n, frame), np.float32)
        if n:
            f(*args, _ptr(out), C.c_long(n))
        return out

    # -- fft ---------------------------------------------------------------------------
    def rfft(self, x, nfft, batch=1, stride_in=1, stride_out=1, out_len=None):
        x = np.ascontiguousarray(x, np.float32)
        nb = nfft // 2 + 1
        if out_len is None:
            out_len = batch * stride_out * nb
        out = np.zeros(out_len * 2, np.float32)
        f = self._fn("rfft")
        f.restype = C.c_int
        rc = f(C.c_int(nfft), C.c_int(batch), C.c_int(stride_in), C.c_int(stride_out), _ptr(x), _ptr(out))
        if rc != 0:
            raise RuntimeError("rfft plan rejected")
        return out.view(np.complex64)

    def irfft(self, X, nfft, batch=1, stride_in=1, stride_out=1, out_len=None):
        X = np.ascontiguousarray(X, np.complex64)
        if out_len is None:
            out_len = batch * stride_out * nfft
        out = np.zeros(out_len, np.float32)
        f = self._fn("irfft")
        f.restype = C.c_int
        rc = f(C.c_int(nfft), C.c_int(batch), C.c_int(stride_in), C.c_int(stride_out),
               _ptr(X.view(np.float32)), _ptr(out))
        if rc != 0:
            raise RuntimeError("irfft plan rejected")
        return out

    def cfft(self, x, nfft, inverse=False, batch=1, stride_in=1, stride_out=1, out_len=None):
        """Complex-domain plan: forward_complex / inverse_complex over complex64 arrays."""
        x = np.ascontiguousarray(x, np.complex64)
        n_out = batch * stride_out * nfft if out_len is None else out_len
        out = np.zeros(n_out, np.complex64)
        f = self._fn("cfft")
        f.restype = C.c_int
        rc = f(C.c_int(nfft), C.c_int(int(inverse)), C.c_int(batch), C.c_int(stride_in), C.c_int(stride_out),
               _ptr(x.view(np.float32)), _ptr(out.view(np.float32)))
        if rc != 0:
            raise RuntimeError("complex plan rejected")
        return out

    def rfft_many(self, frames, nfft):
        frames = np.ascontiguousarray(frames, np.float32).reshape(-1, nfft)
        out = np.empty((frames.shape[0], nfft // 2 + 1), np.complex64)
        f = self._fn("rfft_many")
        f.restype = C.c_int
        assert f(C.c_int(nfft), C.c_long(frames.shape[0]), _ptr(frames), _ptr(out.view(np.float32))) == 0
        return out

    def irfft_many(self, spec, nfft):
        spec = np.ascontiguousarray(spec, np.complex64).reshape(-1, nfft // 2 + 1)
        out = np.empty((spec.shape[0], nfft), np.float32)
        f = self._fn("irfft_many")
        f.restype = C.c_int
        assert f(C.c_int(nfft), C.c_long(spec.shape[0]), _ptr(spec.view(np.float32)), _ptr(out)) == 0
        return out

    # -- norm / kernels ----------------------------------------------------------------
    def build_norm_linear(self, window, ring_len, frame, hop):
        window = np.ascontiguousarray(window, np.float32)
        norm = np.empty(ring_len, np.float32)
        f = self._fn("build_norm_linear")
        f.restype = None
        f(_ptr(norm), _ptr(window), C.c_long(ring_len), C.c_long(frame), C.c_long(hop))
        return norm

    def axpy(self, dst, src, g):
        dst = np.array(dst, np.float32)
        f = self._fn("axpy")
        f.restype = None
        f(_ptr(dst), _ptr(np.ascontiguousarray(src, np.float32)), C.c_float(g), C.c_long(dst.size))
        return dst

    def axpy_windowed(self, dst, src, win, g):
        dst = np.array(dst, np.float32)
        f = self._fn("axpy_windowed")
        f.restype = None
        f(_ptr(dst), _ptr(np.ascontiguousarray(src, np.float32)),
          _ptr(np.ascontiguousarray(win, np.float32)), C.c_float(g), C.c_long(dst.size))
        return dst

    def normalize_and_clear(self, acc, norm, eps):
        acc = np.array(acc, np.float32)
        out = np.empty_like(acc)
        f = self._fn("normalize_and_clear")
        f.restype = None
        f(_ptr(out), _ptr(acc), _ptr(np.ascontiguousarray(norm, np.float32)), C.c_float(eps), C.c_long(acc.size))
        return out, acc

    # -- OLA streaming drive -----------------------------------------------------------
    def ola_stream(self, frames, frame, hop, channels, window, set_window=True, pass_window=False,
                   apply_window_inside=True, gain=1.0, eps=1e-8, shadow_ring=False):
        """frames: [F][channels][frame] -> (out [channels][(F-1)*hop+frame], n_written, peak)."""
        frames = np.ascontiguousarray(frames, np.float32).reshape(-1, channels, frame)
        F = frames.shape[0]
        out = np.zeros((channels, (F - 1) * hop + frame), np.float32)
        peak = C.c_float(0)
        window = None if window is None else np.ascontiguousarray(window, np.float32)
        f = self._fn("ola_stream")
        f.restype = C.c_long
        args = [_ptr(frames), C.c_long(F), C.c_long(frame), C.c_long(hop), C.c_long(channels),
                _ptr(window), C.c_int(int(set_window)), C.c_int(int(pass_window)),
                C.c_int(int(apply_window_inside))]
        if self._p == "ref_":
            args.append(C.c_int(int(shadow_ring)))
        args += [C.c_float(gain), C.c_float(eps), _ptr(out), C.byref(peak)]
        n = f(*args)
        return out, n, peak.value

    # -- whole path --------------------------------------------------------------------
    def roundtrip(self, x, nfft, hop, window, mode=DROP, gain=1.0, eps=1e-8, taps=True):
        """-> dict(n_frames, spectra [F][nb] c64, iframes [F][N], out [n_out])."""
        x = np.ascontiguousarray(x, np.float32)
        window = np.ascontiguousarray(window, np.float32)
        nb = nfft // 2 + 1
        F = (x.size - nfft) // hop + 1 if mode == DROP and x.size >= nfft else (
            0 if mode == DROP else (x.size + hop - 1) // hop)
        spectra = np.zeros((F, nb), np.complex64) if taps else None
        iframes = np.zeros((F, nfft), np.float32) if taps else None
        out = np.zeros(max((F - 1) * hop + nfft, 1), np.float32)
        n_out = C.c_long(0)
        f = self._fn("roundtrip")
        f.restype = C.c_long
        k = f(_ptr(x), C.c_long(x.size), C.c_int(nfft), C.c_long(hop), _ptr(window), C.c_int(mode),
              C.c_float(gain), C.c_float(eps), None if spectra is None else _ptr(spectra.view(np.float32)),
              _ptr(iframes), _ptr(out), C.byref(n_out))
        assert k == F, (k, F)
        return dict(n_frames=F, spectra=spectra, iframes=iframes, out=out[: n_out.value])


class Oracle(_Common):
    _p = "orc_"

    def __init__(self):
        self.lib = C.CDLL(build_oracle())

    def frame_count(self, length, frame, hop, mode):
        f = self.lib.orc_frame_count
        f.restype = C.c_long
        return f(C.c_long(length), C.c_long(frame), C.c_long(hop), C.c_int(mode))

    def ring_len(self, frame, hop):
        f = self.lib.orc_ring_len
        f.restype = C.c_long
        return f(C.c_long(frame), C.c_long(hop))

    def ola_envelope(self, window, frame, hop, apply_window_inside=True, eps=1e-8):
        ring = self.ring_len(frame, hop)
        norm = np.empty(ring, np.float32)
        window = None if window is None else np.ascontiguousarray(window, np.float32)
        self.lib.orc_ola_envelope.restype = None
        self.lib.orc_ola_envelope(_ptr(norm), _ptr(window), C.c_long(ring), C.c_long(frame), C.c_long(hop),
                                  C.c_int(int(apply_window_inside)), C.c_float(eps))
        return norm

    def ola_gather(self, frames, frame, hop, window, norm, gain=1.0, eps=1e-8, first_frame=0,
                   t0=0, n_out=None):
        frames = np.ascontiguousarray(frames, np.float32).reshape(-1, frame)
        F = frames.shape[0]
        if n_out is None:
            n_out = (F - 1) * hop + frame
        out = np.empty(n_out, np.float32)
        norm = np.ascontiguousarray(norm, np.float32)
        window = None if window is None else np.ascontiguousarray(window, np.float32)
        self.lib.orc_ola_gather.restype = None
        self.lib.orc_ola_gather(_ptr(frames), C.c_long(first_frame), C.c_long(F), C.c_long(first_frame + F),
                                C.c_long(frame), C.c_long(hop), _ptr(window), C.c_float(gain), _ptr(norm),
                                C.c_long(norm.size), C.c_float(eps), C.c_long(t0), C.c_long(n_out), _ptr(out))
        return out

    def synth_uniform(self, seed, channel, i0, n):
        out = np.empty(n, np.float32)
        self.lib.orc_synth_uniform.restype = None
        self.lib.orc_synth_uniform(C.c_ulonglong(seed), C.c_ulonglong(channel), C.c_ulonglong(i0),
                                   C.c_long(n), _ptr(out))
        return out

    def slice_roundtrip(self, seed, channel, t_from, n, nfft, hop, window, ring_len, mode=DROP):
        """Slice oracle: the round trip of the synthetic signal (seed, channel) restricted to
        absolute output samples [a, a+n), a = the first multiple of ring_len at or after t_from, plus nfft.
        A fresh accumulator started at a sample offset that is a multiple of ring_len (hence of hop) sees
        the same envelope phase and, after its first nfft samples, the same frames in the same order as the
        full run -- so the values returned are the full run's.  Returns (a, expected[n])."""
        base = -(-t_from // ring_len) * ring_len
        x = self.synth_uniform(seed, channel, base, n + 3 * nfft)
        out = self.roundtrip(x, nfft, hop, window, mode, taps=False)["out"]
        return base + nfft, out[nfft:nfft + n].copy()

    def synth_tones(self, i0, n):
        out = np.empty(n, np.float32)
        self.lib.orc_synth_tones.restype = None
        self.lib.orc_synth_tones(C.c_ulonglong(i0), C.c_long(n), _ptr(out))
        return out

    def bench_roundtrip_mt(self, threads, length, nfft, hop, wtype=HANN, seed=0xC0FFEE):
        f = self.lib.orc_bench_roundtrip_mt
        f.restype = C.c_double
        chk = C.c_double(0)
        t = f(C.c_int(threads), C.c_long(length), C.c_int(nfft), C.c_long(hop), C.c_int(wtype),
              C.c_ulonglong(seed), C.byref(chk))
        return t, chk.value


class Ref(_Common):
    _p = "ref_"

    @staticmethod
    def path(strict=False, isa="v3"):
        name = "libcrlot_ref_strict.so" if strict else ("libcrlot_ref_v4.so" if isa == "v4" else "libcrlot_ref.so")
        return os.path.join(_HERE, "_ref", name)

    @classmethod
    def available(cls, strict=False, isa="v3"):
        return os.path.exists(cls.path(strict, isa))

    @classmethod
    def best_isa(cls):
        """'v4' (AVX-512 F/BW/CD/DQ/VL build) when this host's cpuid has those features and the build exists,
        else 'v3' (AVX2+FMA): what -march=native would pick between the two shipped builds."""
        try:
            flags = set()
            for ln in open("/proc/cpuinfo"):
                if ln.startswith("flags"):
                    flags = set(ln.split(":", 1)[1].split())
                    break
            v4 = {"avx512f", "avx512bw", "avx512cd", "avx512dq", "avx512vl"} <= flags
        except OSError:
            v4 = False
        return "v4" if v4 and cls.available(isa="v4") else "v3"

    def __init__(self, strict=False, isa="v3"):
        self.isa = "v3-strict" if strict else isa
        self.lib = C.CDLL(self.path(strict, isa))

    def ring_len(self, frame, hop):
        f = self.lib.ref_ola_ring_size
        f.restype = C.c_long
        return f(C.c_long(frame), C.c_long(hop))

    def ola_stream_aos(self, frames, frame, hop, channels, window, apply_window_inside=True,
                       gain=1.0, eps=1e-8):
        """frames: [F][frame][channels] interleaved."""
        frames = np.ascontiguousarray(frames, np.float32).reshape(-1, frame, channels)
        F = frames.shape[0]
        out = np.zeros((channels, (F - 1) * hop + frame), np.float32)
        f = self.lib.ref_ola_stream_aos
        f.restype = C.c_long
        n = f(_ptr(frames), C.c_long(F), C.c_long(frame), C.c_long(hop), C.c_long(channels),
              _ptr(np.ascontiguousarray(window, np.float32)), C.c_int(int(apply_window_inside)),
              C.c_float(gain), C.c_float(eps), _ptr(out))
        return out, n
