"""crlot-dsp_b200 -- B200 (sm_100a) implementation of crlot-dsp's STFT -> iSTFT hot path.

The product is the C-ABI shared library ``libcrlot_cuda.so`` (include/crlot_cuda.h) built from
csrc/ by build.py, plus the C++ facades in host/ that keep the reference's class API.  This
Python package is only the thin host-side binding used by tests/, bench.py and
__graft_entry__.py: ctypes over the C-ABI, torch for device memory / streams /
torch.distributed plumbing.  It never computes on the CPU: if the library is missing or no
CUDA device is present, compute calls raise.
"""
from .capi import (  # noqa: F401
    CrlotError, CrlotInvalidArgument, CrlotUnsupported, Plan, FftPlan, OlaAccumulator, Stream, lib, lib_path, frame_count, ola_ring_len, window_make,
    build_norm_linear, ola_envelope, synth_uniform, HANN, HAMMING, BLACKMAN, RECT, ZERO_PAD, DROP,
    bytes_per_sample, framequeue_count, framequeue_extract, PAD_CONSTANT, PAD_REFLECT, PAD_EDGE,
    deinterleave, interleave, axpy, normalize_and_clear, RingBuffer, Framer,
)
