"""CPU-side checks of the product: the C-ABI library loads and exports every symbol that
include/crlot_cuda.h declares, its host-side table builders agree bit-for-bit with the oracle
and the golden vectors, argument validation mirrors the reference, and the per-lane FFT
algebra (csrc/warpfft.cuh) is exercised by the CPU lane emulator.  No compute calls here."""
import os
import re
import subprocess

import numpy as np
import pyoracle as po
import pytest
from conftest import ROOT, bits


def test_abi_exports_every_declared_symbol(dsp):
    hdr = open(os.path.join(ROOT, "include", "crlot_cuda.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    names = sorted(set(re.findall(r"\b(crlot_[a-z0-9_]+)\s*\(", hdr)))
    assert len(names) >= 35
    L = dsp.lib()
    for n in names:
        assert hasattr(L, n), f"{n} declared in include/crlot_cuda.h but not exported"
    assert L.crlot_cuda_abi_version() == 1


def test_no_cpu_fallback_without_device(dsp):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a device is present")
    with pytest.raises(dsp.CrlotError):
        dsp.Plan(1024, 256)


def test_window_tables_match_oracle_and_golden(dsp, oracle, gold):
    g = gold("windows")
    names = ["hann", "hamming", "blackman", "rect"]
    for key in g.files:
        tn, n, per, nm = key.split("_")
        w = dsp.window_make(names.index(tn), int(n), bool(int(per)), int(nm))
        assert np.array_equal(bits(w), bits(g[key])), key
    with pytest.raises(ValueError):
        dsp.window_make(dsp.HANN, 0)                       # WindowLUT.cc:53
    with pytest.raises(ValueError):
        dsp.window_make(4, 16)                             # Blackman-Harris: WindowLUT.cc:242


def test_envelope_and_counts_match_oracle(dsp, oracle, gold):
    g = gold("norm")
    for key in g.files:
        _, N, H, t = key.split("_")
        N, H, t = int(N), int(H), int(t)
        w = oracle.window(t, N)
        ring = dsp.ola_ring_len(N, H)
        assert ring == oracle.ring_len(N, H)
        assert np.array_equal(bits(dsp.build_norm_linear(w, ring, N, H)), bits(g[key]))
    rng = np.random.default_rng(0)
    for N, H in [(1024, 256), (256, 256), (96, 40)]:
        w = rng.uniform(0, 1, N).astype(np.float32)
        for inside in (True, False):
            assert np.array_equal(bits(dsp.ola_envelope(w, N, H, inside)), bits(oracle.ola_envelope(w, N, H, inside)))
        assert np.array_equal(dsp.ola_envelope(None, N, H), np.ones(dsp.ola_ring_len(N, H), np.float32))
    for _ in range(100):
        L, N, H = int(rng.integers(0, 10000)), int(rng.integers(1, 5000)), int(rng.integers(1, 3000))
        for mode in (dsp.ZERO_PAD, dsp.DROP):
            assert dsp.frame_count(L, N, H, mode) == oracle.frame_count(L, N, H, mode)
    # config frame counts quoted in SURVEY.md 8a
    assert dsp.frame_count(285315, 1024, 256, dsp.DROP) == 1111
    assert dsp.frame_count(1 << 30, 2048, 512, dsp.DROP) == 2097149
    assert dsp.frame_count(1 << 24, 1024, 256, dsp.DROP) == 65533
    assert dsp.frame_count(1 << 33, 4096, 1024, dsp.DROP) == 8388605


def test_bytes_per_sample_formula(dsp):
    assert abs(dsp.bytes_per_sample(2048, 512) - 72.03125) < 1e-9
    assert abs(dsp.bytes_per_sample(1024, 256) - 72.0625) < 1e-9
    assert abs(dsp.bytes_per_sample(4096, 1024) - 72.015625) < 1e-9
    assert abs(dsp.bytes_per_sample(256, 64) - 72.25) < 1e-9


def test_lane_emulator(tmp_path):
    exe = str(tmp_path / "emu")
    src = os.path.join(ROOT, "tests", "cpu_emu", "emu_warpfft.cc")
    subprocess.check_call(["g++", "-std=c++17", "-O1", "-o", exe, src])
    out = subprocess.run([exe], capture_output=True, text=True)
    assert out.returncode == 0, out.stdout
    assert out.stdout.count("OK") >= 4
