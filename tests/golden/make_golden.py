"""Generate tests/golden/*.npz from the UNMODIFIED reference (oracle/_ref, built from
/root/reference by oracle/Makefile).  Run in the build container only:

    python tests/golden/make_golden.py

The vectors come from libcrlot_ref_strict.so (reference sources, -ffp-contract=off) so the
plain-C oracle can be pinned to them bit-for-bit anywhere; the default-flags build
(libcrlot_ref.so, FMA contraction allowed inside kissfft as in the reference's own release
build) differs from these by ~1e-7 relative L2, far inside the 1e-5 parity bar, and its
distance is recorded in each file as `fast_vs_strict_rel`.

c1 input: /root/reference/assets/oboe.wav (stereo s16, 44.1 kHz, 285 315 frames) mixed to
mono as (L+R)/2 (main/main.cc:156-160).  Stored losslessly as int32 (L+R); x = sum/65536.
"""
import os
import sys
import wave

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", "..", "oracle"))
import pyoracle as po  # noqa: E402

po.build_ref()
R = po.Ref(strict=True)
RF = po.Ref(strict=False)
rng = np.random.default_rng(20261104)


def rel(a, b):
    return float(np.linalg.norm((a - b).ravel()) / max(np.linalg.norm(b.ravel()), 1e-300))


def save(name, **kw):
    p = os.path.join(HERE, name)
    np.savez_compressed(p, **kw)
    print(f"{name}: {os.path.getsize(p)/1024:.1f} KiB")


# 1. windows ------------------------------------------------------------------------------
d = {}
for t, tn in enumerate(["hann", "hamming", "blackman", "rect"]):
    for n in (1, 2, 255, 256, 1024):
        for per in (0, 1):
            d[f"{tn}_{n}_{per}_0"] = R.window(t, n, per, 0)
    for nm in (1, 2, 3, 4):
        d[f"{tn}_256_0_{nm}"] = R.window(t, 256, 0, nm)
save("windows.npz", **d)

# 2. FFT known answers ----------------------------------------------------------------------
d = {}
for n in (256, 512, 1024, 2048, 4096):
    x = rng.uniform(-1, 1, n).astype(np.float32)
    X = R.rfft(x, n)
    d[f"x_{n}"] = x
    d[f"X_{n}"] = X
    d[f"y_{n}"] = R.irfft(X, n)
    d[f"fast_vs_strict_rel_{n}"] = rel(RF.rfft(x, n), X)
# scrub cases (kissfft_adapter.cc:104-108,159-163)
x = rng.uniform(-1, 1, 512).astype(np.float32)
x[3] = np.nan; x[77] = np.inf; x[200] = -np.inf; x[301] = 1e-35; x[302] = -1e-39
d["x_scrub"] = x
d["X_scrub"] = R.rfft(x, 512)
Xs = R.rfft(rng.uniform(-1, 1, 512).astype(np.float32), 512).copy()
Xs[0] = Xs[0].real + 5j      # imag of DC / Nyquist must be ignored (kiss_fftr.c:131-132)
Xs[256] = Xs[256].real - 3j
d["X_dcny"] = Xs
d["y_dcny"] = R.irfft(Xs, 512)
Xn = Xs.copy(); Xn[10] = np.nan + 0j
d["X_nanbin"] = Xn
d["y_nanbin"] = R.irfft(Xn, 512)
# batch=4 stride=2 layout (tests/fft_test.cc:384-495)
xb = rng.uniform(-1, 1, 4 * 2 * 512).astype(np.float32)
d["xb"] = xb
d["Xb"] = R.rfft(xb, 512, batch=4, stride_in=2, stride_out=2, out_len=4 * 2 * 257)
save("fft_kat.npz", **d)

# 3. envelopes ------------------------------------------------------------------------------
d = {}
for N, H, t in [(1024, 256, 0), (2048, 512, 0), (4096, 1024, 0), (256, 64, 1), (1024, 512, 0), (512, 512, 0), (1000, 300, 2)]:
    w = R.window(t, N)
    ring = R.ring_len(N, H)
    d[f"norm_{N}_{H}_{t}"] = R.build_norm_linear(w, ring, N, H)
save("norm.npz", **d)

# 4. OLA streaming drive --------------------------------------------------------------------
d = {}
for tag, (N, H, Cn, inside, setw, passw, g) in {
    "a": (256, 64, 2, 1, 1, 0, 0.7), "b": (256, 64, 1, 0, 1, 1, 1.0), "c": (256, 64, 1, 0, 0, 0, 1.5),
    "d": (256, 256, 1, 1, 1, 0, 1.0), "e": (512, 128, 3, 1, 1, 0, 1.0)}.items():
    F = 45
    fr = rng.uniform(-1, 1, (F, Cn, N)).astype(np.float32)
    w = R.window(1, N)
    out, n, peak = R.ola_stream(fr, N, H, Cn, w, setw, passw, inside, gain=g)
    d[f"{tag}_cfg"] = np.array([N, H, Cn, inside, setw, passw], np.int64)
    d[f"{tag}_gain"] = np.float32(g)
    d[f"{tag}_frames"] = fr
    d[f"{tag}_window"] = w
    d[f"{tag}_out"] = out
    d[f"{tag}_n"] = np.int64(n)
    d[f"{tag}_peak"] = np.float32(peak)
save("ola.npz", **d)

# 5. round trips, small ---------------------------------------------------------------------
d = {}
for tag, (N, H, L, t, mode) in {"rt1024": (1024, 256, 9000, 0, po.DROP), "rt2048": (2048, 512, 15000, 0, po.DROP),
                                 "rt4096": (4096, 1024, 24000, 0, po.DROP), "rt256": (256, 64, 3000, 1, po.DROP),
                                 "rt256zp": (256, 64, 3000, 1, po.ZERO_PAD)}.items():
    x = rng.uniform(-1, 1, L).astype(np.float32)
    w = R.window(t, N)
    r = R.roundtrip(x, N, H, w, mode)
    rf = RF.roundtrip(x, N, H, w, mode)
    d[f"{tag}_cfg"] = np.array([N, H, L, t, mode], np.int64)
    d[f"{tag}_x"] = x
    d[f"{tag}_spectra"] = r["spectra"]
    d[f"{tag}_iframes"] = r["iframes"]
    d[f"{tag}_out"] = r["out"]
    d[f"{tag}_fast_vs_strict_rel"] = np.array([rel(rf["spectra"], r["spectra"]), rel(rf["out"], r["out"])])
save("roundtrip_small.npz", **d)

# 6. config 1: oboe.wav mono, N=1024 H=256 Hann, DROP ---------------------------------------
with wave.open("/root/reference/assets/oboe.wav", "rb") as wf:
    assert wf.getnchannels() == 2 and wf.getsampwidth() == 2
    pcm = np.frombuffer(wf.readframes(wf.getnframes()), dtype="<i2").reshape(-1, 2).astype(np.int32)
lr = pcm[:, 0] + pcm[:, 1]
x = ((pcm[:, 0].astype(np.float32) / 32768.0) + (pcm[:, 1].astype(np.float32) / 32768.0)) / 2.0
assert np.array_equal(x, (lr.astype(np.float32) / 65536.0))     # lossless re-expression
N, H = 1024, 256
w = R.window(0, N)
r = R.roundtrip(x, N, H, w, po.DROP)
rf = RF.roundtrip(x, N, H, w, po.DROP)
out = r["out"]
blk = 4096
nblk = out.size // blk
pick = np.array([0, 1, 555, 1109, 1110])
steady = slice(N, out.size - N)
snr = 10 * np.log10(np.sum(x[: out.size][steady].astype(np.float64) ** 2) /
                    np.sum((x[: out.size][steady].astype(np.float64) - out[steady]) ** 2))
save("oboe_c1.npz", lr_sum=lr.astype(np.int32), n_frames=np.int64(r["n_frames"]),
     out_head=out[:16384], out_tail=out[-4096:],
     out_block_sum=out[: nblk * blk].astype(np.float64).reshape(nblk, blk).sum(1),
     out_block_energy=(out[: nblk * blk].astype(np.float64) ** 2).reshape(nblk, blk).sum(1),
     out_len=np.int64(out.size), spec_pick_idx=pick, spec_pick=r["spectra"][pick],
     iframes_pick=r["iframes"][pick], snr_db_vs_input=np.float64(snr),
     fast_vs_strict_rel=np.array([rel(rf["spectra"], r["spectra"]), rel(rf["out"], out)]))
print("oboe: frames", r["n_frames"], "out", out.size, "reference round-trip SNR vs input (dB)", snr)
