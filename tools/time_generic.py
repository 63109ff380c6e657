"""Times the any-size kernels (csrc/k_generic.cu) on device-resident input: K1 + K2 + K3 round trip for
frame sizes outside the register-FFT set, next to the register-FFT path at a neighbouring size.
Prints one JSON line per shape.  usage: python tools/time_generic.py [log2_samples]"""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import crlot_dsp_b200 as dsp  # noqa: E402


def hann(n):
    return dsp.window_make(dsp.HANN, n)


def main():
    L = 1 << (int(sys.argv[1]) if len(sys.argv) > 1 else 26)
    x = dsp.synth_uniform(1, L, seed=7)
    for N, H in ((480, 160), (1000, 250), (1024, 250), (1024, 256), (6000, 1500), (8192, 2048)):
        w = hann(N)
        plan = dsp.Plan(N, H, analysis_window=w, ola_window=w, boundary_mode=dsp.DROP)
        F = plan.n_frames(L)
        spec = torch.empty((1, F, plan.bins), dtype=torch.complex64, device="cuda")
        frames = torch.empty((1, F, N), dtype=torch.float32, device="cuda")
        out = torch.empty((1, plan.out_len(L)), dtype=torch.float32, device="cuda")
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
        ms = np.zeros(3)
        reps = 5
        for it in range(reps + 2):
            ev[0].record()
            plan.stft(x, out=spec)
            ev[1].record()
            plan.istft_frames(spec, fuse_synthesis_window=True, out=frames)
            ev[2].record()
            plan.ola(frames, prewindowed=True, out=out)
            ev[3].record()
            torch.cuda.synchronize()
            if it >= 2:
                ms += [ev[i].elapsed_time(ev[i + 1]) for i in range(3)]
        ms /= reps
        print(json.dumps({"nfft": N, "hop": H, "samples": L, "frames": F, "k1_ms": ms[0], "k2_ms": ms[1], "k3_ms": ms[2],
                          "Msamples_per_s": L / ms.sum() / 1e3,
                          "path": "register FFT" if (N in (256, 512, 1024, 2048, 4096) and H % 4 == 0) else "any-size"}), flush=True)


if __name__ == "__main__":
    main()
