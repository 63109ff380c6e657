"""Time kernel 2 alone: python tools/time_k2.py [N [log2len [reps]]]   (CRLOT_K2_NO_TMEM=1 for the A/B)"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import crlot_dsp_b200 as dsp

N = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
LG = int(sys.argv[2]) if len(sys.argv) > 2 else 30
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 20
H = N // 4
w = dsp.window_make(dsp.HANN, N)
plan = dsp.Plan(N, H, analysis_window=w, ola_window=w)
x = dsp.synth_uniform(1, 1 << LG, seed=3)
spec = plan.stft(x)
F = spec.shape[1]
frames = torch.empty((1, F, N), dtype=torch.float32, device="cuda")
for _ in range(3):
    plan.istft_frames(spec, True, out=frames)
torch.cuda.synchronize()
ts = []
for _ in range(reps):
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    plan.istft_frames(spec, True, out=frames)
    b.record()
    torch.cuda.synchronize()
    ts.append(a.elapsed_time(b))
ts.sort()
alg = (8.0 * plan.bins + 4.0 * N) * F
print(f"K2 N={N} 2^{LG}: best {ts[0]:.3f} ms median {ts[len(ts)//2]:.3f} ms -> {alg / ts[len(ts)//2] / 1e6:.0f} GB/s "
      f"tmem={'off' if os.environ.get('CRLOT_K2_NO_TMEM') else 'on'} checksum {float(frames.abs().sum()):.6e}")

if N == 4096 and LG <= 28:      # the two-sub-transform kernel against the first version (different roundings: tolerance, not bits)
    os.environ["CRLOT_K2_4096_V1"] = "1"
    f1 = torch.empty_like(frames)
    plan.istft_frames(spec, True, out=f1)
    torch.cuda.synchronize()
    del os.environ["CRLOT_K2_4096_V1"]
    d = (frames.double() - f1.double())
    print(f"v2 vs v1: rel-L2 {float(d.norm() / f1.double().norm()):.3e} max abs {float(d.abs().max()):.3e}")
