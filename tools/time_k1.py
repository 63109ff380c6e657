"""Time kernel 1 alone (CUDA events, device-resident input): python tools/time_k1.py [N H [log2len [reps]]]
CRLOT_K1_NO_TMEM=1 selects the shared-memory-table variant (A/B in two processes)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import crlot_dsp_b200 as dsp

N = int(sys.argv[1]) if len(sys.argv) > 1 else 2048
H = int(sys.argv[2]) if len(sys.argv) > 2 else 512
LG = int(sys.argv[3]) if len(sys.argv) > 3 else 30
reps = int(sys.argv[4]) if len(sys.argv) > 4 else 20
ch = max(1, (1 << 30) >> LG) if LG < 30 else 1
w = dsp.window_make(dsp.HANN, N)
plan = dsp.Plan(N, H, analysis_window=w, ola_window=w)
x = dsp.synth_uniform(ch, 1 << LG, seed=3)
F = plan.n_frames(1 << LG)
spec = torch.empty((ch, F, plan.bins), dtype=torch.complex64, device="cuda")
for _ in range(3):
    plan.stft(x, out=spec)
torch.cuda.synchronize()
ts = []
for _ in range(reps):
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    plan.stft(x, out=spec)
    b.record()
    torch.cuda.synchronize()
    ts.append(a.elapsed_time(b))
alg = (4.0 * H + 8.0 * plan.bins) * F * ch
ts.sort()
print(f"K1 N={N} H={H} ch={ch} 2^{LG}: best {ts[0]:.3f} ms  median {ts[len(ts)//2]:.3f} ms  -> {alg / ts[len(ts)//2] / 1e6:.0f} GB/s "
      f"tmem={'off' if os.environ.get('CRLOT_K1_NO_TMEM') else 'on'}  checksum {float(torch.view_as_real(spec).abs().sum()):.6e}")
