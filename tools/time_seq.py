"""What a kernel costs inside the K1 -> K2 -> K3 sequence against back to back with itself (CUDA events per launch).
    python tools/time_seq.py [N H [log2len]]"""
import sys
import time

import torch

sys.path.insert(0, ".")
import crlot_dsp_b200 as dsp

N, H = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (2048, 512)
LG = int(sys.argv[3]) if len(sys.argv) > 3 else 30
w = dsp.window_make(dsp.HANN, N)
plan = dsp.Plan(N, H, analysis_window=w, ola_window=w)
x = dsp.synth_uniform(1, 1 << LG, seed=1)
F = plan.n_frames(1 << LG)
spec = torch.empty((1, F, plan.bins), dtype=torch.complex64, device="cuda")
frames = torch.empty((1, F, N), dtype=torch.float32, device="cuda")
out = torch.empty((1, plan.out_len(1 << LG)), dtype=torch.float32, device="cuda")
big = torch.empty(5 << 30, dtype=torch.float32, device="cuda")       # 20 GB

k1 = lambda: plan.stft(x, out=spec)
k2 = lambda: plan.istft_frames(spec, True, out=frames)
k3 = lambda: plan.ola(frames, prewindowed=True, out=out)


def timed(fn):
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    fn()
    b.record()
    return a, b


def run(label, seq, reps=10):
    """seq: list of (name, fn, timed?)"""
    for _ in range(2):
        for _, fn, _ in seq:
            fn()
    torch.cuda.synchronize()
    evs = {}
    for _ in range(reps):
        for name, fn, t in seq:
            if t:
                evs.setdefault(name, []).append(timed(fn))
            else:
                fn()
    torch.cuda.synchronize()
    res = {k: sorted(a.elapsed_time(b) for a, b in v) for k, v in evs.items()}
    print(label, {k: f"best {v[0]:.3f} med {v[len(v)//2]:.3f}" for k, v in res.items()}, flush=True)


run("K1 alone      ", [("k1", k1, True)])
run("K2 alone      ", [("k2", k2, True)])
run("K3 alone      ", [("k3", k3, True)])
run("K1,K2,K3      ", [("k1", k1, True), ("k2", k2, True), ("k3", k3, True)])
run("K3 then K1    ", [("k3", k3, False), ("k1", k1, True)])
run("K2 then K1    ", [("k2", k2, False), ("k1", k1, True)])
run("memset20G, K1 ", [("m", lambda: big.zero_(), False), ("k1", k1, True)])
run("K1 then K2    ", [("k1", k1, False), ("k2", k2, True)])
run("memset20G, K2 ", [("m", lambda: big.zero_(), False), ("k2", k2, True)])
run("sleep, K1     ", [("s", lambda: (torch.cuda.synchronize(), time.sleep(0.05)), False), ("k1", k1, True)])
run("K3, K1a, K1b  ", [("k3", k3, False), ("k1a", k1, True), ("k1b", k1, True)])
run("K2,K3,K1a,K1b ", [("k2", k2, False), ("k3", k3, False), ("k1a", k1, True), ("k1b", k1, True)])
run("K1 x3, K3a,K3b", [("k1", k1, False), ("k1", k1, False), ("k1", k1, False), ("k3a", k3, True), ("k3b", k3, True)])
