"""Quick device timing of K1 / K2 / K3 / fused K2+K3 on the c2 shape (debug helper, not the bench)."""
import sys
import torch
sys.path.insert(0, ".")
import crlot_dsp_b200 as dsp

N, H = int(sys.argv[1]) if len(sys.argv) > 1 else 2048, int(sys.argv[2]) if len(sys.argv) > 2 else 512
logL = int(sys.argv[3]) if len(sys.argv) > 3 else 29
ch = int(sys.argv[4]) if len(sys.argv) > 4 else 1
w = dsp.window_make(dsp.HANN, N)
plan = dsp.Plan(N, H, analysis_window=w, ola_window=w)
x = dsp.synth_uniform(ch, 1 << logL, seed=1)
F = plan.n_frames(1 << logL)
spec = torch.empty((ch, F, plan.bins), dtype=torch.complex64, device="cuda")
frames = torch.empty((ch, F, N), dtype=torch.float32, device="cuda")
out = torch.empty((ch, plan.out_len(1 << logL)), dtype=torch.float32, device="cuda")
out2 = torch.empty_like(out)


def timeit(fn, n=5):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / n


t1 = timeit(lambda: plan.stft(x, out=spec))
t2 = timeit(lambda: plan.istft_frames(spec, True, out=frames))
t3 = timeit(lambda: plan.ola(frames, prewindowed=True, out=out))
try:
    t23 = timeit(lambda: plan.istft_ola(spec, out=out2))
    same = torch.equal(out.view(torch.int32), out2.view(torch.int32))
except dsp.CrlotUnsupported:
    t23, same = float("nan"), None
tall = float("nan")
if plan.roundtrip_fused_all_supported():
    out3 = torch.empty_like(out)
    tall = timeit(lambda: plan.roundtrip_fused_all(x, out=out3))
    same = same and torch.equal(out.view(torch.int32), out3.view(torch.int32))
bins = N // 2 + 1
gb1 = ch * F * (4 * H + 8 * bins) / 1e9
gb23 = ch * F * (8 * bins + 4 * H) / 1e9
print(f"N={N} H={H} L=2^{logL} ch={ch} F={F}: K1 {t1:.3f} ms ({gb1/t1*1e3:.0f} GB/s)  K2 {t2:.3f}  K3 {t3:.3f}  "
      f"K2+K3 {t2+t3:.3f}  fused {t23:.3f} ms ({gb23/t23*1e3:.0f} GB/s alg)  bit-identical {same}")
print(f"round trip: 3-kernel {(1<<logL)*ch/(t1+t2+t3)/1e6:.1f} Gs/s   fused {(1<<logL)*ch/(t1+t23)/1e6:.1f} Gs/s   one kernel {tall:.3f} ms = {(1<<logL)*ch/tall/1e6:.1f} Gs/s")
