"""Bit-identity of the 16-warp tensor-memory kernel 1 (k1_2048tm.cu) with stft_fwd_kernel<2048> on a few shapes."""
import os, sys, torch
sys.path.insert(0, ".")
import crlot_dsp_b200 as dsp
for N, H in ((2048, 512), (1024, 256)):
  w = dsp.window_make(dsp.HANN, N)
  for mode, L, ch in ((dsp.DROP, (1 << 22) + 12344, 2), (dsp.ZERO_PAD, 70 * 512 + 76, 3), (dsp.ZERO_PAD, 70 * 512 + 77, 1), (dsp.DROP, 2048, 1), (dsp.ZERO_PAD, 5, 1)):
      plan = dsp.Plan(N, H, analysis_window=w, ola_window=w, boundary_mode=mode)
      x = dsp.synth_uniform(ch, L, seed=L)
      x[0, L // 3] = float("nan"); x[-1, L // 2] = 1e-35
      a = plan.stft(x).clone()
      os.environ["CRLOT_K1_NO_TM16"] = "1"
      b = plan.stft(x).clone()
      del os.environ["CRLOT_K1_NO_TM16"]
      same = torch.equal(torch.view_as_real(a).view(torch.int32), torch.view_as_real(b).view(torch.int32))
      print("N", N, "mode", mode, "L", L, "ch", ch, "frames", a.shape[1], "bit-identical", same, flush=True)
