"""Stream-safety of the C-ABI objects (round-1 advisor findings).

* One plan, several streams at once: kernel 1 and the fused kernel 2+3 pull their work from a counter that
  each launch zeroes on its own stream; launches that overlap on different streams must not share a counter
  word (csrc/capi.cu crlot_plan_s::counter_for).  Results must equal the serial ones bit for bit.
* The device Framer mixes the legacy default stream (push / pop) with the caller's, possibly non-blocking,
  stream (push_device / stft): its operations are ordered by events (csrc/framer.cu order_after_last), so a
  compaction never overwrites samples an earlier kernel 1 on another stream still reads."""
import numpy as np
import pyoracle as po
import pytest
from conftest import bits

pytestmark = pytest.mark.gpu


def test_one_plan_on_several_streams_at_once(cuda, dsp, oracle):
    torch = cuda
    N, H = 2048, 512
    w = oracle.window(po.HANN, N)
    plan = dsp.Plan(N, H, analysis_window=w, ola_window=w)
    L = 1 << 22
    xs = [dsp.synth_uniform(1, L, seed=100 + i) for i in range(4)]
    want = []
    for x in xs:
        o, s = plan.roundtrip_fused(x)
        want.append((o.clone(), s.clone()))
    torch.cuda.synchronize()
    streams = [torch.cuda.Stream() for _ in xs]            # non-blocking side streams
    for rep in range(6):
        got = []
        for x, st in zip(xs, streams):
            with torch.cuda.stream(st):
                got.append(plan.roundtrip_fused(x))
        torch.cuda.synchronize()
        for i, ((o, s), (wo, ws)) in enumerate(zip(got, want)):
            assert torch.equal(o, wo), (rep, i)
            assert torch.equal(torch.view_as_real(s), torch.view_as_real(ws)), (rep, i)


def test_framer_on_nonblocking_streams(cuda, dsp, oracle):
    torch = cuda
    N, H = 1024, 256
    w = oracle.window(po.HANN, N)
    plan = dsp.Plan(N, H, analysis_window=w, ola_window=w)
    rng = np.random.default_rng(2)
    total = rng.standard_normal(400 * H).astype(np.float32)
    fr = dsp.Framer(N, H, 1, dsp.DROP)
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
    specs, pos = [], 0
    # alternate host pushes (default stream), device pushes and kernel-1 drains on two non-blocking streams;
    # the store compacts / grows several times along the way
    for it in range(12):
        n = int(rng.integers(H, 40 * H))
        n = min(n, total.size - pos)
        if n <= 0:
            break
        chunk = total[pos:pos + n]
        if it % 2 == 0:
            fr.push(chunk)
        else:
            with torch.cuda.stream(s1):
                d = torch.from_numpy(chunk).cuda(non_blocking=False)
                fr.push_device(d)
        pos += n
        with torch.cuda.stream(s2 if it % 3 else s1):
            specs.append(fr.stft(plan))
    torch.cuda.synchronize()
    got = torch.cat([s[0] for s in specs], dim=0).cpu().numpy()
    F = (pos - N) // H + 1
    assert got.shape[0] == F
    want = plan.stft(torch.from_numpy(total[:pos]).cuda())[0].cpu().numpy()
    assert np.array_equal(bits(got), bits(want))
