"""Time-sharded path (config 4) on ONE GPU: the signal is cut into `world` logical shards exactly as
bench.py / the multi-rank path lays them out -- [left halo | owned | right halo] buffers, halos filled by
slicing (what the NCCL neighbour exchange delivers), kernel 1 over the haloed buffer, kernel 3 / the fused
kernel 2+3 at the shard's t_offset (envelope phase) and frame_t0 (recomputed left-halo frames) -- and the
concatenation must be BIT-IDENTICAL to the unsharded GPU run and within 1e-5 rel-L2 of the oracle.

Reference data dependences this rests on: frame k reads only x[kH, kH+N)
(ref: dsp/frame/framer.cc:128-182); accumulator cells never interact across frames beyond the
N-sample span and the envelope is a function of the absolute sample index mod ring_len
(ref: dsp/ola/OLAAccumulator.cc:75-111,194-209)."""
import numpy as np
import pyoracle as po
import pytest
from conftest import bits, rel_l2

pytestmark = pytest.mark.gpu
TOL = 1e-5


def _run_sharded(cuda, dsp, plan, x_full, N, H, world, mode, fused):
    from crlot_dsp_b200 import sharding
    total = x_full.numel()
    outs, specs, owned = [], [], 0
    for r in range(world):
        sh = sharding.make_time_shard(total, N, H, r, world, mode)
        if sh.n_frames == 0:
            continue
        own = sh.own_hi - sh.own_lo
        xbuf = cuda.zeros(sh.left_halo + own + sh.right_halo, dtype=cuda.float32, device="cuda")
        xbuf[sh.left_halo:sh.left_halo + own] = x_full[sh.own_lo:sh.own_hi]
        # the halo exchange, by slicing: what rank r-1 / r+1 would send
        if sh.left_halo:
            xbuf[:sh.left_halo] = x_full[sh.in_lo:sh.own_lo]
        if sh.right_halo > 0:
            xbuf[sh.left_halo + own:] = x_full[sh.own_hi:sh.in_hi]
        length = sh.in_hi - sh.in_lo
        spec = plan.stft(xbuf.unsqueeze(0), length=length, n_frames=sh.n_frames)
        if fused:
            assert plan.istft_ola_supported(t_offset=sh.t0, frame_t0=sh.frame_t0)
            out = plan.istft_ola(spec, t_offset=sh.t0, frame_t0=sh.frame_t0, n_out=sh.n_out)
        else:
            fr = plan.istft_frames(spec, fuse_synthesis_window=True)
            out = plan.ola(fr, prewindowed=True, t_offset=sh.t0, frame_t0=sh.frame_t0, n_out=sh.n_out)
        outs.append((sh, out[0].cpu().numpy()))
        specs.append((sh, spec[0].cpu().numpy()))
        owned += sh.n_owned_frames
    return outs, specs, owned


@pytest.mark.parametrize("N,H", [(1024, 256), (4096, 1024), (2048, 512), (256, 64)])
@pytest.mark.parametrize("world", [4, 8])
def test_logical_shards_equal_unsharded_and_oracle(cuda, dsp, oracle, N, H, world):
    ring = dsp.ola_ring_len(N, H)
    S = 2 * ring + 3 * H                  # not a multiple of ring_len: every shard has its own envelope phase
    total = world * S
    w = oracle.window(po.HANN, N)
    plan = dsp.Plan(N, H, analysis_window=w, ola_window=w, boundary_mode=dsp.DROP)
    x = dsp.synth_uniform(1, total, seed=0xC0FFEE)[0]
    full_out, full_spec, full_frames = plan.roundtrip(x)
    cuda.cuda.synchronize()
    full_out, full_spec = full_out[0].cpu().numpy(), full_spec[0].cpu().numpy()
    F = plan.n_frames(total)

    ref = oracle.roundtrip(x.cpu().numpy(), N, H, w, po.DROP)
    assert rel_l2(full_out, ref["out"]) < TOL

    for fused in (False, True):
        outs, specs, owned = _run_sharded(cuda, dsp, plan, x, N, H, world, dsp.DROP, fused)
        cuda.cuda.synchronize()
        assert owned == F                                              # every frame owned exactly once
        cat = np.concatenate([o for _, o in outs])
        assert cat.size == full_out.size
        # bit-identity with the unsharded GPU run, shard by shard (so a failure names the rank)
        for sh, o in outs:
            assert np.array_equal(bits(o), bits(full_out[sh.t0:sh.t1])), (fused, sh.rank)
        # every locally computed spectrum row (owned and recomputed halo rows) equals the unsharded row
        for sh, s in specs:
            assert np.array_equal(bits(s), bits(full_spec[sh.k_lo:sh.k_lo + sh.n_frames])), (fused, sh.rank)
        # and the concatenation is within the bar of the oracle; each rank's slice on its own as well
        assert rel_l2(cat, ref["out"]) < TOL
        for sh, o in outs:
            assert rel_l2(o, ref["out"][sh.t0:sh.t1]) < TOL, (fused, sh.rank)


def test_logical_shards_zero_pad_tail(cuda, dsp, oracle):
    """ZERO_PAD framing: the last rank also owns the zero-tailed frames past the end of the signal."""
    N, H, world = 1024, 256, 4
    S = 40 * H
    total = world * S
    w = oracle.window(po.HAMMING, N)
    plan = dsp.Plan(N, H, analysis_window=w, ola_window=w, boundary_mode=dsp.ZERO_PAD)
    x = dsp.synth_uniform(1, total, seed=11)[0]
    full_out, _, _ = plan.roundtrip(x)
    full_out = full_out[0].cpu().numpy()
    ref = oracle.roundtrip(x.cpu().numpy(), N, H, w, po.ZERO_PAD)
    assert rel_l2(full_out, ref["out"]) < TOL
    outs, _, owned = _run_sharded(cuda, dsp, plan, x, N, H, world, dsp.ZERO_PAD, False)
    assert owned == plan.n_frames(total)
    cat = np.concatenate([o for _, o in outs])
    assert np.array_equal(bits(cat), bits(full_out))


def test_slice_oracle_at_rank_offset(cuda, dsp, oracle):
    """The per-rank check bench.py's sharded c4 leg uses: a rank far into a long signal compares its
    output with a SLICE oracle -- a fresh reference accumulator started at a sample offset that is a
    multiple of ring_len reproduces the full-run output after its first N samples."""
    from crlot_dsp_b200 import sharding
    N, H, world = 4096, 1024, 8
    S = 1 << 22
    total = world * S
    w = oracle.window(po.HANN, N)
    plan = dsp.Plan(N, H, analysis_window=w, ola_window=w, boundary_mode=dsp.DROP)
    for r in (1, 5, 7):
        sh = sharding.make_time_shard(total, N, H, r, world)
        L = sh.in_hi - sh.in_lo
        xbuf = dsp.synth_uniform(1, L, seed=0xC0FFEE, i0=sh.in_lo)
        spec = plan.stft(xbuf, length=L, n_frames=sh.n_frames)
        out = plan.istft_ola(spec, t_offset=sh.t0, frame_t0=sh.frame_t0, n_out=sh.n_out)[0]
        a0, want = oracle.slice_roundtrip(0xC0FFEE, 0, sh.t0, 1 << 16, N, H, w, plan.ring_len)
        got = out[a0 - sh.t0:a0 - sh.t0 + want.size].cpu().numpy()
        assert rel_l2(got, want) < TOL, r
