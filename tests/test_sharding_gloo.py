"""Multi-rank host logic on CPU: world_size-2 (and 4) gloo process groups run the time-sharded
path end to end -- shard arithmetic, neighbour halo exchange (torch.distributed send/recv), local
compute, envelope phase -- with the CPU oracle standing in for the three kernels, and compare
every rank's output slice bit-for-bit with the unsharded run.  Also the channel-shard split."""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, total_len, nfft, hop, q):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    import torch
    import torch.distributed as dist
    import pyoracle as po
    from crlot_dsp_b200 import sharding
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        orc = po.Oracle()
        w = orc.window(po.HANN, nfft)
        sh = sharding.make_time_shard(total_len, nfft, hop, rank, world)
        own = sh.own_hi - sh.own_lo
        xbuf = torch.zeros(sh.left_halo + own + sh.right_halo, dtype=torch.float32)
        xbuf[sh.left_halo:sh.left_halo + own] = torch.from_numpy(orc.synth_uniform(7, 0, sh.own_lo, own))
        sharding.halo_exchange(xbuf, sh)
        local = xbuf.numpy()
        # what the halo exchange must have delivered
        need = sh.in_hi - sh.in_lo            # (the last rank's owned tail beyond the last DROP frame is not needed)
        assert np.array_equal(local[:need], orc.synth_uniform(7, 0, sh.in_lo, need))
        # local "kernels": frames -> window -> rfft -> irfft -> gather OLA at the right envelope phase
        frames = np.stack([local[k * hop:k * hop + nfft] for k in range(sh.n_frames)]) * w
        ifr = orc.irfft_many(orc.rfft_many(frames.astype(np.float32), nfft), nfft)
        env = orc.ola_envelope(w, nfft, hop)
        out = orc.ola_gather(ifr, nfft, hop, w, env, first_frame=sh.k_lo, t0=sh.t0, n_out=sh.n_out)
        # the unsharded run
        full = orc.roundtrip(orc.synth_uniform(7, 0, 0, total_len), nfft, hop, w, po.DROP)
        ok = np.array_equal(out.view(np.uint32), full["out"][sh.t0:sh.t1].view(np.uint32))
        spec_ok = np.array_equal(orc.rfft_many(frames.astype(np.float32), nfft).view(np.uint64),
                                 full["spectra"][sh.k_lo:sh.k_lo + sh.n_frames].view(np.uint64))
        q.put((rank, bool(ok), bool(spec_ok), sh.n_owned_frames, sh.n_out))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,nfft,hop", [(2, 1024, 256), (2, 4096, 1024), (4, 256, 64), (2, 256, 256), (3, 1000, 300)])
def test_time_sharded_path_equals_unsharded(world, nfft, hop):
    import torch.multiprocessing as mp
    total_len = world * 24 * hop * -(-nfft // hop)
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    import socket
    with socket.socket() as sk:                   # a free port per case: consecutive cases must not collide
        sk.bind(("127.0.0.1", 0))
        port = sk.getsockname()[1]
    procs = [ctx.Process(target=_worker, args=(r, world, port, total_len, nfft, hop, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=240) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    F = (total_len - nfft) // hop + 1
    assert all(ok and sok for _, ok, sok, _, _ in res), res
    assert sum(n for *_, n, _ in res) == F                       # every frame owned exactly once
    assert sum(n for *_, n in res) == (F - 1) * hop + nfft       # every output sample produced exactly once


def test_shard_arithmetic():
    sys.path.insert(0, ROOT)
    from crlot_dsp_b200 import sharding
    # config 4: 2^33 samples, frame 4096 hop 1024, 8 ranks: halo = 3072 samples = 12 288 bytes per direction
    tot = 1 << 33
    shards = [sharding.make_time_shard(tot, 4096, 1024, r, 8) for r in range(8)]
    assert shards[0].left_halo == 0 and shards[7].right_halo == 0
    assert all(s.left_halo == 3072 for s in shards[1:]) and all(s.right_halo == 3072 for s in shards[:7])
    assert sum(s.n_owned_frames for s in shards) == 8388605
    assert all(s.frame_t0 == (-3072 if s.rank else 0) for s in shards)
    assert [s.t0 for s in shards] == [r << 30 for r in range(8)]
    assert all(s.send_left == (3072 if s.rank else 0) and s.send_right == (3072 if s.rank < 7 else 0) for s in shards)
    # hop == nfft: no overlap, no halos, nothing to send in either direction (sends and receives stay symmetric)
    flat = [sharding.make_time_shard(8 * 4096, 512, 512, r, 4) for r in range(4)]
    assert all(s.left_halo == s.right_halo == s.send_left == s.send_right == 0 for s in flat)
    # a signal shorter than one frame: every shard is empty, with empty halos
    empty = [sharding.make_time_shard(4 * 64, 1024, 64, r, 4) for r in range(4)]
    assert all(s.n_frames == 0 and s.n_out == 0 and s.left_halo == 0 and s.right_halo == 0 for s in empty)
    # halo larger than a neighbour's owned span is refused, not silently mis-sliced
    import crlot_dsp_b200 as dsp
    with pytest.raises(dsp.CrlotInvalidArgument):
        sharding.make_time_shard(4 * 256, 1024, 256, 1, 4)
    with pytest.raises(dsp.CrlotInvalidArgument):
        sharding.make_time_shard(1000, 256, 64, 0, 3)          # does not split evenly
    # channel blocks
    assert [sharding.channel_shard(256, r, 8) for r in range(8)] == [(32 * r, 32) for r in range(8)]
    blocks = [sharding.channel_shard(10, r, 4) for r in range(4)]
    assert blocks == [(0, 3), (3, 3), (6, 2), (8, 2)]
    assert sum(n for _, n in blocks) == 10
