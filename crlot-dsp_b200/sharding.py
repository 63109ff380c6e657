"""Partitioning of the hot path across the GPUs of one box.

* channel sharding (config 3, config 5): channels / streams never interact
  (reference: dsp/ola/OLAAccumulator.cc:76-111 loops channels independently) -> contiguous
  channel blocks, no data-path collective.
* time sharding (config 4): rank r owns the OUTPUT range [t0, t1).  Output sample t depends
  only on frames k with k*hop in (t - nfft, t], and frame k only on samples [k*hop, k*hop+nfft),
  so a rank needs (nfft - hop) input samples from BOTH neighbours and recomputes the
  ceil(nfft/hop)-1 frames that start in its left halo.  Every output sample then sums the
  same frames in the same ascending order as a single-GPU run: the sharded result is
  bit-identical to the unsharded one, and the only communication is one neighbour exchange of
  (nfft - hop) * 4 bytes per direction per boundary (NCCL send/recv; latency-bound).

Pure host logic + torch.distributed point-to-point plumbing; works with the gloo backend on
CPU tensors (tests/test_sharding_gloo.py) and the nccl backend on CUDA tensors (bench.py).
"""
from __future__ import annotations

from dataclasses import dataclass

ZERO_PAD, DROP = 0, 1


def frame_count(length, frame, hop, mode):
    if length <= 0:
        return 0
    if mode == DROP:
        return 0 if length < frame else (length - frame) // hop + 1
    return (length + hop - 1) // hop


def channel_shard(channels: int, rank: int, world: int):
    """Contiguous block of channels for `rank`: (first_channel, n_channels)."""
    base, rem = divmod(channels, world)
    first = rank * base + min(rank, rem)
    return first, base + (1 if rank < rem else 0)


@dataclass
class TimeShard:
    rank: int
    world: int
    t0: int            # first output sample owned (absolute)
    t1: int            # one past the last output sample owned
    k_lo: int          # first global frame this rank computes
    n_frames: int      # frames computed locally (incl. the recomputed left-halo frames)
    in_lo: int         # first input sample needed (absolute)
    in_hi: int         # one past the last input sample needed (clipped to the signal)
    own_lo: int        # input samples [own_lo, own_hi) are generated / held by this rank
    own_hi: int
    n_owned_frames: int  # frames whose start lies in [own_lo, own_hi): the spectra this rank owns

    @property
    def left_halo(self):
        return self.own_lo - self.in_lo

    @property
    def right_halo(self):
        return self.in_hi - self.own_hi

    @property
    def n_out(self):
        return self.t1 - self.t0

    @property
    def frame_t0(self):
        """local output index of the first sample of local frame 0 (<= 0)."""
        return self.k_lo * self.hop - self.t0

    hop: int = 0


def time_shard(total_len: int, nfft: int, hop: int, rank: int, world: int, mode: int = DROP) -> TimeShard:
    assert total_len % world == 0, "signal length must split evenly over the ranks"
    S = total_len // world
    assert S % hop == 0, "shard length must be a multiple of hop"
    F = frame_count(total_len, nfft, hop, mode)
    n_out_total = (F - 1) * hop + nfft if F > 0 else 0
    t0 = min(rank * S, n_out_total)
    t1 = n_out_total if rank == world - 1 else min((rank + 1) * S, n_out_total)
    if t1 <= t0 or F == 0:
        return TimeShard(rank, world, t0, t0, 0, 0, rank * S, rank * S, rank * S, (rank + 1) * S, 0, hop)
    k_lo = max(0, -(-(t0 - nfft + 1) // hop))
    k_hi = min(F - 1, (t1 - 1) // hop)
    in_lo = k_lo * hop
    in_hi = min(total_len, k_hi * hop + nfft)
    own_lo, own_hi = rank * S, (rank + 1) * S
    first_owned = max(k_lo, -(-own_lo // hop))
    last_owned = min(k_hi, (own_hi - 1) // hop)
    owned = max(0, last_owned - first_owned + 1)
    return TimeShard(rank, world, t0, t1, k_lo, k_hi - k_lo + 1, in_lo, in_hi, own_lo, own_hi, owned, hop)


def halo_exchange(xbuf, sh: TimeShard, group=None):
    """Fill the halo regions of `xbuf` (a 1-D tensor laid out [left halo | owned | right halo]) from
    the neighbouring ranks: one batched send/recv per boundary and direction."""
    import torch.distributed as dist
    L, own = sh.left_halo, sh.own_hi - sh.own_lo
    ops = []
    if sh.rank > 0:                                   # left neighbour: needs my first (its right halo) samples
        n_send = time_shard_right_halo(sh, sh.rank - 1)
        ops.append(dist.P2POp(dist.isend, xbuf[L:L + n_send], sh.rank - 1, group))
        if L:
            ops.append(dist.P2POp(dist.irecv, xbuf[:L], sh.rank - 1, group))
    if sh.rank < sh.world - 1:                        # right neighbour: needs my last (its left halo) samples
        n_send = time_shard_left_halo(sh, sh.rank + 1)
        ops.append(dist.P2POp(dist.isend, xbuf[L + own - n_send:L + own], sh.rank + 1, group))
        if sh.right_halo:
            ops.append(dist.P2POp(dist.irecv, xbuf[L + own:L + own + sh.right_halo], sh.rank + 1, group))
    if ops:
        for r in dist.batch_isend_irecv(ops):
            r.wait()


# the halo sizes of a neighbour follow from the same arithmetic; kept on the shard so both sides agree
def time_shard_left_halo(sh: TimeShard, other_rank: int) -> int:
    return _peer(sh, other_rank).left_halo


def time_shard_right_halo(sh: TimeShard, other_rank: int) -> int:
    return _peer(sh, other_rank).right_halo


def _peer(sh: TimeShard, other_rank: int) -> TimeShard:
    return time_shard(sh._total_len, sh._nfft, sh.hop, other_rank, sh.world, sh._mode)


def make_time_shard(total_len, nfft, hop, rank, world, mode=DROP) -> TimeShard:
    sh = time_shard(total_len, nfft, hop, rank, world, mode)
    sh._total_len, sh._nfft, sh._mode = total_len, nfft, mode
    return sh
