"""Partitioning of the hot path across the GPUs of one box -- Python face of the C-ABI
entry points in csrc/shard.cu (crlot_channel_shard, crlot_shard_plan, crlot_shard_group_*,
crlot_halo_exchange, crlot_sharded_roundtrip_execute).  One implementation: everything here forwards to them.

* channel sharding (config 3, config 5): channels / streams never interact
  (reference: dsp/ola/OLAAccumulator.cc:76-111 loops channels independently) -> contiguous
  channel blocks, no data-path collective.
* time sharding (config 4): rank r owns the OUTPUT range [t0, t1).  Output sample t depends
  only on frames k with k*hop in (t - nfft, t], and frame k only on samples [k*hop, k*hop+nfft),
  so a rank needs (nfft - hop) input samples from BOTH neighbours and recomputes the
  ceil(nfft/hop)-1 frames that start in its left halo.  Every output sample then sums the
  same frames in the same ascending order as a single-GPU run: the sharded result is
  bit-identical to the unsharded one, and the only communication is one neighbour exchange of
  (nfft - hop) * 4 bytes per direction per boundary (ncclSend/ncclRecv inside one group; latency-bound).

The halo exchange has two transports behind one function: the C-ABI's NCCL group for CUDA buffers
(`ShardGroup`), and torch.distributed point-to-point for CPU tensors under gloo -- the latter exists so the
multi-rank host logic (who sends what to whom, zero-length halves skipped on both sides) is testable without
GPUs (tests/test_sharding_gloo.py); both take their counts from the same crlot_time_shard.
"""
from __future__ import annotations

import ctypes as C

from . import capi

ZERO_PAD, DROP = 0, 1


def frame_count(length, frame, hop, mode):
    return capi.frame_count(length, frame, hop, mode)


def channel_shard(channels: int, rank: int, world: int):
    """Contiguous block of channels for `rank`: (first_channel, n_channels)."""
    first, count = C.c_longlong(0), C.c_longlong(0)
    capi._check(capi.lib().crlot_channel_shard(C.c_longlong(channels), C.c_int(rank), C.c_int(world),
                                               C.byref(first), C.byref(count)))
    return first.value, count.value


class TimeShard(C.Structure):
    """crlot_time_shard (include/crlot_cuda.h), field for field."""
    _fields_ = [("rank", C.c_int), ("world", C.c_int), ("nfft", C.c_int), ("hop", C.c_int), ("boundary_mode", C.c_int),
                ("total_len", C.c_longlong),
                ("t0", C.c_longlong), ("t1", C.c_longlong), ("n_out", C.c_longlong),
                ("k_lo", C.c_longlong), ("n_frames", C.c_longlong), ("n_owned_frames", C.c_longlong),
                ("in_lo", C.c_longlong), ("in_hi", C.c_longlong), ("own_lo", C.c_longlong), ("own_hi", C.c_longlong),
                ("left_halo", C.c_longlong), ("right_halo", C.c_longlong),
                ("send_left", C.c_longlong), ("send_right", C.c_longlong), ("frame_t0", C.c_longlong)]

    @property
    def buffer_len(self):
        """samples of the [left halo | owned | right halo] buffer"""
        return self.left_halo + (self.own_hi - self.own_lo) + self.right_halo


def make_time_shard(total_len, nfft, hop, rank, world, mode=DROP) -> TimeShard:
    sh = TimeShard()
    capi._check(capi.lib().crlot_shard_plan(C.c_longlong(total_len), C.c_int(nfft), C.c_int(hop), C.c_int(rank),
                                            C.c_int(world), C.c_int(mode), C.byref(sh)))
    return sh


time_shard = make_time_shard


class ShardGroup:
    """crlot_shard_group_*: this rank's NCCL communicator for the halo exchange.  The 128-byte unique id
    travels from rank 0 over torch.distributed (any transport would do)."""

    def __init__(self, rank, world, device, dist=None):
        self._h = C.c_void_p()
        self.rank, self.world, self.device = rank, world, device
        ident = (C.c_char * 128)()
        if world > 1:
            import torch
            if rank == 0:
                capi._check(capi.lib().crlot_shard_group_unique_id(ident))
            box = [bytes(ident.raw) if rank == 0 else None]
            (dist or __import__("torch.distributed").distributed).broadcast_object_list(box, src=0)
            ident = (C.c_char * 128).from_buffer_copy(box[0])
            del torch
        capi._check(capi.lib().crlot_shard_group_create(C.c_int(world), C.c_int(rank), ident, C.c_int(device),
                                                        C.byref(self._h)))

    def close(self):
        if self._h:
            capi.lib().crlot_shard_group_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def halo_exchange(xbuf, sh: TimeShard, group=None):
    """Fill the halo regions of `xbuf` (a 1-D tensor laid out [left halo | owned | right halo]) from the
    neighbouring ranks.  group: a ShardGroup (CUDA buffers: crlot_halo_exchange on the current stream), or
    None / a torch.distributed process group (CPU tensors under gloo)."""
    if isinstance(group, ShardGroup):
        capi._check(capi.lib().crlot_halo_exchange(group._h, C.byref(sh), C.c_void_p(xbuf.data_ptr()),
                                                   capi._stream_ptr(group.device)))
        return
    import torch.distributed as dist
    L, own = sh.left_halo, sh.own_hi - sh.own_lo
    ops = []
    if sh.rank > 0:
        if sh.send_left:
            ops.append(dist.P2POp(dist.isend, xbuf[L:L + sh.send_left], sh.rank - 1, group))
        if L:
            ops.append(dist.P2POp(dist.irecv, xbuf[:L], sh.rank - 1, group))
    if sh.rank < sh.world - 1:
        if sh.send_right:
            ops.append(dist.P2POp(dist.isend, xbuf[L + own - sh.send_right:L + own], sh.rank + 1, group))
        if sh.right_halo:
            ops.append(dist.P2POp(dist.irecv, xbuf[L + own:L + own + sh.right_halo], sh.rank + 1, group))
    if ops:
        for r in dist.batch_isend_irecv(ops):
            r.wait()


def sharded_roundtrip(plan, group, sh: TimeShard, xbuf, spec, frames, out):
    """crlot_sharded_roundtrip_execute: halo exchange + K1 + (K2, K3 | fused when frames is None) on the current stream."""
    capi._check(capi.lib().crlot_sharded_roundtrip_execute(
        plan._h, group._h if group is not None else None, C.byref(sh), C.c_void_p(xbuf.data_ptr()),
        C.c_void_p(spec.data_ptr()), C.c_void_p(frames.data_ptr() if frames is not None else 0),
        C.c_void_p(out.data_ptr()), capi._stream_ptr(plan.device)))
    return out
