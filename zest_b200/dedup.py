"""Device-resident dedup index: Xet chunk hash -> 64-bit location (include/zest_gpu.h, zg_dedup_*).

Stands in, on the seed/dedup path, for zest's per-hash probes of the chunk cache (storage.hasChunk,
src/storage.zig:137) and the XorbRegistry scan (src/storage.zig:145-228): one launch answers a whole batch.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _native
from ._native import check
from .context import Context, default_context

MISSING = np.uint64(0xFFFFFFFFFFFFFFFF)


class DedupIndex:
    """`max_entries` distinct hashes at most.  Per hash the index keeps the smallest location ever inserted."""

    def __init__(self, max_entries: int, ctx: Context | None = None):
        self.ctx = ctx or default_context()
        self._h = C.c_void_p()
        check(_native.lib().zg_dedup_create(self.ctx.handle, int(max_entries), C.byref(self._h)))

    @property
    def handle(self):
        return self._h

    def insert(self, hashes, locations) -> np.ndarray:
        """-> was_new (bool per input): the hash was absent before this call and this occurrence's location was kept."""
        h = np.ascontiguousarray(hashes, dtype=np.uint8).reshape(-1, 32)
        loc = np.ascontiguousarray(locations, dtype=np.uint64)
        if loc.size != h.shape[0]:
            raise ValueError("hashes and locations differ in length")
        new = np.zeros(h.shape[0], np.uint8)
        check(_native.lib().zg_dedup_insert(self._h, h.ctypes.data, loc.ctypes.data, h.shape[0], new.ctypes.data))
        return new.astype(bool)

    def lookup(self, hashes) -> np.ndarray:
        """-> locations (u64), MISSING where the hash is not in the index."""
        h = np.ascontiguousarray(hashes, dtype=np.uint8).reshape(-1, 32)
        out = np.empty(h.shape[0], np.uint64)
        check(_native.lib().zg_dedup_lookup(self._h, h.ctypes.data, h.shape[0], out.ctypes.data))
        return out

    def __len__(self) -> int:
        n = C.c_uint64(0)
        check(_native.lib().zg_dedup_count(self._h, C.byref(n)))
        return int(n.value)

    def close(self):
        if self._h:
            _native.lib().zg_dedup_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:  # noqa: BLE001
            pass


def shard_of(hashes, n_shards: int) -> np.ndarray:
    """Which of n_shards GPUs owns each hash when the index is split across GPUs: the little-endian u32 at bytes
    16..19 of the hash, modulo n_shards (bytes 0..15 pick slot and tag inside a shard, so the split leaves them
    uniform)."""
    h = np.ascontiguousarray(hashes, dtype=np.uint8).reshape(-1, 32)
    return (h[:, 16:20].copy().view("<u4").reshape(-1) % np.uint32(n_shards)).astype(np.int32)


class _DeviceShard:
    """This rank's part of a split index, driven with torch CUDA tensors (no host round trip)."""

    def __init__(self, max_entries: int, ctx: Context | None = None):
        import torch
        self.torch = torch
        self.idx = DedupIndex(max_entries, ctx)
        self.device = torch.device("cuda", self.idx.ctx.device)

    def _stream(self):
        return C.c_void_p(self.torch.cuda.current_stream(self.device).cuda_stream)

    def insert_t(self, hashes_t, locs_t):
        n = hashes_t.shape[0]
        new = self.torch.zeros(n, dtype=self.torch.uint8, device=self.device)
        if n:
            check(_native.lib().zg_dev_dedup_insert(self.idx.handle, hashes_t.data_ptr(), locs_t.data_ptr(), n, new.data_ptr(), self._stream()))
        return new

    def lookup_t(self, hashes_t):
        n = hashes_t.shape[0]
        out = self.torch.empty(n, dtype=self.torch.int64, device=self.device)
        if n:
            check(_native.lib().zg_dev_dedup_lookup(self.idx.handle, hashes_t.data_ptr(), n, out.data_ptr(), self._stream()))
        return out

    def __len__(self):
        return len(self.idx)


class ShardedDedupIndex:
    """The dedup index split over the ranks of a torch.distributed group, one shard per GPU (owner = shard_of).

    This is the one place on the path with a real exchange step: every rank chunks and hashes its own shard files,
    but a chunk may already be held by any rank, so each batch of digests is routed to its owners and the answers
    are routed back -- two all-to-alls of 32-byte digests / 8-byte locations (NCCL over NVLink on a B200 node; gloo
    in the CPU tests, where `local` is a dict-backed stand-in for the device shard).  Locations are int64 tensors;
    -1 (all ones) means absent."""

    def __init__(self, max_entries_per_rank: int, ctx: Context | None = None, local=None, group=None):
        import torch
        import torch.distributed as dist
        self.torch, self.dist, self.group = torch, dist, group
        self.world = dist.get_world_size(group)
        self.rank = dist.get_rank(group)
        self.local = local if local is not None else _DeviceShard(max_entries_per_rank, ctx)

    def _route(self, hashes_t):
        """Sort the batch by owner and exchange counts: -> (order, send_counts, recv_counts)."""
        t = self.torch
        b = hashes_t[:, 16:20].to(t.int64)
        owner = (b[:, 0] | (b[:, 1] << 8) | (b[:, 2] << 16) | (b[:, 3] << 24)) % self.world
        order = t.sort(owner, stable=True).indices
        send = t.bincount(owner, minlength=self.world).to(t.int64)
        recv = t.empty_like(send)
        self.dist.all_to_all_single(recv, send, group=self.group)
        return order, [int(x) for x in send.tolist()], [int(x) for x in recv.tolist()]

    def _exchange(self, x, out_counts, in_counts):
        out = self.torch.empty((sum(out_counts),) + tuple(x.shape[1:]), dtype=x.dtype, device=x.device)
        self.dist.all_to_all_single(out, x.contiguous(), output_split_sizes=out_counts, input_split_sizes=in_counts, group=self.group)
        return out

    def insert(self, hashes_t, locs_t):
        """-> was_new (uint8 tensor, input order)."""
        order, send, recv = self._route(hashes_t)
        h = self._exchange(hashes_t[order], recv, send)
        l = self._exchange(locs_t[order], recv, send)
        new = self._exchange(self.local.insert_t(h, l), send, recv)
        out = self.torch.empty_like(new)
        out[order] = new
        return out

    def lookup(self, hashes_t):
        """-> locations (int64 tensor, input order; -1 = absent)."""
        order, send, recv = self._route(hashes_t)
        h = self._exchange(hashes_t[order], recv, send)
        loc = self._exchange(self.local.lookup_t(h), send, recv)
        out = self.torch.empty_like(loc)
        out[order] = loc
        return out
