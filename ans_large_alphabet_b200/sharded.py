"""Multi-GPU semi-static coding: one process per GPU, the input split into per-rank block ranges.

SURVEY.md section 8(e): blocks are independent given the model, so every rank encodes / decodes its
own contiguous shard with no data-path collective.  The single semi-static model comes from ONE
all-reduce (sum) of the symbol histogram -- plus an all-reduce (max) of the largest value for
the methods whose alphabet is the raw value range -- after which every rank runs the same
deterministic host normalisation on the same global histogram and therefore builds the same
model without a broadcast.

torch / torch.distributed are plumbing here (device buffers, the NCCL communicator); all coding
work goes through the C-ABI of include/b200ans.h.
"""
import os

import numpy as np

from ._capi import METHOD_FOLD, METHOD_INT, METHOD_MSB, METHOD_RFOLD


OPT_BINS = 3071  # plain ANS: bins counted optimistically in one pass (four shared-memory replicas; b200ans.cu kOptBins)


def shard_range(n, block_syms, rank, world):
    """Contiguous block range of `rank`: blocks [rank*nb/world, (rank+1)*nb/world)  ->  symbol range."""
    nb = (n + block_syms - 1) // block_syms
    b0 = rank * nb // world
    b1 = (rank + 1) * nb // world
    return min(n, b0 * block_syms), min(n, b1 * block_syms)


def alphabet_bins(method, f, max_value=0):
    if method == METHOD_INT:
        return int(max_value) + 1
    if method == METHOD_MSB:
        return 1280
    return 1 << (f + 9)


def _all_reduce(t, op, group):
    import torch.distributed as dist

    if group is not None and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(t, op=op, group=group)


def reduce_histogram(bins, group=None):
    """ONE all-reduce (sum) of a rank-local histogram (torch int64 tensor, any device the group's
    backend supports).  Returns the global histogram as host uint64.  Counts are < 2^63, so the
    int64 transport is exact."""
    import torch.distributed as dist

    _all_reduce(bins, dist.ReduceOp.SUM, group)
    if bins.is_cuda:  # one asynchronous copy into page-locked memory + one stream synchronisation
        import torch

        host = _pinned_like(bins)
        host.copy_(bins, non_blocking=True)
        torch.cuda.current_stream(bins.device).synchronize()
        return host.numpy().view(np.uint64).copy()
    return bins.cpu().numpy().view(np.uint64)


_PINNED = {}


def _pinned_like(t):
    """page-locked staging tensor of t's shape (cached: a step's histogram has the same shape every time)"""
    import torch

    key = (tuple(t.shape), t.dtype)
    buf = _PINNED.get(key)
    if buf is None:
        buf = _PINNED[key] = torch.empty(t.shape, dtype=t.dtype, pin_memory=True)
    return buf


def reduce_bins_and_max(buf, opt_bins, group=None):
    """ONE all-reduce for counts AND the largest value: `buf` holds this rank's optimistic counts in [0, opt_bins) and
    its own maximum in slot opt_bins + rank (zeros in the other ranks' slots), so the sum is an all-gather of the
    maxima behind the global counts.  Returns (global counts as host uint64, global maximum)."""
    host = reduce_histogram(buf, group)
    return host[:opt_bins], int(host[opt_bins:].max())


def reduce_max(value, device, group=None):
    """all-reduce (max) of the largest raw value (the alphabet size of ANS / ANSrfold pass 1)"""
    import torch
    import torch.distributed as dist

    t = torch.tensor([int(value)], dtype=torch.int64, device=device)
    _all_reduce(t, dist.ReduceOp.MAX, group)
    return int(t.item())


def global_histogram(codec, method, f, d_src, n, group=None, sync=None):
    """Device histogram of this rank's shard, all-reduced over `group`. Returns host uint64 bins.

    d_src: torch uint32/int32 CUDA tensor (16-byte aligned). `sync` is called between library
    calls and torch ops (the codec may run on its own stream)."""
    import torch
    import torch.distributed as dist

    dev = d_src.device
    sync = sync or (lambda: torch.cuda.synchronize(dev))
    world = dist.get_world_size(group) if (group is not None and dist.is_initialized()) else 1

    def hist(m, nbins):
        bins = torch.zeros(nbins, dtype=torch.int64, device=dev)
        sync()
        codec.histogram_device(m, f, d_src.data_ptr(), n, bins.data_ptr(), nbins)
        sync()
        return reduce_histogram(bins, group if world > 1 else None)

    if method == METHOD_INT:
        # The reference takes the maximum first, then counts (ans_int.hpp:42-49).  One pass and ONE all-reduce instead:
        # count optimistically into OPT_BINS bins while the kernel keeps the running maximum, and ship every rank's
        # maximum in its own slot behind the bins (sum of one-hot slots = all-gather).  Only an alphabet beyond
        # OPT_BINS needs the second pass.
        rank = dist.get_rank(group) if world > 1 else 0
        buf = getattr(codec, "_shard_bins", None)  # reused from step to step: one memset instead of an allocation
        if buf is None or buf.numel() != OPT_BINS + world or buf.device != dev:
            buf = codec._shard_bins = torch.empty(OPT_BINS + world, dtype=torch.int64, device=dev)
        buf.zero_()
        sync()
        # the kernel's 32-bit maximum lands in the low half of this rank's (zeroed, little-endian) int64 slot
        codec.histogram_max_device(METHOD_INT, f, d_src.data_ptr(), n, buf.data_ptr(), OPT_BINS,
                                   buf.data_ptr() + 8 * (OPT_BINS + rank))
        sync()
        counts, max_value = reduce_bins_and_max(buf, OPT_BINS, group if world > 1 else None)
        if max_value < OPT_BINS:
            return counts[: max_value + 1].copy()
        return hist(METHOD_INT, max_value + 1)
    if method == METHOD_RFOLD:
        mx = torch.zeros(1, dtype=torch.int32, device=dev)
        sync()
        codec.max_device(d_src.data_ptr(), n, mx.data_ptr())
        sync()
        max_value = reduce_max(int(mx.item()) & 0xFFFFFFFF, dev, group if world > 1 else None)
        raw = hist(METHOD_INT, max_value + 1)
        codec.rfold_prepare(f, raw)  # identical on every rank: same global histogram, same tie-break
        return hist(METHOD_RFOLD, alphabet_bins(method, f))
    return hist(method, alphabet_bins(method, f))


def init_comm(codec, group=None):
    """Give `codec` its NCCL communicator over the ranks of `group`: rank 0 draws the unique id (b200ans_comm_unique_id),
    torch.distributed only carries the 128 bytes to the others, b200ans_comm_init is the collective ncclCommInitRank.
    After this b200ans_compress_sharded[_device] runs the histogram exchange inside the library, on the codec's
    stream.  Works with any torch backend (the bytes travel as a CPU tensor under gloo, a CUDA tensor under nccl)."""
    import torch
    import torch.distributed as dist

    from ._capi import comm_unique_id

    rank, world = dist.get_rank(group), dist.get_world_size(group)
    uid = comm_unique_id() if rank == 0 else np.zeros(128, dtype=np.uint8)
    t = torch.from_numpy(uid)
    on_gpu = dist.get_backend(group) == "nccl"
    if on_gpu:
        t = t.cuda()
    dist.broadcast(t, src=dist.get_global_rank(group, 0) if group is not None else 0, group=group)
    codec.comm_init(t.cpu().numpy(), rank, world)
    if world > 1 and os.environ.get("B200ANS_NO_IPC", "0") in ("", "0"):
        init_ipc(codec, group)
    return rank, world


def init_ipc(codec, group=None):
    """The histogram exchange over NVLink peer memory (b200ans_comm_ipc_export / _attach): every rank exports its
    exchange buffer, torch.distributed carries the 64-byte handles, every rank maps the others'.  For ranks that are
    processes on one node; returns False (and leaves the NCCL path in place) where the buffers cannot be mapped."""
    import torch
    import torch.distributed as dist

    rank, world = dist.get_rank(group), dist.get_world_size(group)
    ok = 1
    try:
        handle = codec.comm_ipc_export()
    except Exception:  # noqa: BLE001 -- no IPC here: this rank still takes part in the collectives below, then everybody stays with NCCL
        handle, ok = np.zeros(64, dtype=np.uint8), 0
    mine = torch.from_numpy(handle)
    if dist.get_backend(group) == "nccl":
        mine = mine.cuda()
    every = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(every, mine, group=group)  # (also the barrier between everybody's export and anybody's attach)
    if ok:
        try:
            codec.comm_ipc_attach(rank, world, torch.stack(every).cpu().numpy())
        except Exception:  # noqa: BLE001 -- not one node, or no peer access
            ok = 0
    flag = torch.tensor([ok], dtype=torch.int32, device=mine.device)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=group)
    if int(flag.item()) == 0:
        codec.comm_ipc_release()
        return False
    return True


def compress_shard(codec, method, f, d_src, n, d_dst, group=None, sync=None):
    """Semi-static encode of this rank's shard with the GLOBAL model (b200ans_compress_sharded_device: histogram ->
    one NCCL all-reduce inside the library -> normalise -> encode). Returns container bytes.  The codec needs its
    communicator (init_comm) when the group has more than one rank."""
    return codec.compress_sharded_device(method, f, d_dst.data_ptr(), d_dst.numel() * d_dst.element_size(), d_src.data_ptr(), n)


def compress_shard_staged(codec, method, f, d_src, n, d_dst, group=None, sync=None):
    """The same through the staged entry points with torch.distributed doing the all-reduce (round 1's path; kept as the
    cross-check of the library's own exchange)."""
    bins = global_histogram(codec, method, f, d_src, n, group, sync)
    codec.model_build(method, f, bins)
    return codec.encode_device(d_dst.data_ptr(), d_dst.numel() * d_dst.element_size(), d_src.data_ptr(), n)


def decompress_shard(codec, method, f, d_blob, nbytes, d_out, n):
    """Decode needs no communication: each shard's container carries the (global) prelude."""
    codec.decompress_device(method, f, d_out.data_ptr(), n, d_blob.data_ptr(), nbytes)
