# SPDX-License-Identifier: WTFPL
"""Samples-per-pixel sharding across one-process-per-GPU ranks.

Every pixel-sample is independent and its random numbers are keyed on (pixel, GLOBAL sample index),
so rank r of R renders sample indices [offset_r, offset_r + count_r) of every pixel into its own sum
buffer; the one exchange of the path is the NCCL reduce of those sums inside cudaRender.
torch.distributed is used for rendezvous only (shipping the 128-byte NCCL id)."""
import ctypes

from . import cabi


def shard_samples(total_spp, world_size, rank):
    """(first global sample index, sample count) of `rank`; counts differ by at most one and cover [0, total_spp)."""
    if world_size < 1 or not 0 <= rank < world_size or total_spp < 0:
        raise ValueError("bad shard arguments")
    base, extra = divmod(total_spp, world_size)
    count = base + (1 if rank < extra else 0)
    offset = rank * base + min(rank, extra)
    return offset, count


def exchange_unique_id(dist, device=None):
    """Rank 0 creates the NCCL unique id through the C ABI; torch.distributed broadcasts its 128 bytes."""
    import torch
    lib = cabi.load()
    buf = (ctypes.c_ubyte * 128)()
    if dist.get_rank() == 0:
        cabi.check(lib.cudaCommGetUniqueId(buf))
    t = torch.tensor(list(buf), dtype=torch.uint8, device=device or "cpu")
    dist.broadcast(t, src=0)
    return bytes(t.cpu().tolist())


def init_communicator(dist, device=None):
    """Collective: creates the layer's own NCCL communicator over the ranks of the default process group."""
    id128 = exchange_unique_id(dist, device)
    buf = (ctypes.c_ubyte * 128).from_buffer_copy(id128)
    cabi.check(cabi.load().cudaCommInitRank(buf, dist.get_rank(), dist.get_world_size()))
