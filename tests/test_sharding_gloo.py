# SPDX-License-Identifier: WTFPL
"""Host side of the multi-GPU path on CPU: the spp partition, and (world_size 2, gloo) the rendezvous
that ships the NCCL id plus the sum-then-normalise algebra of the reduce, with the oracle standing in
for each rank's render."""
import os
import socket
import sys

import numpy as np
import pytest

from vulkancomputeraytracing_b200.sharding import shard_samples

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_shards_cover_the_sample_range_exactly():
    for total in (0, 1, 7, 16, 500, 1000, 1001):
        for world in (1, 2, 3, 4, 8):
            spans = [shard_samples(total, world, r) for r in range(world)]
            assert spans[0][0] == 0 and sum(c for _, c in spans) == total
            for (o0, c0), (o1, _) in zip(spans, spans[1:]):
                assert o0 + c0 == o1
            counts = [c for _, c in spans]
            assert max(counts) - min(counts) <= 1
    assert shard_samples(1000, 8, 3) == (375, 125) and shard_samples(1000, 2, 1) == (500, 500)   # BASELINE.json configs[2]
    with pytest.raises(ValueError):
        shard_samples(10, 2, 2)


def _worker(rank, world, port, out):
    sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
    import torch
    import torch.distributed as dist
    import oracle_harness as oh
    from vulkancomputeraytracing_b200 import sharding
    dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{port}", rank=rank, world_size=world)
    try:
        # the id exchange: without a GPU rank 0 cannot create a real id (that is a loud error, tested elsewhere),
        # so ship a stand-in through the same broadcast code path
        import ctypes
        from vulkancomputeraytracing_b200 import cabi
        real = cabi.load().cudaCommGetUniqueId
        class FakeLib:
            def __getattr__(self, name): return getattr(cabi._lib, name)
            def cudaCommGetUniqueId(self, buf):
                for i in range(128): buf[i] = (i * 7 + 3) & 255
                return 0
        cabi_load = cabi.load
        cabi.load = lambda: FakeLib()
        try:
            got = sharding.exchange_unique_id(dist)
        finally:
            cabi.load = cabi_load
        assert got == bytes((i * 7 + 3) & 255 for i in range(128))
        # each rank renders its shard of the samples (oracle = stand-in for cudaRender with VCRT_FLAG_NO_REDUCE)
        orc = oh.Oracle()
        scene = orc.scene()
        total_spp, w, h = 5, 40, 24
        offset, count = sharding.shard_samples(total_spp, world, rank)
        sums, st = orc.render(scene, w, h, count, sample_offset=offset, normalise=False, threads=2)
        t = torch.from_numpy(sums[..., :3].copy())
        dist.reduce(t, dst=0, op=dist.ReduceOp.SUM)                  # the path's one exchange
        rays = torch.tensor([st["rays"]], dtype=torch.int64)
        dist.reduce(rays, dst=0, op=dist.ReduceOp.SUM)
        if rank == 0:
            full, fst = orc.render(scene, w, h, total_spp, normalise=True, threads=2)
            np.testing.assert_allclose(t.numpy() / np.float32(total_spp), full[..., :3], rtol=1e-6, atol=1e-7)
            assert int(rays[0]) == fst["rays"]                       # the set of paths traced does not depend on the rank count
            open(out, "w").write("ok")
    finally:
        dist.destroy_process_group()


def test_two_rank_gloo_reduce_equals_single_render(tmp_path):
    import torch.multiprocessing as mp
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]
    out = str(tmp_path / "ok")
    mp.spawn(_worker, args=(2, port, out), nprocs=2, join=True)
    assert open(out).read() == "ok"
