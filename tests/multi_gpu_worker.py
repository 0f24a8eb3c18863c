# SPDX-License-Identifier: WTFPL
"""One rank of the multi-GPU parity check (launched by tests/test_multi_gpu.py and by hand:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29511 tests/multi_gpu_worker.py

Checks that the spp-sharded render + the NCCL reduce inside cudaRender gives the image of ONE GPU rendering all samples:
the invariant is the reference's one imageStore per pixel of the mean over all samples (shader.comp:53-57).  Default
accumulator (32.32 fixed point, int64 reduce): bit-identical for any rank count and any split.  Strict float sums
(sampleChunks = 1, float reduce): equal up to the rounding of the partial sums (SURVEY.md 8e "determinism note"), bounded
here by 2^-20 absolute on radiance in [0, 1].  Prints one line per case and MULTI_GPU_CHECK PASS / FAIL."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import torch.distributed as dist
from vulkancomputeraytracing_b200 import cabi, host, sharding

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
lib = cabi.load()
cabi.check(lib.cudaSelectDevice(local))
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
sharding.init_communicator(dist, device=torch.device("cuda", local))
W, H = 320, 180
ok = True
cases = 0
for scene_name, scene in (("rtiow", host.generate_scene()), ("stress", host.stress_scene())):
    host.upload_scene(scene)
    for trav in (cabi.VCRT_TRAVERSAL_BRUTE, cabi.VCRT_TRAVERSAL_LBVH):
        # 37: unequal shards; 1: fewer samples than ranks (the reference's default spp; the other ranks only join the reduce)
        for spp in (8 * world, 37, 1):
            for flags in (0, cabi.VCRT_FLAG_ALL_REDUCE):
                for chunks in (0, 1):
                    off, cnt = sharding.shard_samples(spp, world, rank)
                    have = rank == 0 or flags != 0
                    img, st = host.render(W, H, cnt, sample_offset=off, traversal=trav, total_samples=spp, flags=flags, sample_chunks=chunks, download=have)
                    full, sf = host.render(W, H, spp, traversal=trav, flags=cabi.VCRT_FLAG_NO_REDUCE, sample_chunks=chunks)   # every rank alone, communicator bypassed
                    rays = torch.tensor([st.rays], dtype=torch.int64, device=f"cuda:{local}"); dist.all_reduce(rays)
                    good = int(rays.item()) == sf.rays
                    worst = 0.0
                    if have:
                        worst = float(np.abs(img - full).max())
                        good &= np.array_equal(img, full) if chunks == 0 else worst <= 2.0 ** -20
                    cases += 1
                    if have:
                        print(f"rank {rank}/{world} {scene_name} traversal {trav} spp {spp} (mine {cnt}) flags {flags} chunks {chunks}: "
                              f"sharded {'==' if chunks == 0 else '~='} single: {good}  max|d| {worst:.3e}", flush=True)
                    ok &= bool(good)
    # progressive frames across ranks: two sharded calls accumulated == one single-GPU render of all samples
    off, cnt = sharding.shard_samples(12, world, rank)
    host.render(W, H, cnt, sample_offset=off, traversal=cabi.VCRT_TRAVERSAL_LBVH, total_samples=24, download=False)
    off2, cnt2 = sharding.shard_samples(12, world, rank)
    img, _ = host.render(W, H, cnt2, sample_offset=12 + off2, traversal=cabi.VCRT_TRAVERSAL_LBVH, total_samples=24, flags=cabi.VCRT_FLAG_ACCUMULATE, download=rank == 0)
    full, _ = host.render(W, H, 24, traversal=cabi.VCRT_TRAVERSAL_LBVH, flags=cabi.VCRT_FLAG_NO_REDUCE)
    if rank == 0:
        good = np.array_equal(img, full)
        print(f"rank 0/{world} {scene_name} progressive 12 + 12 samples sharded == single 24: {good}", flush=True)
        ok &= bool(good)
t = torch.tensor([1 if ok else 0], device=f"cuda:{local}")
dist.all_reduce(t, op=dist.ReduceOp.MIN)
if rank == 0:
    print(f"MULTI_GPU_CHECK {'PASS' if int(t.item()) == 1 else 'FAIL'} ranks {world} cases {cases}", flush=True)
host.destroy()
dist.destroy_process_group()
sys.exit(0 if int(t.item()) == 1 else 1)
