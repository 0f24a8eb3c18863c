# dev/verification: run under torchrun on N GPUs. Checks that the spp-sharded render + NCCL reduce equals a single-GPU render.
import os, sys, ctypes
import numpy as np
sys.path.insert(0, '.')
import torch, torch.distributed as dist
from vulkancomputeraytracing_b200 import cabi, host, sharding
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
lib = cabi.load(); cabi.check(lib.cudaSelectDevice(local))
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
sharding.init_communicator(dist, device=torch.device("cuda", local))
scene = host.generate_scene(); host.upload_scene(scene)
W, H, SPP = 320, 180, 37
ok = True
for trav in (cabi.VCRT_TRAVERSAL_BRUTE, cabi.VCRT_TRAVERSAL_LBVH):
    for flags in (0, cabi.VCRT_FLAG_ALL_REDUCE):
        off, cnt = sharding.shard_samples(SPP, world, rank)
        img, st = host.render(W, H, cnt, sample_offset=off, traversal=trav, total_samples=SPP, flags=flags, download=(rank == 0 or flags != 0))
        # single-GPU reference on every rank, communicator bypassed
        full, _ = host.render(W, H, SPP, traversal=trav, flags=cabi.VCRT_FLAG_NO_REDUCE)
        if img is not None:
            same = np.array_equal(img, full)   # fixed-point sums: integer reduce, so bit-identical for any rank count
            print(f"rank {rank} trav {trav} flags {flags}: sharded == single: {same}  max|d| {np.abs(img-full).max():.3e}", flush=True)
            ok &= same
t = torch.tensor([1 if ok else 0], device=f"cuda:{local}"); dist.all_reduce(t, op=dist.ReduceOp.MIN)
if rank == 0: print("MULTI_GPU_CHECK", "PASS" if int(t.item()) == 1 else "FAIL", flush=True)
host.destroy(); dist.destroy_process_group()
