# dev: how long does the path's one exchange take? run under torchrun. Prints per-rank kernel ms vs total ms of cudaRender.
import os, sys
sys.path.insert(0, '.')
import torch, torch.distributed as dist
from vulkancomputeraytracing_b200 import cabi, host, sharding
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
lib = cabi.load(); cabi.check(lib.cudaSelectDevice(local)); torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
sharding.init_communicator(dist, device=torch.device("cuda", local))
host.upload_scene(host.generate_scene())
W, H, SPP = 3840, 2160, 8 * world
for flags, name in ((0, "reduce"), (cabi.VCRT_FLAG_ALL_REDUCE, "allreduce"), (cabi.VCRT_FLAG_NO_REDUCE, "none")):
    for i in range(3):
        off, cnt = sharding.shard_samples(SPP, world, rank)
        lib.cudaCommBarrier()
        _, st = host.render(W, H, cnt, sample_offset=off, total_samples=SPP, flags=flags, download=False, traversal=2)
        if i == 2: print(f"rank {rank} {name}: kernel {st.renderKernelMs:.2f} ms, whole cudaRender {st.totalMs:.2f} ms, after-kernel {st.totalMs - st.renderKernelMs:.2f} ms", flush=True)
host.destroy(); dist.destroy_process_group()
