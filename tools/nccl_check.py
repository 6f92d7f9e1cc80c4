"""Dev tool (run under torchrun): times the 24 MB packed-gradient all-reduce and the 4-float bounds all-reduce."""
import os
import torch
import torch.distributed as dist

local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
rank, world = dist.get_rank(), dist.get_world_size()
for n, name in ((6_000_000, "packed gradients 24 MB"), (4, "bounds 16 B")):
    t = torch.ones(n, device="cuda")
    for _ in range(5):
        dist.all_reduce(t)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(20):
        dist.all_reduce(t)
    b.record()
    torch.cuda.synchronize()
    ms = a.elapsed_time(b) / 20
    if rank == 0:
        print(f"all_reduce {name}: {ms:.3f} ms  busbw {2 * (world - 1) / world * n * 4 / ms / 1e6:.1f} GB/s  world {world}", flush=True)
dist.destroy_process_group()
