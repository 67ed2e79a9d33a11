"""All-reduce probe for the band-sum exchange of the sharded dual update:
   torchrun --nproc-per-node N profiles/nccl_probe.py
prints time and algorithm bandwidth for the slab (1/5 of C1) and the full C1 message."""
import os
import torch
import torch.distributed as dist

dist.init_process_group("nccl")
rank = dist.get_rank()
torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", 0)))
for mb in (17, 68, 339):
    x = torch.ones(mb * 1000 * 1000 // 4, dtype=torch.float32, device="cuda")
    for _ in range(3):
        dist.all_reduce(x)
    torch.cuda.synchronize()
    dist.barrier()
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    e0.record()
    reps = 10
    for _ in range(reps):
        dist.all_reduce(x)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    if rank == 0:
        print(f"all_reduce {mb} MB x{dist.get_world_size()}: {ms:.3f} ms  algbw {mb / ms:.1f} GB/s", flush=True)
dist.destroy_process_group()
