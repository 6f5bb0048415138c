"""Dev tool (torchrun): time the per-step gradient exchange alone -- NCCL all-reduce of the flat 236 B/Gaussian buffer."""
import os
import sys

import torch
import torch.distributed as dist

rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
P = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
n = P * 59
x = torch.randn(n, device="cuda")
src = torch.randn(n, device="cuda")
for name, fn in (("all_reduce %d MB" % (n * 4 // 10**6), lambda: dist.all_reduce(x)),
                 ("copy into flat", lambda: x.copy_(src))):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    dist.barrier()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(20):
        fn()
    b.record()
    torch.cuda.synchronize()
    ms = a.elapsed_time(b) / 20
    t = torch.tensor([ms], device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    if rank == 0:
        print(f"{name}: {float(t):.3f} ms  (world {world}; algbw {n * 4 / float(t) / 1e6:.0f} GB/s)")
dist.barrier()
dist.destroy_process_group()
