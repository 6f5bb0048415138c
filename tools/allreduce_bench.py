"""Dev tool (torchrun): time the per-step gradient exchange collectives alone -- NCCL all-reduce of the flat
236 B/Gaussian buffer (plain exchange), of the 44 B/Gaussian small-leaf buffer and the all-gather of 12 B/Gaussian
SH factors (factored exchange).  Run under different NCCL_ALGO / NCCL_PROTO to see what NCCL's tuner leaves behind."""
import os
import sys

import torch
import torch.distributed as dist

rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
P = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
big, small = torch.randn(P * 59, device="cuda"), torch.randn(P * 11, device="cuda")
fac, fac_all = torch.randn(P * 3, device="cuda"), torch.empty(world * P * 3, device="cuda")
tag = "ALGO=%s PROTO=%s" % (os.environ.get("NCCL_ALGO", "-"), os.environ.get("NCCL_PROTO", "-"))
for name, nbytes, fn in (("all_reduce 236 B/G", big.numel() * 4, lambda: dist.all_reduce(big)),
                         ("all_reduce  44 B/G", small.numel() * 4, lambda: dist.all_reduce(small)),
                         ("all_gather  12 B/G", fac_all.numel() * 4, lambda: dist.all_gather_into_tensor(fac_all, fac))):
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
        print(f"[{tag}] {name}: {float(t) * 1e3:.0f} us  (world {world}; {nbytes / 1e6:.0f} MB, "
              f"{nbytes / float(t) / 1e6:.0f} GB/s)", flush=True)
dist.barrier()
dist.destroy_process_group()
