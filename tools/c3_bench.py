"""BASELINE config 3: scene S with 3 M Gaussians, a batch of 8 camera views on a 30-degree arc
(yaw -15 ... +15 degrees, looking at (0, 0, 6)) at 1920x1080, forward + backward of every view, per-Gaussian gradients
summed over the batch.  STRONG scaling: the 8 views are split over G ranks (gsr_dp.shard_views), each rank renders its
block one view after the other, then the exchange of gsr_dp.dp_step_factored (all-gather of 12 B/Gaussian/view SH
factors + all-reduce of the 44 B/Gaussian small gradients, SH gradient rebuilt locally).  G = 1 has no exchange but
sums the batch the same way (factors kept, one rebuild); --full-sum accumulates full gradients instead.

  python tools/c3_bench.py                                                     # 1 GPU
  python -m torch.distributed.run --nnodes=1 --nproc-per-node G --master-addr 127.0.0.1 tools/c3_bench.py

Timing: CUDA events around `--steps` steps, barrier + synchronize on both sides, max over ranks; rank 0 prints one JSON
line (views/s = 8 * steps / time; Gaussians/s = P * views/s).
"""
import argparse
import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "gaussian-transformer_b200"))
import gsr_dp  # noqa: E402
import gsr_scenes as S  # noqa: E402
from diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer  # noqa: E402

LEAVES = gsr_dp.LEAVES


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--P", type=int, default=3_000_000)
    ap.add_argument("--views", type=int, default=8)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--full-sum", action="store_true", help="1 GPU: sum full gradients instead of SH factors")
    args = ap.parse_args()
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    dev = torch.device("cuda", int(os.environ.get("LOCAL_RANK", 0)))
    torch.cuda.set_device(dev)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    W, H, M, D = 1920, 1080, 16, 3
    sc = S.scene_S(args.P, W, H, seed=0)
    t = {k: torch.from_numpy(np.ascontiguousarray(sc[k])).to(dev) for k in LEAVES + ("bg",)}
    dpix = torch.from_numpy(S.upstream_grad(W, H, 1)).to(dev)
    rasters = []
    for yaw in np.linspace(-15.0, 15.0, args.views):
        cam = S.look_at_camera(W, H, 60.0, float(yaw))
        c = {k: torch.from_numpy(cam[k]).to(dev) for k in ("viewmatrix", "projmatrix", "campos")}
        rasters.append(GaussianRasterizer(GaussianRasterizationSettings(
            H, W, cam["tanfovx"], cam["tanfovy"], t["bg"], 1.0, c["viewmatrix"], c["projmatrix"], D, c["campos"],
            False, False)))
    mine = gsr_dp.shard_views(args.views, rank, world)
    rendered = []

    def render_backward(v):
        leaves = {k: t[k].detach().requires_grad_(True) for k in LEAVES}
        m2 = torch.zeros_like(leaves["means3D"], requires_grad=True)
        color, radii = rasters[v](means3D=leaves["means3D"], means2D=m2, opacities=leaves["opacities"],
                                  shs=leaves["sh"], scales=leaves["scales"], rotations=leaves["rotations"])
        color.backward(dpix)
        rendered.append(radii)
        return {k: leaves[k].grad for k in LEAVES}

    if args.full_sum:
        # plain accumulation of the full 236 B/Gaussian gradients (what dp_step does), single GPU only
        flat = gsr_dp.FlatGrads({k: tuple(t[k].shape) for k in LEAVES}, dev)

        def step():
            flat.zero_()
            for v in mine:
                with gsr_dp.grads_into(flat):
                    g = render_backward(v)
                flat.add_(g)
            return flat.views
    else:
        # factored sum at every G (at G = 1 the "exchange" is local: the factors of the 8 views are kept and dL/dsh is
        # rebuilt once per step instead of 7 read-modify-write passes over 192 B/Gaussian)
        flat = gsr_dp.FlatGrads({k: tuple(t[k].shape) for k in gsr_dp.SMALL_LEAVES}, dev)
        per_rank = (args.views + world - 1) // world
        exchange = gsr_dp.FactorExchange(per_rank, args.P, dev, world)

        def rb(v, _ex):
            with gsr_dp.capture_sh_factors(exchange), gsr_dp.grads_into(flat):
                return render_backward(v)

        def step():
            return gsr_dp.dp_step_factored(rb, args.views, flat, rank, world, t["means3D"], D, M, exchange=exchange)

    def sync_all():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        g = step()
    sync_all()
    rendered.clear()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(args.steps):
        g = step()
    b.record()
    sync_all()
    ms = torch.tensor([a.elapsed_time(b)], device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    checksum = float(sum(g[k].double().abs().sum() for k in LEAVES))
    if rank == 0:
        sec = float(ms) / 1e3
        vps = args.views * args.steps / sec
        print(json.dumps({
            "workload": "C3: scene S, %d Gaussians SH3, %d views on a 30-degree arc, 1920x1080, fwd+bwd, gradients "
                        "summed over the batch" % (args.P, args.views),
            "n_gpus": world, "scaling": "strong", "views_per_rank": len(mine), "steps": args.steps,
            "ms_per_step": round(1e3 * sec / args.steps, 3), "views_per_s": round(vps, 1),
            "gaussians_per_s": round(args.P * vps), "grad_abs_sum": checksum,
            "visible_last_view": int((rendered[-1] > 0).sum()),
            "sum": "full gradients" if args.full_sum else "SH factors + rebuild",
            "exchange": "none" if world == 1 else "all-gather 12 B/Gaussian/view + all-reduce 44 B/Gaussian"}))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
