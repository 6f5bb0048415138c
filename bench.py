#!/usr/bin/env python
"""bench.py -- headline benchmark of the hot path (rasterize_gaussians forward+backward).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
  (N > 1: launched by `python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N ...`)

Workload (BASELINE.json configs[1], "C2"): synthetic scene S, 1,000,000 Gaussians, SH degree 3,
1920x1080, one view; a step = one forward+backward of rasterize_gaussians through the public API
(GaussianRasterizer + autograd).  With N > 1 every rank renders its own view of the replicated Gaussians (cameras on
a 30-degree arc) and the per-Gaussian gradients are summed with an NCCL all-reduce: weak scaling, value = views/s over
all ranks.  Prints ONE JSON line (rank 0).
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path[:0] = [os.path.join(ROOT, "gaussian-transformer_b200"), os.path.join(ROOT, "tests")]

P_GAUSS, WIDTH, HEIGHT, SH_DEG, M_COEF = 1_000_000, 1920, 1080, 3, 16
METRIC = "fwd+bwd iters/s (rasterize_gaussians, 1M Gaussians SH3, 1920x1080)"
WORKLOAD = "C2: scene S (seed 0), 1M Gaussians, SH degree 3, 1920x1080, 1 view per GPU, fwd+bwd"
UNIT = "iters/s"
TENSORS = ("means3D", "scales", "rotations", "opacities", "sh", "bg", "viewmatrix", "projmatrix", "campos")
LEAVES = ("means3D", "sh", "opacities", "scales", "rotations")


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="ours", choices=("ours", "reference"))
    ap.add_argument("--no-cpu-baseline", action="store_true", help="skip the ~10-20 s CPU oracle timing")
    ap.add_argument("--no-c3", action="store_true", help="N > 1: skip the C3 (3M Gaussians, 8-view batch) record")
    ap.add_argument("--timeline", default=None, help="dev: write a kernel timeline of 2 steps (torch.profiler) and exit")
    ap.add_argument("--dp-exchange", default="peer", choices=("peer", "factors", "allreduce"),
                    help="N > 1: our kernels over NVLink peer memory (default: fused factor gather + SH rebuild, "
                         "two-shot all-reduce of 44 B/Gaussian), the same exchange with NCCL (all-gather of 12 "
                         "B/Gaussian/view SH factors + all-reduce of 44 B/Gaussian), or one NCCL all-reduce of the "
                         "full 236 B/Gaussian")
    ap.add_argument("--peer-gather", default="push", choices=("nccl", "copy", "fused", "push"),
                    help="peer exchange: how the SH-factor records travel (gsr_dp.PeerExchange)")
    ap.add_argument("--multimem", action="store_true",
                    help="peer exchange: all-reduce through the allocation's NVSwitch multicast mapping "
                         "(multimem.ld_reduce / multimem.st) instead of the two-shot P2P kernel (measured at 8 GPUs: "
                         "2.12 vs 2.10 ms/step, profiles/r02_multigpu_exchange_experiments.txt)")
    ap.add_argument("--c3-exchange", default="factors", choices=("factors", "peer"),
                    help="exchange of the C3 record (measured at 8 GPUs: 3.11 ms/batch with the NCCL factor "
                         "exchange, 3.20 with the peer exchange)")
    return ap.parse_args()


def peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"], "measured"
    except Exception:
        return 6650.0, "fallback"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 20 ms while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.f = tempfile.NamedTemporaryFile("w+", suffix=".csv", delete=False)
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(index), "--query-gpu=" + self.Q,
                                       "--format=csv,noheader,nounits", "-lms", "20"], stdout=self.f,
                                      stderr=subprocess.DEVNULL)
        except Exception:
            self.p = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self.p is None:
            return out
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except Exception:
            self.p.kill()
        self.f.flush()
        self.f.seek(0)
        sm, mx, reasons = [], [], set()
        for line in self.f.read().splitlines():
            c = [x.strip() for x in line.split(",")]
            if len(c) < 7:
                continue
            try:
                sm.append(float(c[0]))
                mx.append(float(c[1]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), c[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        os.unlink(self.f.name)
        if sm:
            out.update(sm_mhz=float(np.median(sm)), sm_max_mhz=float(max(mx)), reasons=sorted(reasons),
                       samples=len(sm))
        return out


def algorithmic_bytes(P, V, R, N, K, M):
    """SURVEY.md 8(d) / DESIGN.md: compulsory HBM bytes of one fwd+bwd iteration, per stage and in total."""
    a_fwd = 16 * P + (68 + 12 * K) * V + 60 * R + 20 * N
    a_bwd = 20 * N + 40 * R + (116 + 12 * K) * V + (56 + 12 * M) * P
    per_stage = {
        "preprocess_fwd": 16 * P + (32 + 12 * K) * V + 36 * V,
        "render_fwd": 4 * R + 36 * R + 20 * N,
        "render_bwd": 20 * N + 4 * R + 36 * R + 36 * V,
        "preprocess_bwd": 36 * V + (44 + 12 * K) * V + (56 + 12 * M) * P,
    }
    return a_fwd + a_bwd, per_stage


# dram__bytes_read.sum + dram__bytes_write.sum per launch from the `ncu --set full` capture of the yaw-0 C2 view
# (V = 860,991, R = 12,071,224) summarised in profiles/r02_ncu_full_metrics.txt.  Quoted as roofline.traffic only when
# the run's own V and R equal the capture's (N = 1); otherwise traffic is null.
NCU_CAPTURE = {"V": 860991, "R": 12071224, "source": "profiles/r02_ncu_full_metrics.txt"}
NCU_DRAM_BYTES_C2 = {
    "preprocess_fwd": 236.03e6 + 37.63e6,
    "render_fwd": 65.47e6 + 17.93e6,
    "render_bwd": 116.66e6 + 6.51e6,
    "preprocess_bwd": 285.04e6 + 244.50e6,
}


def make_scene(rank, world):
    import gsr_scenes as S
    yaw = 0.0 if world == 1 else -15.0 + 30.0 * rank / (world - 1)
    sc = S.scene_S(P_GAUSS, WIDTH, HEIGHT, seed=0, M=M_COEF, sh_degree=SH_DEG, yaw_deg=yaw)
    return sc, S.upstream_grad(WIDTH, HEIGHT, seed=1)


def cpu_baseline(sc, dpix):
    """The oracle port (oracle/oracle.c, OpenMP over all host cores) timed on ONE full iteration of the same workload
    (~10-20 s of CPU work)."""
    from checkers import Oracle
    orc = Oracle()
    t0 = time.time()
    orc.forward(sc, internals=False)
    orc.backward(dpix)
    dt = time.time() - t0
    return {"value": 1.0 / dt, "unit": UNIT, "cores": orc.num_threads(), "kind": "port",
            "sample": "1 full fwd+bwd iteration of the same workload (oracle/oracle.c, OpenMP), %.1f s" % dt}


def cpu_baseline_pytorch(dev):
    """BASELINE.json's second baseline and configs[0]: the pure-PyTorch CPU implementation of rasterize_gaussians
    (oracle/torch_cpu_rasterizer.py, autograd backward) on config C1 -- the bundled tiramisu_ds COLMAP points as
    Gaussians, 800x450 -- one forward+backward on all host cores; next to it this library on the same scene."""
    import torch
    import gsr_scenes as S
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import torch_cpu_rasterizer as T
    from diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer
    sc = S.scene_C1(os.path.join(ROOT, "tests", "golden", "data", "c1_tiramisu_points.npz"))
    dpix_np = S.upstream_grad(sc["W"], sc["H"], 1)
    torch.set_num_threads(os.cpu_count() or 1)
    t0 = time.time()
    color_cpu, _, _, info = T.fwd_bwd(sc, dpix_np)
    dt = time.time() - t0
    t = {k: torch.from_numpy(np.ascontiguousarray(sc[k])).to(dev) for k in TENSORS}
    dpix = torch.from_numpy(dpix_np).to(dev)
    raster = GaussianRasterizer(GaussianRasterizationSettings(
        sc["H"], sc["W"], sc["tanfovx"], sc["tanfovy"], t["bg"], 1.0, t["viewmatrix"], t["projmatrix"],
        sc["sh_degree"], t["campos"], False, False))

    def one():
        leaves = {k: t[k].detach().requires_grad_(True) for k in LEAVES}
        color, _ = raster(means3D=leaves["means3D"], means2D=torch.zeros_like(leaves["means3D"]),
                          opacities=leaves["opacities"], shs=leaves["sh"], scales=leaves["scales"],
                          rotations=leaves["rotations"])
        color.backward(dpix)
        return color
    for _ in range(5):
        color = one()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(50):
        one()
    b.record()
    torch.cuda.synchronize()
    ms = a.elapsed_time(b) / 50
    return {"value": 1.0 / dt, "unit": UNIT, "cores": torch.get_num_threads(), "kind": "port",
            "implementation": "pure PyTorch fp32 on the CPU, autograd backward (oracle/torch_cpu_rasterizer.py)",
            "sample": "1 fwd+bwd of config C1 (33,730 COLMAP points as Gaussians, SH3, 800x450, %d instances), %.1f s"
                      % (info["num_rendered"], dt),
            "this_library_same_scene": {"value": 1000.0 / ms, "unit": UNIT, "ms_per_step": ms,
                                        "max_abs_image_diff_vs_pytorch": float(
                                            (color.detach().cpu() - color_cpu).abs().max())}}


def run_reference(args, rank, world):
    """Reference arm: the UNMODIFIED reference rasterizer (oracle/_ref, its own CUDA kernels) on the same box and
    workload; if that library is absent, the CPU oracle port."""
    if rank != 0:
        return None
    sc, dpix = make_scene(0, 1)
    line = {"impl": "reference", "metric": METRIC, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic", "config": {"workload": WORKLOAD, "P": P_GAUSS, "W": WIDTH, "H": HEIGHT}}
    from checkers import REF_SO
    import torch
    if os.path.exists(REF_SO) and torch.cuda.is_available():
        from checkers import Reference
        from util import to_device
        torch.cuda.set_device(0)
        d = to_device(sc)
        g = torch.from_numpy(dpix).cuda()
        ref = Reference()

        def step():
            ref.forward(d, internals=False, sync=False)
            ref.backward(g, sync=False)
        for _ in range(max(args.warmup, 3)):
            step()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(args.steps):
            step()
        b.record()
        torch.cuda.synchronize()
        ms = a.elapsed_time(b) / args.steps
        v = 1000.0 / ms
        line.update(value=v, ms_per_step=ms, gaussians_per_s=v * P_GAUSS,
                    cpu_baseline={"value": v, "unit": UNIT, "cores": 0, "kind": "reference",
                                  "sample": "the unmodified reference CUDA rasterizer (oracle/_ref, built from "
                                            "/root/reference sources) on the same B200, full workload, inputs "
                                            "resident; it has no CPU implementation"})
    else:
        cb = cpu_baseline(sc, dpix)
        line.update(value=cb["value"], ms_per_step=1000.0 / cb["value"], gaussians_per_s=cb["value"] * P_GAUSS,
                    cpu_baseline=cb)
    line["e2e"] = {"value": line["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    return line


def dp_check(step, t, dpix, rank, world, dev):
    """N > 1, once before timing: the gradients every rank holds after one exchanged step against the sum of the N
    views' gradients rendered by THIS rank alone, one view after the other (the batch sum of
    train_stacked_transformer.py:203-222 on one device).  -> max over ranks of max|a - b| / max|b| per tensor."""
    import torch
    import torch.distributed as dist
    import gsr_scenes as S
    from diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer
    _, got = step(t)
    got = [g.clone() for g in got]
    rasters = []
    for r in range(world):
        cam = S.look_at_camera(WIDTH, HEIGHT, 60.0, -15.0 + 30.0 * r / (world - 1))
        c = {k: torch.from_numpy(cam[k]).to(dev) for k in ("viewmatrix", "projmatrix", "campos")}
        rasters.append(GaussianRasterizer(GaussianRasterizationSettings(
            HEIGHT, WIDTH, cam["tanfovx"], cam["tanfovy"], t["bg"], 1.0, c["viewmatrix"], c["projmatrix"], SH_DEG,
            c["campos"], False, False)))

    def batch_sum():
        want = None
        for raster in rasters:
            leaves = {k: t[k].detach().requires_grad_(True) for k in LEAVES}
            color, _ = raster(means3D=leaves["means3D"], means2D=torch.zeros_like(leaves["means3D"]),
                              opacities=leaves["opacities"], shs=leaves["sh"], scales=leaves["scales"],
                              rotations=leaves["rotations"])
            color.backward(dpix)
            g = [leaves[k].grad for k in LEAVES]
            want = g if want is None else [a + b for a, b in zip(want, g)]
        return want
    # The blend backward's fp32 atomics are order-nondeterministic: two single-rank sums of the same batch differ
    # by up to ~1e-3 of max on the ill-conditioned tensors (scales, rotations) at this size.  The bar per tensor is
    # therefore max(1e-3, 3 x that self-deviation), as in tests/test_gpu_parity.py.
    want, again = batch_sum(), batch_sum()
    rel = lambda a, b: (a - b).abs().max() / b.abs().max().clamp_min(1e-30)
    errs = torch.stack([rel(a, b) for a, b in zip(got, want)])
    self_dev = torch.stack([rel(a, b) for a, b in zip(again, want)])
    dist.all_reduce(errs, op=dist.ReduceOp.MAX)
    dist.all_reduce(self_dev, op=dist.ReduceOp.MAX)
    return {"max_rel_err": {k: float(e) for k, e in zip(LEAVES, errs)},
            "single_rank_self_deviation": {k: float(e) for k, e in zip(LEAVES, self_dev)},
            "bar": {k: max(1e-3, 3.0 * float(e)) for k, e in zip(LEAVES, self_dev)}}


def c3_record(rank, world, dev, steps=6, warmup=2, peer=True, multimem=True, gather="push"):
    """BASELINE configs[2] ("C3") next to the C2 line: 3 M Gaussians, a batch of 8 views on a 30-degree arc at
    1080p, STRONG scaling -- rank 0 alone renders all 8 views (t_1), then the batch is split over the N ranks with
    the gsr_dp exchange (t_N); efficiency = t_1 / (N * t_N).  Device time, max over ranks."""
    import torch
    import torch.distributed as dist
    import gsr_dp
    import gsr_scenes as S
    from diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer
    P, views = 3_000_000, 8
    sc = S.scene_S(P, WIDTH, HEIGHT, seed=0)
    t = {k: torch.from_numpy(np.ascontiguousarray(sc[k])).to(dev) for k in LEAVES + ("bg",)}
    dpix = torch.from_numpy(S.upstream_grad(WIDTH, HEIGHT, 1)).to(dev)
    rasters = []
    for yaw in np.linspace(-15.0, 15.0, views):
        cam = S.look_at_camera(WIDTH, HEIGHT, 60.0, float(yaw))
        c = {k: torch.from_numpy(cam[k]).to(dev) for k in ("viewmatrix", "projmatrix", "campos")}
        rasters.append(GaussianRasterizer(GaussianRasterizationSettings(
            HEIGHT, WIDTH, cam["tanfovx"], cam["tanfovy"], t["bg"], 1.0, c["viewmatrix"], c["projmatrix"], SH_DEG,
            c["campos"], False, False)))

    def render_backward(v):
        leaves = {k: t[k].detach().requires_grad_(True) for k in LEAVES}
        color, _ = rasters[v](means3D=leaves["means3D"], means2D=torch.zeros_like(leaves["means3D"]),
                              opacities=leaves["opacities"], shs=leaves["sh"], scales=leaves["scales"],
                              rotations=leaves["rotations"])
        color.backward(dpix)
        return {k: leaves[k].grad for k in LEAVES}

    def timed(r_, w_):
        shapes = {k: tuple(t[k].shape) for k in gsr_dp.SMALL_LEAVES}
        if peer and w_ > 1:
            ex = gsr_dp.PeerExchange(views, P, shapes, dev, r_, w_, use_multimem=multimem, gather=gather)
            flat = ex.flat
        else:
            flat = gsr_dp.FlatGrads(shapes, dev)
            ex = gsr_dp.FactorExchange((views + w_ - 1) // w_, P, dev, w_)

        def rb(v, _ex):
            with gsr_dp.capture_sh_factors(ex), gsr_dp.grads_into(flat):
                return render_backward(v)

        def one():
            if peer and w_ > 1:
                return gsr_dp.dp_step_peer(rb, views, ex, t["means3D"], SH_DEG, M_COEF)
            return gsr_dp.dp_step_factored(rb, views, flat, r_, w_, t["means3D"], SH_DEG, M_COEF, exchange=ex)
        for _ in range(warmup):
            g = one()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(steps):
            g = one()
        b.record()
        torch.cuda.synchronize()
        return a.elapsed_time(b) / steps, float(sum(g[k].double().abs().sum() for k in LEAVES))

    dist.barrier()
    t1 = torch.zeros(2, device=dev, dtype=torch.float64)
    if rank == 0:  # the single-GPU reference time of the same batch (no collective: world = 1 inside gsr_dp)
        ms1, sum1 = timed(0, 1)
        t1 += torch.tensor([ms1, sum1], device=dev, dtype=torch.float64)
    dist.barrier()
    dist.broadcast(t1, 0)
    msN, sumN = timed(rank, world)
    tt = torch.tensor([msN], device=dev)
    dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    msN = float(tt[0])
    return {"workload": "C3: scene S, 3M Gaussians SH3, batch of 8 views on a 30-degree arc, 1920x1080, fwd+bwd, "
                        "batch split over the ranks (strong scaling)",
            "ms_per_batch": msN, "ms_per_batch_1gpu": float(t1[0]), "views_per_s": 8 * 1000.0 / msN,
            "gaussians_per_s": P * 8 * 1000.0 / msN, "efficiency_vs_1": float(t1[0]) / (world * msN),
            "grad_abs_sum": sumN, "grad_abs_sum_1gpu": float(t1[1]), "steps": steps,
            "exchange": "gsr_dp.PeerExchange" if peer else "gsr_dp.FactorExchange (NCCL all-gather of SH factors per "
                                                           "view slot + all-reduce of 44 B/Gaussian)"}


def run_ours(args, rank, world, local_rank):
    import torch
    import torch.distributed as dist
    from diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer, _lib
    dev = torch.device("cuda", local_rank)
    torch.cuda.set_device(dev)
    sc, dpix_np = make_scene(rank, world)
    t = {k: torch.from_numpy(np.ascontiguousarray(sc[k])).to(dev) for k in TENSORS}
    # e2e leg: the Gaussians are replicated, so rank r keeps only ITS 1/N slice of every parameter array in pinned
    # host memory; per step it copies that slice and the ranks all-gather the rest over NVLink (N = 1: the whole array)
    rows = (P_GAUSS + world - 1) // world
    lo = rank * rows
    host = {k: torch.from_numpy(np.ascontiguousarray(sc[k][lo:lo + rows])).pin_memory() for k in LEAVES}
    dpix = torch.from_numpy(dpix_np).to(dev)
    settings = GaussianRasterizationSettings(HEIGHT, WIDTH, sc["tanfovx"], sc["tanfovy"], t["bg"], 1.0,
                                             t["viewmatrix"], t["projmatrix"], SH_DEG, t["campos"], False, False)
    raster = GaussianRasterizer(settings)
    loss_host = torch.zeros(1).pin_memory()
    stats = {}
    flat = None
    pg_param = None
    if world > 1:
        import gsr_dp
        keep = LEAVES if args.dp_exchange == "allreduce" else gsr_dp.SMALL_LEAVES
        peer = None
        if args.dp_exchange == "peer":
            # symmetric memory (peer mapping over NVLink) may be unavailable on a box: every rank must take the same
            # path, so the ranks agree on the outcome and fall back to the NCCL factor exchange together
            try:
                peer = gsr_dp.PeerExchange(world, P_GAUSS, {k: tuple(t[k].shape) for k in keep}, dev, rank, world,
                                           use_multimem=args.multimem, gather=args.peer_gather)
                ok = torch.ones(1, device=dev)
            except Exception as exc:  # noqa: BLE001
                print("bench.py: peer exchange unavailable on rank %d (%s); using --dp-exchange factors" % (rank, exc),
                      file=sys.stderr)
                ok = torch.zeros(1, device=dev)
            dist.all_reduce(ok, op=dist.ReduceOp.MIN)
            if float(ok) == 0.0:
                peer, args.dp_exchange = None, "factors"
        if peer is not None:
            flat = peer.flat
        else:
            flat = gsr_dp.FlatGrads({k: tuple(t[k].shape) for k in keep}, dev)
        exchange = gsr_dp.FactorExchange(1, P_GAUSS, dev, world) if args.dp_exchange == "factors" else None
        pg_param = dist.new_group(backend="nccl")  # own communicator: parameter all-gathers of the e2e leg

    want = {"loss": False}

    def render_backward(src):
        leaves = {k: src[k].detach().requires_grad_(True) for k in LEAVES}
        m2 = torch.zeros_like(leaves["means3D"], requires_grad=True)
        color, radii = raster(means3D=leaves["means3D"], means2D=m2, opacities=leaves["opacities"], shs=leaves["sh"],
                              scales=leaves["scales"], rotations=leaves["rotations"])
        # The upstream gradient dL/dpixel is handed to the backward directly, as the reference arm does
        # (ref.backward(dL_dpix)); the step's scalar result <color, dL_dpix> is only formed where it is read back
        # (the e2e leg).
        loss = torch.dot(color.detach().view(-1), dpix.view(-1)) if want["loss"] else None
        color.backward(dpix)
        stats["radii"] = radii
        return loss, {k: leaves[k].grad for k in LEAVES}

    def step(src):
        """One step: this rank's view forward + backward, then (N > 1) the gradient exchange of gsr_dp.py."""
        if world == 1:
            loss, g = render_backward(src)
            return loss, [g[k] for k in LEAVES]
        box = {}
        if args.dp_exchange == "allreduce":
            def rb(_v):
                with gsr_dp.grads_into(flat):
                    box["loss"], g = render_backward(src)
                return g
            g = gsr_dp.dp_step(rb, world, flat, rank, world)
        elif args.dp_exchange == "peer":
            def rb(_v, _ex):
                with gsr_dp.capture_sh_factors(peer), gsr_dp.grads_into(flat):
                    box["loss"], g = render_backward(src)
                return g
            g = gsr_dp.dp_step_peer(rb, world, peer, src["means3D"], SH_DEG, M_COEF)
        else:
            def rb(_v, _ex):
                with gsr_dp.capture_sh_factors(exchange), gsr_dp.grads_into(flat):
                    box["loss"], g = render_backward(src)
                return g
            g = gsr_dp.dp_step_factored(rb, world, flat, rank, world, src["means3D"], SH_DEG, M_COEF,
                                        exchange=exchange)
        return box["loss"], [g[k] for k in LEAVES]

    def sync_all():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(args.warmup, 3)):
        step(t)
    sync_all()
    check = dp_check(step, t, dpix, rank, world, dev) if world > 1 else None
    sync_all()
    if args.timeline:
        from torch.profiler import ProfilerActivity, profile
        with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as tp:
            for _ in range(2):
                step(t)
            sync_all()
        if rank == 0:
            ev = sorted((e for e in tp.events() if e.device_type.name == "CUDA"), key=lambda e: e.time_range.start)
            t0 = ev[0].time_range.start
            with open(args.timeline, "w") as f:
                for e in ev:
                    f.write("%9.1f %8.1f  %s\n" % (e.time_range.start - t0, e.time_range.end - e.time_range.start,
                                                    e.name[:90]))
        return None
    # ---- timed region 1: inputs resident in HBM ("value"); per-stage device times recorded by the library.  The
    # clock sampler runs over BOTH timed regions (region 1 alone is shorter than a few nvidia-smi periods).
    clocks = ClockSampler(local_rank) if rank == 0 else None
    _lib.profile_enable(True)
    _lib.launch_count(reset=True)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(args.steps):
        step(t)
    b.record()
    sync_all()
    launches = _lib.launch_count(reset=True)  # counted inside the library at every launch site
    ms = a.elapsed_time(b) / args.steps
    prof = _lib.profile_read()
    _lib.profile_enable(False)
    # ---- timed region 2: end to end from pinned host buffers through the same public API.  Every step's inputs are
    # copied host->device inside the timed region; the copy (+ all-gather) of step s+1 (copy stream, second buffer
    # set) overlaps the compute of step s, and every step ends with a device->host read of its loss.
    copy_stream = torch.cuda.Stream(device=dev)
    bufs = [{k: torch.empty((rows * world,) + tuple(t[k].shape[1:]), device=dev) for k in LEAVES} for _ in range(2)]
    ready = [torch.cuda.Event() for _ in range(2)]
    consumed = [torch.cuda.Event() for _ in range(2)]

    def prefetch(s_):
        b = s_ & 1
        with torch.cuda.stream(copy_stream):
            copy_stream.wait_event(consumed[b])  # the step that last used this buffer set has finished
            for k in LEAVES:
                bufs[b][k][lo:lo + host[k].shape[0]].copy_(host[k], non_blocking=True)
            if world > 1:
                for k in LEAVES:
                    dist.all_gather_into_tensor(bufs[b][k], bufs[b][k][lo:lo + rows], group=pg_param)
            ready[b].record(copy_stream)

    def e2e_run(n_):
        cur = torch.cuda.current_stream()
        for b in range(2):
            consumed[b].record(cur)
        prefetch(0)
        for s_ in range(n_):
            if s_ + 1 < n_:
                prefetch(s_ + 1)
            b = s_ & 1
            cur.wait_event(ready[b])
            want["loss"] = True
            loss, _ = step({k: bufs[b][k][:P_GAUSS] for k in LEAVES})
            consumed[b].record(cur)
            loss_host.copy_(loss.detach().reshape(1), non_blocking=True)
            cur.synchronize()
    e2e_run(3)
    sync_all()
    a2, b2 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    n_e2e = max(3, args.steps // 2)
    a2.record()
    e2e_run(n_e2e)
    b2.record()
    sync_all()
    want["loss"] = False
    clk = clocks.stop() if clocks else None
    ms_e2e = a2.elapsed_time(b2) / n_e2e
    if world > 1:
        tt = torch.tensor([ms, ms_e2e], device=dev)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        ms, ms_e2e = float(tt[0]), float(tt[1])
    c3 = c3_record(rank, world, dev, peer=args.c3_exchange == "peer", multimem=args.multimem,
                   gather=args.peer_gather) \
        if (world > 1 and not args.no_c3) else None
    if rank != 0:
        return None
    # ---- bookkeeping for the JSON line
    radii = stats["radii"]
    V = int((radii > 0).sum())
    R = int(_lib_last_R(raster, t))
    N = WIDTH * HEIGHT
    K = (SH_DEG + 1) ** 2
    a_total, a_stage = algorithmic_bytes(P_GAUSS, V, R, N, K, M_COEF)
    peak, peak_kind = peaks()
    per_stage_ms = {s: (v[0] / max(v[1], 1)) for s, v in prof.items()}
    dom = max(a_stage, key=lambda s: per_stage_ms.get(s, 0.0))
    dom_ms = per_stage_ms[dom]
    achieved = a_stage[dom] / (dom_ms * 1e-3) / 1e9
    h2d = sum(host[k].numel() * 4 for k in LEAVES)
    same_view = (V == NCU_CAPTURE["V"] and R == NCU_CAPTURE["R"])
    value = world * 1000.0 / ms
    # the HBM-bound kernels against the same roof (the dominant kernel is an issue-bound blend kernel)
    hbm_kernels = {s: {"algorithmic_bytes": a_stage[s], "kernel_ms": per_stage_ms[s],
                       "achieved": a_stage[s] / (per_stage_ms[s] * 1e-3) / 1e9,
                       "frac": a_stage[s] / (per_stage_ms[s] * 1e-3) / 1e9 / peak}
                   for s in ("preprocess_fwd", "preprocess_bwd") if per_stage_ms.get(s)}
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic", "gaussians_per_s": value * P_GAUSS,
        "config": {"workload": WORKLOAD,
                   "exchange": "none" if world == 1 else "NCCL all-reduce of 236 B/Gaussian gradients"
                   if args.dp_exchange == "allreduce" else
                   "NCCL all-reduce of 44 B/Gaussian + all-gather of 12 B/Gaussian/view SH factors, SH gradient "
                   "rebuilt locally" if args.dp_exchange == "factors" else
                   "SH factor records (12 B/Gaussian/view) gathered by %s and summed by our rebuild kernel; all-reduce "
                   "of 44 B/Gaussian by our %s kernel over the symmetric allocation"
                   % ({"nccl": "one NCCL all-gather", "copy": "peer-to-peer copies", "fused": "remote reads inside "
                       "the rebuild kernel", "push": "our push kernel (NVLink stores into every peer's buffer)"}[
                          args.peer_gather],
                      "NVSwitch multimem" if (peer is not None and peer.flat_mc) else "two-shot P2P"),
                   "P": P_GAUSS, "visible": V, "num_rendered": R, "W": WIDTH, "H": HEIGHT,
                   "l2": "inputs (236 MB) + scratch exceed the 126 MB L2; no explicit flush"},
        "clocks": clk,
        "e2e": {"value": world * 1000.0 / ms_e2e, "unit": UNIT, "h2d_bytes_per_step": h2d,
                "h2d_bytes_per_step_all_ranks": h2d * world, "d2h_bytes_per_step": 4, "ms_per_step": ms_e2e,
                "inputs": "each rank copies its 1/N slice of the replicated parameters from pinned host memory; "
                          "NCCL all-gather over NVLink fills in the rest" if world > 1 else
                          "all parameters copied from pinned host memory every step"},
        "gpu_launches": launches,
        "gpu_launches_per_step": launches / args.steps,
        "roofline": {"bound": "hbm", "kernel": dom, "achieved": achieved, "peak": peak, "peak_kind": peak_kind,
                     "unit": "GB/s", "frac": achieved / peak,
                     "traffic": NCU_DRAM_BYTES_C2.get(dom) if same_view else None,
                     "traffic_source": NCU_CAPTURE["source"] + " (ncu --set full, same view)" if same_view else None,
                     "algorithmic_bytes": a_stage[dom], "kernel_ms": dom_ms,
                     "note": "alpha blending is FP32/MUFU-pipe bound, not HBM bound; the HBM fraction is reported "
                             "because BASELINE.json's metric asks for it; roofline_hbm_kernels lists the two "
                             "HBM-bound kernels against the same roof"},
        "roofline_hbm_kernels": hbm_kernels,
        "roofline_step": {"algorithmic_bytes": a_total, "achieved": a_total / (ms * 1e-3) / 1e9, "peak": peak,
                          "unit": "GB/s", "frac": a_total / (ms * 1e-3) / 1e9 / peak,
                          "frac_of_nominal_8000": a_total / (ms * 1e-3) / 1e9 / 8000.0},
        "stage_ms": per_stage_ms,
    }
    if check is not None:
        check["ok"] = all(check["max_rel_err"][k] <= check["bar"][k] for k in LEAVES)
        check["against"] = "sum of the N views' gradients rendered on one rank (bar: max(1e-3, 3 x the deviation " \
                           "between two such single-rank sums))"
        line["dp_check"] = check
    if c3 is not None:
        line["c3"] = c3
    if world == 1 and not args.no_cpu_baseline:
        line["cpu_baseline"] = cpu_baseline(sc, dpix_np)
        line["cpu_baseline_pytorch"] = cpu_baseline_pytorch(dev)
    return line


def _lib_last_R(raster, t):
    """num_rendered of the benchmark view (one extra forward through the _C-shaped shim)."""
    import torch
    from diff_gaussian_rasterization import _C
    e = torch.Tensor([])
    s = raster.raster_settings
    return _C.rasterize_gaussians(s.bg, t["means3D"], e, t["opacities"], t["scales"], t["rotations"], 1.0, e,
                                  s.viewmatrix, s.projmatrix, s.tanfovx, s.tanfovy, s.image_height, s.image_width,
                                  t["sh"], s.sh_degree, s.campos, False, False)[0]


def main():
    args = parse()
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local_rank = int(os.environ.get("LOCAL_RANK", 0))
    if args.impl == "reference":
        line = run_reference(args, rank, world)
        if line is not None:
            print(json.dumps(line), flush=True)
        return
    import torch
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a GPU (the product path has no CPU fallback)")
    if world > 1:
        import torch.distributed as dist
        torch.cuda.set_device(local_rank)
        # high-priority NCCL stream: the SH-factor all-gather is launched between the two backward stages and must
        # get SMs while preprocess_bwd's CTAs are still being dispatched
        opts = dist.ProcessGroupNCCL.Options(is_high_priority_stream=True)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank), pg_options=opts)
    line = run_ours(args, rank, world, local_rank)
    if line is not None:
        print(json.dumps(line), flush=True)
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
