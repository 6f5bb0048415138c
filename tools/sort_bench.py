"""Depth sort vs the library (VERDICT r1 item 6 / SURVEY 2.4 K5): cub::DeviceRadixSort::SortPairs<u32,u32> (CUDA 12.9
CUB, its onesweep policy on sm_100) against the product's sorts on the SAME keys, same GPU, CUDA events.

  python tools/sort_bench.py [--json out.jsonl]

Key sets: the depth keys preprocess_fwd produces for scene S at the C2 (1 M, 1080p), C3 (3 M) and C5 (6 M, 4K) sizes
(bit pattern of the view-space depth of visible Gaussians, 0xffffffff for culled ones), plus C1's 33,730.
Results are checked against torch.sort(stable=True) before anything is timed.  oracle/_build/libcubsort.so is bench
infrastructure (oracle/cub_sort.cu); the product never loads it."""
import argparse
import ctypes
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "gaussian-transformer_b200"), os.path.join(ROOT, "tests")]
import gsr_scenes as S  # noqa: E402
from diff_gaussian_rasterization import _lib  # noqa: E402


def depth_keys(P, W, H):
    sc = S.scene_S(P, W, H, seed=0)
    m, v = sc["means3D"], sc["viewmatrix"]
    z = (m[:, 0] * v[0, 2] + m[:, 1] * v[1, 2] + m[:, 2] * v[2, 2] + v[3, 2]).astype(np.float32)
    tx = np.abs(m[:, 0] / np.maximum(z, 1e-6)) < sc["tanfovx"] * 1.3
    keys = z.view(np.uint32).copy()
    keys[~((z > 0.2) & tx)] = 0xFFFFFFFF
    return torch.from_numpy(keys.view(np.int32)).cuda()


def timeit(fn, reset, n=30):
    ts = []
    for i in range(n + 5):
        reset()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        if i >= 5:
            ts.append(a.elapsed_time(b) * 1e3)
    return float(np.median(ts)), float(np.min(ts))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--json", default=None)
    args = ap.parse_args()
    cub = ctypes.CDLL(os.path.join(ROOT, "oracle", "_build", "libcubsort.so"))
    cub.cub_sort_temp_bytes.restype = ctypes.c_size_t
    cub.cub_sort_temp_bytes.argtypes = [ctypes.c_int]
    cub.cub_sort_pairs.argtypes = [ctypes.c_void_p, ctypes.c_size_t] + [ctypes.c_void_p] * 4 + [ctypes.c_int,
                                                                                                ctypes.c_int,
                                                                                                ctypes.c_void_p]
    stream = torch.cuda.current_stream().cuda_stream
    rows = []
    for name, P, W, H in (("C1", 33_730, 800, 450), ("C2", 1_000_000, 1920, 1080), ("C3", 3_000_000, 1920, 1080),
                          ("C5", 6_000_000, 3840, 2160)):
        src = depth_keys(P, W, H)
        want_k, want_v = torch.sort(src.view(torch.int32).to(torch.int64) & 0xFFFFFFFF, stable=True)
        ident = torch.arange(P, dtype=torch.int32, device="cuda")
        # --- CUB
        tb = cub.cub_sort_temp_bytes(P)
        temp = torch.empty(tb, dtype=torch.uint8, device="cuda")
        ko, vo = torch.empty_like(src), torch.empty_like(src)

        def run_cub():
            rc = cub.cub_sort_pairs(temp.data_ptr(), tb, src.data_ptr(), ko.data_ptr(), ident.data_ptr(),
                                    vo.data_ptr(), P, 32, stream)
            assert rc == 0, rc
        run_cub()
        torch.cuda.synchronize()
        assert torch.equal(vo.to(torch.int64), want_v), "CUB result differs from torch.sort"
        t_cub = timeit(run_cub, lambda: None)
        # --- ours
        nb = ctypes.c_size_t()
        _lib.check(_lib.lib.gsr_sort_scratch_bytes(P, ctypes.byref(nb)))
        scratch = torch.empty(nb.value, dtype=torch.uint8, device="cuda")
        k0, k1, v0, v1 = (torch.empty_like(src) for _ in range(4))
        res = {}
        for variant, label in ((1, "onesweep"), (0, "hist_scan_scatter")):
            def run():
                _lib.check(_lib.lib.gsr_sort_pairs(P, k0.data_ptr(), k1.data_ptr(), v0.data_ptr(), v1.data_ptr(),
                                                   scratch.data_ptr(), variant, stream))
            k0.copy_(src)
            run()
            torch.cuda.synchronize()
            assert torch.equal(v0.to(torch.int64), want_v), label + ": result differs from torch.sort"
            assert torch.equal(k0.to(torch.int64) & 0xFFFFFFFF, want_k), label + ": keys differ"
            res[label] = timeit(run, lambda: k0.copy_(src))
        row = dict(keys=name, n=P, cub_us=t_cub[0], cub_min_us=t_cub[1], onesweep_us=res["onesweep"][0],
                   onesweep_min_us=res["onesweep"][1], hist_scan_scatter_us=res["hist_scan_scatter"][0],
                   speedup_vs_cub=t_cub[0] / res["onesweep"][0])
        rows.append(row)
        print(json.dumps(row), flush=True)
    if args.json:
        with open(args.json, "w") as f:
            for r in rows:
                f.write(json.dumps(r) + "\n")


if __name__ == "__main__":
    main()
