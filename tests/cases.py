"""Named parity cases: small seeded scenes covering the reference's argument combinations and edge cases
(SURVEY.md A.8).  Each returns (scene dict of numpy arrays, options)."""
import math
import os

import numpy as np

import gsr_scenes as S


def _base(P=1500, W=168, H=100, **kw):
    return S.scene_S(P, W, H, **kw)


def case(name):
    opt = dict(colors_precomp=False, cov3D_precomp=False)
    if name == "s_deg3":
        sc = _base(seed=1)
    elif name == "s_deg0":
        sc = _base(seed=2, sh_degree=0)
    elif name == "s_deg1_m4":
        sc = _base(seed=3, M=4, sh_degree=1)
    elif name == "s_deg2_m9":
        sc = _base(seed=4, M=9, sh_degree=2)
    elif name == "s_deg0_m1":
        sc = _base(seed=5, M=1, sh_degree=0)
    elif name == "s_deg2_m16":
        sc = _base(seed=6, M=16, sh_degree=2)
    elif name == "s_bg_yaw":
        sc = _base(seed=7, yaw_deg=20.0, bg=(0.3, 0.6, 0.9))
    elif name == "s_colors":
        sc = _base(seed=8)
        opt["colors_precomp"] = True
    elif name == "s_cov3d":
        sc = _base(seed=9)
        opt["cov3D_precomp"] = True
    elif name == "c1_tiramisu":  # BASELINE config 1: bundled COLMAP points as Gaussians, 800x450
        sc = S.scene_C1(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "data",
                                     "c1_tiramisu_points.npz"))
    elif name == "s_big":  # large splats: long per-tile lists, early termination, big-sort path
        sc = _base(P=12000, W=160, H=96, seed=10, log_scale_mean=math.log(0.15))
    elif name == "s_huge_lists":  # lists longer than the shared-memory sort capacity (global fallback)
        sc = _base(P=30000, W=64, H=48, seed=17, log_scale_mean=math.log(1.0))
    elif name == "s_dense_opaque":  # many opaque splats: T threshold cut-off exercised everywhere
        sc = _base(P=4000, W=128, H=128, seed=11, log_scale_mean=math.log(0.05))
        sc["opacities"] = np.full_like(sc["opacities"], 0.95)
    elif name == "s_scale_mod":
        sc = _base(seed=12)
        sc["scale_modifier"] = 1.7
    elif name == "s_exact_tiles":  # W,H multiples of 16
        sc = _base(P=2000, W=128, H=64, seed=13)
    elif name == "all_culled":
        sc = _base(P=500, seed=14)
        sc["means3D"][:, 2] = -np.abs(sc["means3D"][:, 2])
    elif name == "one_huge":  # one Gaussian covering every tile
        sc = _base(P=1, W=100, H=60, seed=15)
        sc["means3D"][:] = (0.0, 0.0, 3.0)
        sc["scales"][:] = 5.0
        sc["opacities"][:] = 0.8
    elif name == "nonfinite":  # NaN / Inf parameters must be culled or propagate like the reference
        sc = _base(P=400, seed=16)
        sc["means3D"][0, 0] = np.nan
        sc["means3D"][1, 2] = np.nan
        sc["opacities"][4, 0] = np.nan
        sc["sh"][5, 0, 0] = np.nan
        sc["means3D"][6, 1] = np.inf
    elif name == "nonfinite_cov":  # NaN covariance: undefined in the reference (uninitialised sort key); ours culls
        sc = _base(P=400, seed=18)
        sc["scales"][2, 1] = np.nan
        sc["rotations"][3, 0] = np.nan
        sc["scales"][7, 0] = np.inf
    else:
        raise KeyError(name)
    return sc, opt


ALL = ["s_deg3", "s_deg0", "s_deg1_m4", "s_deg2_m9", "s_deg0_m1", "s_deg2_m16", "s_bg_yaw", "s_colors", "s_cov3d",
       "c1_tiramisu", "s_big", "s_huge_lists", "s_dense_opaque", "s_scale_mod", "s_exact_tiles", "all_culled", "one_huge", "nonfinite"]
GOLDEN = ["s_deg3", "s_deg1_m4", "s_bg_yaw", "s_colors", "s_cov3d", "one_huge"]


def extras(sc, opt):
    """colors_precomp / cov3D_precomp arrays for a case (deterministic)."""
    rng = np.random.default_rng(1234)
    colors = rng.random((sc["P"], 3)).astype(np.float32) if opt["colors_precomp"] else None
    cov = S.covariance_from_scale_rot(sc["scales"], sc["rotations"], sc["scale_modifier"]) if opt["cov3D_precomp"] else None
    return colors, cov
