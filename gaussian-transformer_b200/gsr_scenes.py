"""Seeded synthetic scenes for parity tests and bench.py (numpy, CPU).

Scene "S" is the workload of BASELINE.json configs 2/3/5 as made concrete in
SURVEY.md section 8(d); camera matrices follow the reference's conventions
(scene/cameras.py:54-57, utils/graphics_utils.py:38-71): viewmatrix = W2C^T,
projmatrix = W2C^T @ P^T, both read column-major by the kernels.
"""
import math

import numpy as np

SH_C0 = 0.28209479177387814


def projection_matrix(znear, zfar, tanx, tany):
    """utils/graphics_utils.py:51-71 getProjectionMatrix (as numpy float32)."""
    P = np.zeros((4, 4), dtype=np.float32)
    top, right = tany * znear, tanx * znear
    P[0, 0] = 2.0 * znear / (2 * right)
    P[1, 1] = 2.0 * znear / (2 * top)
    P[3, 2] = 1.0
    P[2, 2] = zfar / (zfar - znear)
    P[2, 3] = -(zfar * znear) / (zfar - znear)
    return P


def look_at_camera(W, H, fovx_deg=60.0, yaw_deg=0.0, target=(0.0, 0.0, 6.0), znear=0.01, zfar=100.0):
    """Camera on a circle around `target` (radius = |target|), yaw 0 = at the origin looking +z."""
    tanx = math.tan(math.radians(fovx_deg) / 2)
    tany = tanx * H / W
    yaw = math.radians(yaw_deg)
    tgt = np.asarray(target, dtype=np.float64)
    dist = float(np.linalg.norm(tgt))
    # camera centre rotated about the target's vertical axis
    c = tgt + dist * np.array([-math.sin(yaw), 0.0, -math.cos(yaw)])
    fwd = tgt - c
    fwd /= np.linalg.norm(fwd)
    up = np.array([0.0, 1.0, 0.0])
    right = np.cross(up, fwd)
    right /= np.linalg.norm(right)
    up2 = np.cross(fwd, right)
    Rwc = np.stack([right, up2, fwd])  # rows: camera axes in world coords (world -> camera)
    W2C = np.eye(4)
    W2C[:3, :3] = Rwc
    W2C[:3, 3] = -Rwc @ c
    view = W2C.T.astype(np.float32)  # world_view_transform
    proj = (view @ projection_matrix(znear, zfar, tanx, tany).T).astype(np.float32)  # full_proj_transform
    campos = np.linalg.inv(view.astype(np.float64))[3, :3].astype(np.float32)
    return dict(W=W, H=H, tanfovx=float(tanx), tanfovy=float(tany), viewmatrix=view, projmatrix=proj,
                campos=campos)


def scene_S(P, W, H, seed=0, M=16, sh_degree=3, yaw_deg=0.0, log_scale_mean=math.log(0.01), bg=(0.0, 0.0, 0.0)):
    """SURVEY.md 8(d) synthetic scene: returns a dict of float32 numpy arrays + camera."""
    rng = np.random.default_rng(seed)
    cam = look_at_camera(W, H, 60.0, yaw_deg)
    tanx, tany = cam["tanfovx"], cam["tanfovy"]
    z = rng.uniform(2.0, 10.0, P)
    near = rng.random(P) < 0.02
    z = np.where(near, rng.uniform(-1.0, 0.2, P), z)
    zz = np.where(near, 1.0, z)
    x = zz * tanx * rng.uniform(-1.1, 1.1, P)
    y = zz * tany * rng.uniform(-1.1, 1.1, P)
    means3D = np.stack([x, y, z], 1).astype(np.float32)
    scales = np.exp(rng.normal(log_scale_mean, 0.7, (P, 3))).astype(np.float32)
    q = rng.normal(0, 1, (P, 4))
    rotations = (q / np.linalg.norm(q, axis=1, keepdims=True)).astype(np.float32)
    opacities = (1.0 / (1.0 + np.exp(-rng.normal(-1.0, 2.0, (P, 1))))).astype(np.float32)
    sh = np.zeros((P, M, 3), dtype=np.float32)
    sh[:, 0, :] = (rng.random((P, 3)) - 0.5) / SH_C0
    if M > 1:
        sh[:, 1:, :] = rng.normal(0, 0.1, (P, M - 1, 3))
    out = dict(cam)
    out.update(P=P, M=M, sh_degree=sh_degree, means3D=means3D, scales=scales, rotations=rotations,
               opacities=opacities, sh=sh, bg=np.asarray(bg, dtype=np.float32), scale_modifier=1.0)
    return out


def upstream_grad(W, H, seed=1):
    """dL/dcolor = U(-1,1)/(H*W), dense."""
    rng = np.random.default_rng(seed)
    return (rng.uniform(-1, 1, (3, H, W)) / (H * W)).astype(np.float32)


def covariance_from_scale_rot(scales, rotations, mod=1.0):
    """(P,6) upper triangle of R S^2 R^T (scene/gaussian_model.py:27-31) for the cov3D_precomp path."""
    r, x, y, z = (rotations[:, i].astype(np.float64) for i in range(4))
    R = np.empty((len(r), 3, 3))
    R[:, 0, 0] = 1 - 2 * (y * y + z * z); R[:, 0, 1] = 2 * (x * y - r * z); R[:, 0, 2] = 2 * (x * z + r * y)
    R[:, 1, 0] = 2 * (x * y + r * z); R[:, 1, 1] = 1 - 2 * (x * x + z * z); R[:, 1, 2] = 2 * (y * z - r * x)
    R[:, 2, 0] = 2 * (x * z - r * y); R[:, 2, 1] = 2 * (y * z + r * x); R[:, 2, 2] = 1 - 2 * (x * x + y * y)
    L = R * (mod * scales.astype(np.float64))[:, None, :]
    S = L @ L.transpose(0, 2, 1)
    return np.stack([S[:, 0, 0], S[:, 0, 1], S[:, 0, 2], S[:, 1, 1], S[:, 1, 2], S[:, 2, 2]], 1).astype(np.float32)


def scene_C1(points_npz, W=800, H=450, M=16, sh_degree=3):
    """BASELINE config 1: the bundled tiramisu_ds COLMAP points as Gaussians, initialised exactly like
    GaussianModel.create_from_pcd (scene/gaussian_model.py:124-147): DC = RGB2SH(colour) (utils/sh_utils.py:114-115),
    rest 0, scale = sqrt(mean 3-NN distance^2) on all axes (clamped at 1e-7 before the root), rotation (1,0,0,0),
    opacity 0.1.  Camera: COLMAP world frame, identity pose (the cloud sits in front of it), focal 3287.46 px at
    4032 px width (cameras.bin) rescaled to W (utils/camera_utils.py:35-39).  `points_npz`: the fixture written by
    tests/golden/make_c1_points.py."""
    z = np.load(points_npz)
    xyz, rgb, dist2 = z["xyz"].astype(np.float32), z["rgb"], z["dist2"].astype(np.float32)
    P = xyz.shape[0]
    fovx = 2.0 * math.atan(4032.0 / (2.0 * 3287.46))
    cam = look_at_camera(W, H, math.degrees(fovx), 0.0)
    sh = np.zeros((P, M, 3), dtype=np.float32)
    sh[:, 0, :] = ((rgb.astype(np.float32) / 255.0) - 0.5) / SH_C0
    s = np.sqrt(np.maximum(dist2, 1e-7)).astype(np.float32)
    rot = np.zeros((P, 4), dtype=np.float32)
    rot[:, 0] = 1.0
    out = dict(cam)
    out.update(P=P, M=M, sh_degree=sh_degree, means3D=xyz, scales=np.repeat(s[:, None], 3, 1).copy(), rotations=rot,
               opacities=np.full((P, 1), 0.1, dtype=np.float32), sh=sh, bg=np.zeros(3, dtype=np.float32),
               scale_modifier=1.0)
    return out
