"""GPU tests of the Python surface (GaussianRasterizer / rasterize_gaussians / markVisible): every argument
combination of DGR/diff_gaussian_rasterization/__init__.py:187-220, autograd wiring, edge cases."""
import contextlib

import numpy as np
import pytest
import torch

import cases
import gsr_scenes as S
from util import np_, rel_err, settings_from, to_device

pytestmark = pytest.mark.gpu


def _module(d, **kw):
    from diff_gaussian_rasterization import GaussianRasterizer
    return GaussianRasterizer(settings_from(d, **kw))


def _leaf(t):
    return t.clone().requires_grad_(True)


def test_autograd_matches_reference(reference):
    sc, _ = cases.case("s_deg3")
    d = to_device(sc)
    m3, sh, op, scl, rot = (_leaf(d[k]) for k in ("means3D", "sh", "opacities", "scales", "rotations"))
    m2 = torch.zeros_like(m3, requires_grad=True)
    color, radii = _module(d)(means3D=m3, means2D=m2, opacities=op, shs=sh, scales=scl, rotations=rot)
    assert color.shape == (3, sc["H"], sc["W"]) and radii.dtype == torch.int32 and radii.shape == (sc["P"],)
    dpix = torch.from_numpy(S.upstream_grad(sc["W"], sc["H"], 5)).cuda()
    (color * dpix).sum().backward()
    ref = reference.forward(d, internals=False)
    g = reference.backward(dpix)
    assert np.abs(np_(color) - np_(ref["color"])).max() <= 1e-5
    for t, k in ((m3, "means3D"), (m2, "means2D"), (sh, "sh"), (op, "opacity"), (scl, "scales"), (rot, "rot")):
        assert t.grad is not None and t.grad.shape == t.shape, k
        assert rel_err(np_(t.grad), np_(g[k]).reshape(t.shape)) <= 1e-3, k
    assert float(m2.grad[:, 2].abs().max()) == 0.0


def test_sh_factor_exchange_matches_plain_sh_gradient():
    """gsr_dp's SH path: backward without dL/dsh + (campos, factors) per view + local rebuild over two views must
    equal the sum of the two views' ordinary SH gradients; the other gradients are unchanged."""
    import gsr_dp
    grads, items_all, means = [], [], None
    for yaw in (-10.0, 12.0):
        sc = S.scene_S(3000, 320, 208, seed=5, M=16, sh_degree=3, yaw_deg=yaw)
        d = to_device(sc)
        dpix = torch.from_numpy(S.upstream_grad(sc["W"], sc["H"], 5)).cuda()
        out = []
        for capture in (False, True):
            m3, sh, op, scl, rot = (_leaf(d[k]) for k in ("means3D", "sh", "opacities", "scales", "rotations"))
            items = []
            if capture:
                with gsr_dp.capture_sh_factors() as items:
                    color, _ = _module(d)(means3D=m3, means2D=torch.zeros_like(m3), opacities=op, shs=sh, scales=scl,
                                          rotations=rot)
                    (color * dpix).sum().backward()
                assert sh.grad is None and len(items) == 1
                items_all.append(items[0])
            else:
                color, _ = _module(d)(means3D=m3, means2D=torch.zeros_like(m3), opacities=op, shs=sh, scales=scl,
                                      rotations=rot)
                (color * dpix).sum().backward()
            out.append((m3.grad, op.grad, scl.grad, rot.grad, sh.grad))
        for a, b in zip(out[0][:4], out[1][:4]):
            assert rel_err(np_(b), np_(a)) <= 1e-3  # atomics order only
        grads.append(out[0][4])
        means = d["means3D"]
    cam, fac = gsr_dp.gather_sh_factors(items_all, 2, 3000, means.device, 1)
    from diff_gaussian_rasterization import _C
    rebuilt = _C.sh_grad_from_factors(means, cam, fac, 3, 16)
    want = grads[0] + grads[1]
    assert float(want.abs().max()) > 0
    assert rel_err(np_(rebuilt), np_(want)) <= 1e-3
    again = _C.sh_grad_from_factors(means, cam, fac, 3, 16, out=rebuilt.clone())  # accumulate mode
    assert rel_err(np_(again), np_(2 * rebuilt)) <= 1e-6
    # generic (non-staged) variant: M = 4, degree 1
    r4 = _C.sh_grad_from_factors(means, cam, fac, 1, 4)
    assert rel_err(np_(r4), np_(_C.sh_grad_from_factors(means, cam, fac, 1, 16)[:, :4])) <= 1e-6


def test_factor_exchange_sink_and_staged_backward():
    """A FactorExchange as the sink: factors land in its buffer between the two backward stages and equal both
    the list sink's and gsr_sh_color_factors(dL_dcolors); the staged backward equals the one-call backward."""
    import gsr_dp
    from diff_gaussian_rasterization import _C
    sc = S.scene_S(3000, 320, 208, seed=6, M=16, sh_degree=3, yaw_deg=4.0)
    d = to_device(sc)
    dpix = torch.from_numpy(S.upstream_grad(sc["W"], sc["H"], 6)).cuda()
    got = []
    for sink in (None, [], gsr_dp.FactorExchange(1, 3000, d["means3D"].device, 1)):
        m3, sh, op, scl, rot = (_leaf(d[k]) for k in ("means3D", "sh", "opacities", "scales", "rotations"))
        with gsr_dp.capture_sh_factors(sink) if sink is not None else contextlib.nullcontext():
            color, _ = _module(d)(means3D=m3, means2D=torch.zeros_like(m3), opacities=op, shs=sh, scales=scl,
                                  rotations=rot)
            (color * dpix).sum().backward()
        got.append((m3.grad, op.grad, scl.grad, rot.grad))
    for k in (1, 2):
        for a, b in zip(got[0], got[k]):
            assert rel_err(np_(b), np_(a)) <= 1e-3
    ex = sink
    cam, fac = ex.finish()
    assert cam.shape == (1, 3) and fac.shape == (1, 3000, 3)
    np.testing.assert_array_equal(np_(cam[0]), sc["campos"].reshape(3))
    # list sink of the second run: same factors up to the order of the blend atomics
    # (both come from the accumulators of their own backward)
    e = torch.Tensor([])
    R, c, radii, geom, binning, img = _C.rasterize_gaussians(
        d["bg"], d["means3D"], e, d["opacities"], d["scales"], d["rotations"], 1.0, e, d["viewmatrix"],
        d["projmatrix"], d["tanfovx"], d["tanfovy"], sc["H"], sc["W"], d["sh"], 3, d["campos"], False, False)
    g = _C.rasterize_gaussians_backward(d["bg"], d["means3D"], radii, e, d["scales"], d["rotations"], 1.0, e,
                                        d["viewmatrix"], d["projmatrix"], d["tanfovx"], d["tanfovy"], dpix, d["sh"],
                                        3, d["campos"], geom, R, binning, img, False)
    want = _C.sh_color_factors(geom, g[1])
    assert float(want.abs().max()) > 0
    assert rel_err(np_(fac[0]), np_(want)) <= 1e-3


def test_dp_step_factored_two_views_one_rank():
    """gsr_dp.dp_step_factored with the real rasterizer, two views on one rank (slots 0 and 1 of the exchange, the
    first view written straight into the flat buffer, the second added): equals the plain sum of both backward
    passes for all five parameters."""
    import gsr_dp
    P, W, H = 3000, 320, 208
    scs = [S.scene_S(P, W, H, seed=7, M=16, sh_degree=3, yaw_deg=y) for y in (-8.0, 9.0)]
    ds = [to_device(sc) for sc in scs]
    dpix = [torch.from_numpy(S.upstream_grad(W, H, 20 + v)).cuda() for v in range(2)]
    names = {"means3D": "means3D", "sh": "sh", "opacities": "opacities", "scales": "scales", "rotations": "rotations"}

    def run(v, sink_ctx):
        d = ds[v]
        leaves = {k: _leaf(d[k]) for k in names}
        with sink_ctx:
            color, _ = _module(d)(means3D=leaves["means3D"], means2D=torch.zeros_like(leaves["means3D"]),
                                  opacities=leaves["opacities"], shs=leaves["sh"], scales=leaves["scales"],
                                  rotations=leaves["rotations"])
            color.backward(dpix[v])
        return {k: leaves[k].grad for k in names}

    plain = [run(v, contextlib.nullcontext()) for v in range(2)]
    flat = gsr_dp.FlatGrads({k: tuple(ds[0][k].shape) for k in gsr_dp.SMALL_LEAVES}, "cuda")
    ex = gsr_dp.FactorExchange(2, P, ds[0]["means3D"].device, 1)

    class both:
        def __init__(self, *c): self.c = c
        def __enter__(self): [c.__enter__() for c in self.c]
        def __exit__(self, *e): [c.__exit__(*e) for c in reversed(self.c)]

    for _ in range(2):  # second step: buffers are reused
        out = gsr_dp.dp_step_factored(
            lambda v, e: {k: g for k, g in run(v, both(gsr_dp.capture_sh_factors(e), gsr_dp.grads_into(flat))).items()
                          if k in gsr_dp.SMALL_LEAVES},
            2, flat, 0, 1, ds[0]["means3D"], 3, 16, exchange=ex)
    for k in names:
        want = plain[0][k] + plain[1][k]
        assert float(want.abs().max()) > 0
        assert rel_err(np_(out[k]).reshape(want.shape), np_(want)) <= 1e-3, k


def test_cta_scatter_kernel_writes_the_same_lists():
    """bin_scatter_cta_kernel (taken automatically once point_list outgrows the L2) against bin_pass<true>: identical
    per-tile lists, image and n_contrib."""
    from diff_gaussian_rasterization import _C, _lib
    sc = S.scene_S(80000, 640, 360, seed=9, M=16, sh_degree=3)
    d = to_device(sc)
    e = torch.Tensor([])
    res = []
    for force in (0, 1, -1):
        _lib.set_option("bin_cta", force)
        if force < 0:
            break
        res.append(_C.rasterize_gaussians(d["bg"], d["means3D"], e, d["opacities"], d["scales"], d["rotations"], 1.0, e,
                                          d["viewmatrix"], d["projmatrix"], d["tanfovx"], d["tanfovy"], sc["H"],
                                          sc["W"], d["sh"], 3, d["campos"], False, False))
    (R0, c0, _, _, b0, i0), (R1, c1, _, _, b1, i1) = res
    N = sc["W"] * sc["H"]
    assert R0 == R1 and R0 > 100000
    assert torch.equal(b0[:4 * R0], b1[:4 * R0]) and torch.equal(c0, c1) and torch.equal(i0[:8 * N], i1[:8 * N])


def test_speculative_forward_hit_miss_and_overflow_guard():
    """gsr_forward launches phase 2 before num_rendered reaches the host, against a buffer sized from a hint: a
    hit, a miss (hint far too small: the device-side guard must keep the kernels from touching the short buffer)
    and the no-hint first call all return the same bits."""
    from diff_gaussian_rasterization import _C
    sc = S.scene_S(150000, 640, 360, seed=8, M=16, sh_degree=3)
    d = to_device(sc)
    e = torch.Tensor([])
    dev = (d["means3D"].device.index, sc["P"], sc["W"], sc["H"])

    def fwd():
        return _C.rasterize_gaussians(d["bg"], d["means3D"], e, d["opacities"], d["scales"], d["rotations"], 1.0, e,
                                      d["viewmatrix"], d["projmatrix"], d["tanfovx"], d["tanfovy"], sc["H"], sc["W"],
                                      d["sh"], 3, d["campos"], False, False)
    _C._r_hint.pop(dev, None)
    R0, c0, radii0, _, b0, i0 = fwd()                  # no hint: phase 1, round trip, phase 2
    assert R0 > 1000 and _C._r_hint[dev] == R0
    R1, c1, radii1, _, b1, i1 = fwd()                  # hit: speculative launch
    assert b1.numel() > b0.numel()                     # over-allocated by the 25 % margin
    _C._r_hint[dev] = 10                               # miss: capacity of ~65k entries < R
    if R0 <= 10 + 2 + 65536:
        pytest.skip("scene too sparse to overflow the minimum capacity")
    R2, c2, radii2, _, b2, i2 = fwd()
    assert R0 == R1 == R2
    N = sc["W"] * sc["H"]
    for c, r, b, i in ((c1, radii1, b1, i1), (c2, radii2, b2, i2)):
        assert torch.equal(c, c0)
        assert torch.equal(r, radii0)
        assert torch.equal(b[:4 * R0], b0[:4 * R0])    # the per-tile lists
        assert torch.equal(i[:8 * N], i0[:8 * N])      # final_T, n_contrib (the rest has uninitialised padding)
    assert _C._r_hint[dev] == R0


@pytest.mark.parametrize("P,M,deg", [(3001, 16, 3), (3001, 16, 1), (2048, 4, 1), (700, 9, 2), (513, 16, 0)])
def test_split_sh_matches_concatenated(P, M, deg):
    """shs=(features_dc, features_rest): same bits forward, same gradients as through torch.cat (SURVEY 8f rank 3;
    M = 9 is not built and must fall back to the concatenation transparently)."""
    sc = S.scene_S(P, 320, 208, seed=12, M=M, sh_degree=deg)
    d = to_device(sc)
    dpix = torch.from_numpy(S.upstream_grad(sc["W"], sc["H"], 12)).cuda()
    res = []
    for split in (False, True):
        m3, op, scl, rot = (_leaf(d[k]) for k in ("means3D", "opacities", "scales", "rotations"))
        dc, rest = _leaf(d["sh"][:, :1].contiguous()), _leaf(d["sh"][:, 1:].contiguous())
        shs = (dc, rest) if split else torch.cat((dc, rest), dim=1)
        color, radii = _module(d)(means3D=m3, means2D=torch.zeros_like(m3), opacities=op, shs=shs, scales=scl,
                                  rotations=rot)
        (color * dpix).sum().backward()
        res.append((color.detach(), radii, dc.grad, rest.grad, m3.grad, op.grad, scl.grad, rot.grad))
    a, b = res
    assert torch.equal(a[0], b[0]) and torch.equal(a[1], b[1])
    assert float(a[2].abs().max()) > 0 and (deg == 0 or float(a[3].abs().max()) > 0)
    for x, y in zip(a[2:], b[2:]):
        assert x.shape == y.shape
        assert rel_err(np_(y), np_(x)) <= 1e-3  # order of the blend atomics only (same bar as the parity tests)


def test_colors_and_cov_paths():
    sc, _ = cases.case("s_deg3")
    d = to_device(sc)
    P = sc["P"]
    cols = _leaf(torch.rand(P, 3, device="cuda"))
    cov = _leaf(torch.from_numpy(S.covariance_from_scale_rot(sc["scales"], sc["rotations"])).cuda())
    m3, op = _leaf(d["means3D"]), _leaf(d["opacities"])
    color, radii = _module(d)(means3D=m3, means2D=torch.zeros_like(m3), opacities=op, colors_precomp=cols,
                              cov3D_precomp=cov)
    color.sum().backward()
    assert cols.grad.shape == (P, 3) and cov.grad.shape == (P, 6) and m3.grad.shape == (P, 3)
    assert torch.isfinite(cols.grad).all() and float(cols.grad.abs().sum()) > 0 and float(cov.grad.abs().sum()) > 0


def test_argument_validation():
    sc, _ = cases.case("s_deg0_m1")
    d = to_device(sc)
    r = _module(d)
    m3, op = d["means3D"], d["opacities"]
    with pytest.raises(Exception, match="excatly one of either SHs or precomputed colors"):
        r(means3D=m3, means2D=m3, opacities=op, scales=d["scales"], rotations=d["rotations"])
    with pytest.raises(Exception, match="excatly one of either SHs or precomputed colors"):
        r(means3D=m3, means2D=m3, opacities=op, shs=d["sh"], colors_precomp=m3, scales=d["scales"],
          rotations=d["rotations"])
    with pytest.raises(Exception, match="scale/rotation pair or precomputed 3D covariance"):
        r(means3D=m3, means2D=m3, opacities=op, shs=d["sh"], scales=d["scales"])
    with pytest.raises(Exception, match="scale/rotation pair or precomputed 3D covariance"):
        r(means3D=m3, means2D=m3, opacities=op, shs=d["sh"], scales=d["scales"], rotations=d["rotations"],
          cov3D_precomp=torch.zeros(sc["P"], 6, device="cuda"))
    with pytest.raises(RuntimeError, match=r"means3D must have dimensions \(num_points, 3\)"):
        r(means3D=m3[:, :2], means2D=m3, opacities=op, shs=d["sh"], scales=d["scales"], rotations=d["rotations"])
    with pytest.raises(RuntimeError):
        r(means3D=m3.cpu(), means2D=m3, opacities=op, shs=d["sh"], scales=d["scales"], rotations=d["rotations"])


def test_empty_input_gives_zero_image_not_bg():
    sc, _ = cases.case("s_bg_yaw")
    d = to_device(sc)
    z = lambda *s: torch.zeros(*s, device="cuda")
    color, radii = _module(d)(means3D=z(0, 3), means2D=z(0, 3), opacities=z(0, 1), shs=z(0, 16, 3), scales=z(0, 3),
                              rotations=z(0, 4))
    assert radii.numel() == 0 and float(color.abs().max()) == 0.0  # rasterize_points.cu:66,80-114


def test_all_culled_is_background_and_zero_grads():
    sc, _ = cases.case("all_culled")
    sc["bg"] = np.array([0.2, 0.4, 0.6], np.float32)
    d = to_device(sc)
    m3 = _leaf(d["means3D"])
    color, radii = _module(d)(means3D=m3, means2D=torch.zeros_like(m3), opacities=d["opacities"], shs=d["sh"],
                              scales=d["scales"], rotations=d["rotations"])
    assert int(radii.max()) == 0
    assert torch.allclose(color, d["bg"].view(3, 1, 1).expand_as(color))
    color.sum().backward()
    assert float(m3.grad.abs().max()) == 0.0


def test_strided_views_like_box_sort_consumers(reference):
    """model/box_sort.py:17-27 hands the rasterizer column slices of one (P,26) tensor."""
    sc, _ = cases.case("s_deg1_m4")
    d = to_device(sc)
    P = sc["P"]
    flat = torch.cat([d["means3D"], d["scales"], d["rotations"], d["opacities"], d["sh"].reshape(P, 12)], 1)
    flat = _leaf(flat)
    m3, scl, rot, op, sh = flat[:, 0:3], flat[:, 3:6], flat[:, 6:10], flat[:, 10:11], flat[:, 11:23].reshape(P, 4, 3)
    color, _ = _module(d)(means3D=m3, means2D=torch.zeros(P, 3, device="cuda"), opacities=op, shs=sh, scales=scl,
                          rotations=rot)
    ref = reference.forward(d, internals=False)
    assert np.abs(np_(color) - np_(ref["color"])).max() <= 1e-5
    color.sum().backward()
    assert flat.grad.shape == (P, 23) and torch.isfinite(flat.grad).all()


def test_mark_visible(reference, oracle):
    sc, _ = cases.case("s_bg_yaw")
    d = to_device(sc)
    ours = _module(d).markVisible(d["means3D"])
    assert ours.dtype == torch.bool
    assert torch.equal(ours, reference.mark_visible(d["means3D"], d["viewmatrix"], d["projmatrix"]))
    assert (np_(ours) == oracle.mark_visible(sc["means3D"], sc["viewmatrix"], sc["projmatrix"])).all()


def test_non_default_stream_and_repeat_calls():
    sc, _ = cases.case("s_deg3")
    d = to_device(sc)
    r = _module(d)
    kw = dict(means3D=d["means3D"], means2D=torch.zeros_like(d["means3D"]), opacities=d["opacities"], shs=d["sh"],
              scales=d["scales"], rotations=d["rotations"])
    base, _ = r(**kw)
    st = torch.cuda.Stream()
    torch.cuda.synchronize()
    with torch.cuda.stream(st):
        other, _ = r(**kw)
    st.synchronize()
    assert torch.equal(base, other)


def test_sh_degree_growth_uses_stride_M():
    """active_sh_degree grows during training while the stride stays M (train.py:72-73)."""
    sc, _ = cases.case("s_deg3")
    d = to_device(sc)
    imgs = []
    for deg in (0, 1, 2, 3):
        d["sh_degree"] = deg
        c, _ = _module(d)(means3D=d["means3D"], means2D=torch.zeros_like(d["means3D"]), opacities=d["opacities"],
                          shs=d["sh"], scales=d["scales"], rotations=d["rotations"])
        imgs.append(c)
    assert not torch.equal(imgs[0], imgs[3])


def _render_leaves(d, **kw):
    m3, sh, op, scl, rot = (_leaf(d[k]) for k in ("means3D", "sh", "opacities", "scales", "rotations"))
    color, radii = _module(d, **kw)(means3D=m3, means2D=torch.zeros_like(m3), opacities=op, shs=sh, scales=scl,
                                    rotations=rot)
    return color, radii, (m3, sh, op, scl, rot)


def test_prefiltered_violation_raises():
    """prefiltered=True promises that no point is behind the near plane; the reference prints and __trap()s the
    kernel (auxiliary.h:154-162), which surfaces as a CUDA error = RuntimeError.  Here the forward raises
    RuntimeError after the kernel flagged the violation (defined deviation 3), and a scene
    that keeps the promise renders the same bits with and without the flag."""
    sc = S.scene_S(2000, 160, 96, seed=21)  # 2 % of the points sit behind the near plane
    d = to_device(sc)
    with pytest.raises(RuntimeError, match="point is filtered although prefiltered is set"):
        _render_leaves(d, prefiltered=True)
    keep = sc["means3D"][:, 2] > 0.5
    for k in ("means3D", "sh", "opacities", "scales", "rotations"):
        sc[k] = np.ascontiguousarray(sc[k][keep])
    sc["P"] = int(keep.sum())
    d = to_device(sc)
    c0, r0, _ = _render_leaves(d, prefiltered=False)
    c1, r1, _ = _render_leaves(d, prefiltered=True)
    assert torch.equal(c0, c1) and torch.equal(r0, r1)


def test_debug_snapshots_on_failure(tmp_path, monkeypatch):
    """debug=True contract of DGR/diff_gaussian_rasterization/__init__.py:83-90,132-139: when the native call fails,
    the argument tuple is written to snapshot_fw.dump / snapshot_bw.dump in the working directory and the error is
    re-raised; the dumps replay through the _C-shaped entry points.  Without debug nothing is written."""
    from diff_gaussian_rasterization import _C
    monkeypatch.chdir(tmp_path)
    sc = S.scene_S(1500, 160, 96, seed=22)
    d = to_device(sc)
    # forward failure: a genuine one (prefiltered violation detected by the kernel)
    with pytest.raises(RuntimeError):
        _render_leaves(d, prefiltered=True, debug=False)
    assert not (tmp_path / "snapshot_fw.dump").exists()
    with pytest.raises(RuntimeError):
        _render_leaves(d, prefiltered=True, debug=True)
    fw = torch.load(tmp_path / "snapshot_fw.dump", weights_only=False)
    assert len(fw) == 19 and fw[1].shape == (1500, 3) and fw[17] is True and fw[18] is True
    args = [a.cuda() if isinstance(a, torch.Tensor) else a for a in fw]
    args[17] = False  # the same tuple with the promise withdrawn replays through the pybind-shaped entry
    R, color, radii, *_ = _C.rasterize_gaussians(*args)
    c_ok, r_ok, _ = _render_leaves(d)
    assert R > 0 and torch.equal(color, c_ok.detach()) and torch.equal(radii, r_ok)
    # backward failure: the native backward is made to fail (an invalid library status)
    color, radii, leaves = _render_leaves(d, debug=True)
    real = _C.rasterize_gaussians_backward

    def failing(*a, **k):
        raise RuntimeError("gsrast: forced failure")
    monkeypatch.setattr(_C, "rasterize_gaussians_backward", failing)
    with pytest.raises(RuntimeError, match="forced failure"):
        color.sum().backward()
    monkeypatch.setattr(_C, "rasterize_gaussians_backward", real)
    bw = torch.load(tmp_path / "snapshot_bw.dump", weights_only=False)
    assert len(bw) == 21 and bw[2].dtype == torch.int32 and bw[12].shape == color.shape
    args = [a.cuda() if isinstance(a, torch.Tensor) else a for a in bw]
    grads = real(*args)  # (dL_dmeans2D, dL_dcolors, dL_dopacity, dL_dmeans3D, dL_dcov3D, dL_dsh, dL_dscales, dL_drot)
    color2, _, leaves2 = _render_leaves(d)
    color2.sum().backward()
    assert rel_err(np_(grads[3]), np_(leaves2[0].grad)) <= 1e-3 and rel_err(np_(grads[5]), np_(leaves2[1].grad)) <= 1e-3


def test_launch_counter_counts_kernels():
    from diff_gaussian_rasterization import _lib
    sc = S.scene_S(1500, 160, 96, seed=23)
    d = to_device(sc)
    _lib.launch_count(reset=True)
    color, _, _ = _render_leaves(d)
    n_fwd = _lib.launch_count()
    color.sum().backward()
    n_all = _lib.launch_count(reset=True)
    assert 4 <= n_fwd <= 24 and n_fwd + 2 <= n_all <= n_fwd + 4, (n_fwd, n_all)
    assert _lib.launch_count() == 0


def test_packed_expf_is_bit_identical_to_expf():
    """render_fwd2_kernel evaluates exp() for its two pixels with a hand-scheduled copy of nvcc's expf sequence (so
    that half of it packs into FFMA2 / FADD2 / FMUL2).  n_contrib and final_T are bit-exact contracts, so the copy
    must return expf's bits: every float in [-104, -1e-30] and [1e-30, 104] (by bit pattern) and their negations."""
    import struct
    from diff_gaussian_rasterization import _lib
    bits = lambda f: struct.unpack("<I", struct.pack("<f", f))[0]
    out = torch.zeros(1, dtype=torch.int64, device="cuda")
    for lo, hi in ((bits(1e-30), bits(104.0)), (0, 4096), (bits(104.0), bits(1e30))):
        _lib.check(_lib.lib.gsr_selftest_expf(lo, hi, out.data_ptr(), torch.cuda.current_stream().cuda_stream))
        torch.cuda.synchronize()
        assert int(out.item()) == 0, (hex(lo), hex(hi), int(out.item()))


def test_two_pixel_forward_equals_one_pixel_forward():
    from diff_gaussian_rasterization import _lib
    from util import ours_forward
    for name in ("s_deg3", "s_big", "s_dense_opaque", "nonfinite", "one_huge", "s_huge_lists"):
        sc, _ = cases.case(name)
        d = to_device(sc)
        res = []
        for v in (1, 2):
            _lib.set_option("render_fwd", v)
            res.append(ours_forward(d))
        _lib.set_option("render_fwd", 2)
        a = res[0]
        for b in res[1:]:
            assert torch.equal(a["n_contrib"], b["n_contrib"]), name
            assert torch.equal(a["final_T"].view(torch.int32), b["final_T"].view(torch.int32)), name
            c1, c2 = a["color"], b["color"]
            assert torch.equal(torch.nan_to_num(c1), torch.nan_to_num(c2)) and torch.equal(c1.isnan(), c2.isnan()), name


def test_pair_backward_equals_default_backward():
    """render_bwd2_kernel<false, true> (small scenes: two list entries per step of the hit loop, option render_bwd = 4)
    against the default instantiation (5 = PAIR off) on the same forward: same operations per entry, so the gradients
    differ by the order of the atomics only."""
    from diff_gaussian_rasterization import _lib
    from test_gpu_parity import _check_grads
    from util import ours_backward, ours_forward
    for name in ("s_deg3", "s_big", "s_dense_opaque", "one_huge", "s_huge_lists", "c1_tiramisu"):
        sc, _ = cases.case(name)
        d = to_device(sc)
        dpix = torch.from_numpy(S.upstream_grad(sc["W"], sc["H"], seed=7)).cuda()
        fwd = ours_forward(d)
        res = []
        for v in (5, 4):
            _lib.set_option("render_bwd", v)
            res.append(ours_backward(d, fwd, dpix))
        _lib.set_option("render_bwd", 2)
        _check_grads(res[1], res[0], sc=sc)


def test_bulk_async_preprocess_equals_one_tile_kernel():
    """preprocess_fwd_bulk_kernel (persistent CTAs, cp.async.bulk + mbarrier double buffering; option pre_fwd = 2)
    against the default one-tile-per-CTA kernel: every forward output bit-identical, incl. a partial last tile."""
    from diff_gaussian_rasterization import _lib
    from util import ours_forward
    for P, W, H in ((128 * 37 + 5, 320, 200), (128 * 500, 640, 360), (100, 160, 96)):
        sc = S.scene_S(P, W, H, seed=41)
        d = to_device(sc)
        res = []
        for v in (1, 2):
            _lib.set_option("pre_fwd", v)
            res.append(ours_forward(d))
        _lib.set_option("pre_fwd", 1)
        a, b = res
        for k in ("radii", "means2D", "depths", "conic_opacity", "rgb", "clamped", "tiles_touched", "point_list",
                  "ranges", "n_contrib", "color"):
            assert torch.equal(torch.nan_to_num(a[k].float()), torch.nan_to_num(b[k].float())), (P, k)
