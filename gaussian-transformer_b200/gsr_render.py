"""`gsr_render.render`: `gaussian_renderer.render()` (gaussian_renderer/__init__.py:18-100 of the reference) with the
step in front of the rasterizer folded into it: the rasterizer is handed the model's RAW
`_opacity`, `_scaling`, `_rotation` (sigmoid / exp / normalize of scene/gaussian_model.py:95-106 run inside
preprocess_fwd, their chain rule inside preprocess_bwd) and the two SH arrays `_features_dc`, `_features_rest` in
place (no `torch.cat`, gradients arrive on the two parameters directly).  Same signature and return dict; rendered
images are bit-identical to the reference's render() on this library.  Any configuration the fold does not cover
(`pipe.compute_cov3D_python`, `pipe.convert_SHs_python`, `override_color`, SH layouts other than 16 / 4 coefficients)
takes the reference's own statements.

    from gsr_render import render            # instead of: from gaussian_renderer import render
"""
import math

import torch

from diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer


def render(viewpoint_camera, pc, pipe, bg_color, scaling_modifier=1.0, override_color=None):
    screenspace_points = torch.zeros_like(pc.get_xyz, dtype=pc.get_xyz.dtype, requires_grad=True, device="cuda") + 0
    try:
        screenspace_points.retain_grad()
    except Exception:
        pass
    settings = GaussianRasterizationSettings(
        image_height=int(viewpoint_camera.image_height), image_width=int(viewpoint_camera.image_width),
        tanfovx=math.tan(viewpoint_camera.FoVx * 0.5), tanfovy=math.tan(viewpoint_camera.FoVy * 0.5), bg=bg_color,
        scale_modifier=scaling_modifier, viewmatrix=viewpoint_camera.world_view_transform,
        projmatrix=viewpoint_camera.full_proj_transform, sh_degree=pc.active_sh_degree,
        campos=viewpoint_camera.camera_center, prefiltered=False, debug=pipe.debug)
    rasterizer = GaussianRasterizer(raster_settings=settings)
    folded = (override_color is None and not pipe.compute_cov3D_python and not pipe.convert_SHs_python
              and pc._features_rest.size(1) + 1 in (4, 16))
    if folded:
        image, radii = rasterizer(means3D=pc.get_xyz, means2D=screenspace_points,
                                  shs=(pc._features_dc, pc._features_rest), colors_precomp=None,
                                  opacities=pc._opacity, scales=pc._scaling, rotations=pc._rotation,
                                  cov3D_precomp=None, raw_params=True)
    else:
        import gaussian_renderer  # the reference's own statements for everything else
        return gaussian_renderer.render(viewpoint_camera, pc, pipe, bg_color, scaling_modifier, override_color)
    return {"render": image, "viewspace_points": screenspace_points, "visibility_filter": radii > 0, "radii": radii}
