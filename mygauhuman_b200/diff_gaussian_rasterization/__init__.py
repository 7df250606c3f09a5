"""Operator API of the tile rasteriser, interface-compatible with the reference
package ``diff_gaussian_rasterization``
(/root/reference/submodules/diff-gaussian-rasterization/diff_gaussian_rasterization/__init__.py):

* ``GaussianRasterizationSettings``  (reference :160-172) -- same 12 fields, same order
* ``GaussianRasterizer``             (reference :174-224) -- ``forward`` / ``markVisible``
* ``rasterize_gaussians``            (reference :21-42)
* ``_C``                             -- ``rasterize_gaussians`` / ``rasterize_gaussians_backward`` / ``mark_visible``

so ``gaussian_renderer.render()`` (reference gaussian_renderer/__init__.py:203-272)
and ``baking.bake_set`` (reference baking.py:259-282) run unchanged on top of
libghr.so.  forward returns ``(color[3,H,W], radii[P], depth[1,H,W], alpha[1,H,W])``;
backward consumes ``(grad_color, grad_radii, grad_depth, grad_alpha)``.
"""
from typing import NamedTuple

import torch
import torch.nn as nn

from . import _C

_SNAPSHOT = {"fw": "snapshot_fw.dump", "bw": "snapshot_bw.dump"}


class GaussianRasterizationSettings(NamedTuple):
    image_height: int
    image_width: int
    tanfovx: float
    tanfovy: float
    bg: torch.Tensor
    scale_modifier: float
    viewmatrix: torch.Tensor
    projmatrix: torch.Tensor
    sh_degree: int
    campos: torch.Tensor
    prefiltered: bool
    debug: bool


def _host_copy(args):
    return tuple(a.detach().cpu().clone() if torch.is_tensor(a) else a for a in args)


def _call_native(fn, args, debug, which):
    """Run a native entry point; in debug mode dump the inputs when it fails
    (reference __init__.py:83-90,135-142)."""
    if not debug:
        return fn(*args)
    saved = _host_copy(args)
    try:
        return fn(*args)
    except Exception:
        torch.save(saved, _SNAPSHOT[which])
        print(f"\nAn error occured in {'forward' if which == 'fw' else 'backward'}. "
              f"Writing {_SNAPSHOT[which]} for debugging.\n")
        raise


class _RasterizeGaussians(torch.autograd.Function):
    @staticmethod
    def forward(ctx, means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp,
                raster_settings):
        rs = raster_settings
        native_args = (rs.bg, means3D, colors_precomp, opacities, scales, rotations, rs.scale_modifier,
                       cov3Ds_precomp, rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy, rs.image_height,
                       rs.image_width, sh, rs.sh_degree, rs.campos, rs.prefiltered, rs.debug)
        (num_rendered, color, depth, alpha, radii, geom_state, binning_state,
         image_state) = _call_native(_C.rasterize_gaussians, native_args, rs.debug, "fw")
        ctx.raster_settings = rs
        ctx.num_rendered = num_rendered
        ctx.save_for_backward(colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh, geom_state,
                              binning_state, image_state, alpha)
        ctx.mark_non_differentiable(radii)
        return color, radii, depth, alpha

    @staticmethod
    def backward(ctx, grad_color, grad_radii, grad_depth, grad_alpha):
        rs = ctx.raster_settings
        (colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh, geom_state, binning_state,
         image_state, alpha) = ctx.saved_tensors
        native_args = (rs.bg, means3D, radii, colors_precomp, scales, rotations, rs.scale_modifier, cov3Ds_precomp,
                       rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy, grad_color, grad_depth, grad_alpha, sh,
                       rs.sh_degree, rs.campos, geom_state, ctx.num_rendered, binning_state, image_state, alpha,
                       rs.debug)
        (g_means2D, g_colors, g_opacities, g_means3D, g_cov3D, g_sh, g_scales,
         g_rotations) = _call_native(_C.rasterize_gaussians_backward, native_args, rs.debug, "bw")
        # one gradient per forward input, in forward's order; raster_settings gets None
        return g_means3D, g_means2D, g_sh, g_colors, g_opacities, g_scales, g_rotations, g_cov3D, None


class _RasterizeGaussiansMulti(torch.autograd.Function):
    """Several attributes over one set of Gaussians: projection, sort and tile binning once, one blend per attribute
    (what the reference's render(), gaussian_renderer/__init__.py:203-272, does with one full rasteriser call each)."""

    @staticmethod
    def forward(ctx, raster_settings, means3D, means2D, sh, colors_precomp, opacities, scales, rotations,
                cov3Ds_precomp, *extra_colors):
        rs = raster_settings
        native_args = (rs.bg, means3D, colors_precomp, opacities, scales, rotations, rs.scale_modifier,
                       cov3Ds_precomp, rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy, rs.image_height,
                       rs.image_width, sh, rs.sh_degree, rs.campos, rs.prefiltered, rs.debug, extra_colors)
        (num_rendered, color, depth, alpha, radii, geom_state, binning_state, image_state, extra_images,
         first_colors, packed) = _call_native(_C.rasterize_gaussians_multi, native_args, rs.debug, "fw")
        ctx.raster_settings = rs
        ctx.num_rendered = num_rendered
        ctx.has_means2D = torch.is_tensor(means2D)
        ctx.save_for_backward(colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh, geom_state,
                              binning_state, image_state, alpha, first_colors, packed, *extra_colors)
        ctx.mark_non_differentiable(radii)
        ctx.set_materialize_grads(False)
        return (color, radii, depth, alpha) + tuple(extra_images)

    @staticmethod
    def backward(ctx, grad_color, grad_radii, grad_depth, grad_alpha, *grad_extra):
        rs = ctx.raster_settings
        (colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh, geom_state, binning_state,
         image_state, alpha, first_colors, packed) = ctx.saved_tensors[:13]
        extra_colors = ctx.saved_tensors[13:]
        zeros = lambda c: torch.zeros(c, rs.image_height, rs.image_width, dtype=torch.float32, device=means3D.device)
        grad_color = zeros(3) if grad_color is None else grad_color
        grad_depth = zeros(1) if grad_depth is None else grad_depth
        grad_alpha = zeros(1) if grad_alpha is None else grad_alpha
        native_args = (rs.bg, means3D, radii, colors_precomp, scales, rotations, rs.scale_modifier, cov3Ds_precomp,
                       rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy, grad_color, grad_depth, grad_alpha, sh,
                       rs.sh_degree, rs.campos, geom_state, ctx.num_rendered, binning_state, image_state, alpha,
                       rs.debug, first_colors, packed, extra_colors, grad_extra)
        (g_means2D, g_colors, g_opacities, g_means3D, g_cov3D, g_sh, g_scales, g_rotations,
         g_extra) = _call_native(_C.rasterize_gaussians_multi_backward, native_args, rs.debug, "bw")
        return (None, g_means3D, g_means2D if ctx.has_means2D else None, g_sh, g_colors, g_opacities, g_scales,
                g_rotations, g_cov3D) + tuple(g_extra)


def rasterize_gaussians(means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp,
                        raster_settings):
    return _RasterizeGaussians.apply(means3D, means2D, sh, colors_precomp, opacities, scales, rotations,
                                     cov3Ds_precomp, raster_settings)


def _absent():
    # the reference's convention for a missing optional input: an empty CPU tensor
    return torch.Tensor([])


class GaussianRasterizer(nn.Module):
    def __init__(self, raster_settings):
        super().__init__()
        self.raster_settings = raster_settings

    def markVisible(self, positions):
        """Boolean mask of the points in front of the camera (view-space z > 0.2)."""
        rs = self.raster_settings
        with torch.no_grad():
            return _C.mark_visible(positions, rs.viewmatrix, rs.projmatrix)

    def forward(self, means3D, means2D, opacities, shs=None, colors_precomp=None, scales=None, rotations=None,
                cov3D_precomp=None):
        if (shs is None) == (colors_precomp is None):
            raise Exception('Please provide excatly one of either SHs or precomputed colors!')
        has_scale_rot = scales is not None or rotations is not None
        incomplete_scale_rot = scales is None or rotations is None
        if (incomplete_scale_rot and cov3D_precomp is None) or (has_scale_rot and cov3D_precomp is not None):
            raise Exception('Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!')
        shs = _absent() if shs is None else shs
        colors_precomp = _absent() if colors_precomp is None else colors_precomp
        scales = _absent() if scales is None else scales
        rotations = _absent() if rotations is None else rotations
        cov3D_precomp = _absent() if cov3D_precomp is None else cov3D_precomp
        return rasterize_gaussians(means3D, means2D, shs, colors_precomp, opacities, scales, rotations, cov3D_precomp,
                                   self.raster_settings)

    def forward_multi(self, means3D, means2D, opacities, extra_colors, shs=None, colors_precomp=None, scales=None,
                      rotations=None, cov3D_precomp=None):
        """forward() plus one more image per [P,3] tensor in `extra_colors`, all blended over the same projected,
        sorted and tile-binned Gaussians.  Equivalent to calling forward() once per attribute with
        colors_precomp=extra_colors[k] (reference gaussian_renderer/__init__.py:214-272), bit for bit in the images.
        Returns (color, radii, depth, alpha, [extra images])."""
        if (shs is None) == (colors_precomp is None):
            raise Exception('Please provide excatly one of either SHs or precomputed colors!')
        has_scale_rot = scales is not None or rotations is not None
        incomplete_scale_rot = scales is None or rotations is None
        if (incomplete_scale_rot and cov3D_precomp is None) or (has_scale_rot and cov3D_precomp is not None):
            raise Exception('Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!')
        shs = _absent() if shs is None else shs
        colors_precomp = _absent() if colors_precomp is None else colors_precomp
        scales = _absent() if scales is None else scales
        rotations = _absent() if rotations is None else rotations
        cov3D_precomp = _absent() if cov3D_precomp is None else cov3D_precomp
        out = _RasterizeGaussiansMulti.apply(self.raster_settings, means3D, means2D, shs, colors_precomp, opacities,
                                             scales, rotations, cov3D_precomp, *extra_colors)
        return out[0], out[1], out[2], out[3], list(out[4:])
