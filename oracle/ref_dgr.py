"""ORACLE -- TEST INFRASTRUCTURE ONLY.

The reference rasteriser as the package `diff_gaussian_rasterization`, on the GPU box: its OWN Python front-end
(submodules/diff-gaussian-rasterization/diff_gaussian_rasterization/__init__.py, copied unmodified into
baseline/_ref/ref_dgr by oracle/install_ref_py.sh) over a `_C` module with the three functions of its compiled
extension (ext.cpp:15-19), implemented here on oracle/_ref (the unmodified CUDA sources behind oracle/ref_shim.cu)
instead of its torch binding rasterize_points.cu.  Used by the drop-in tests to run the reference's render() /
bake_set once on the reference rasteriser and once on libghr, and compare.
"""
import importlib
import os
import sys
import types

import torch

from . import ref_cuda, ref_py

_handles = {}   # id -> RefRasterizer (one per outstanding forward: render() keeps seven before the first backward)
_next = [1]


def _opt(t):
    return None if t is None or t.numel() == 0 else t.contiguous()


def _rasterize_gaussians(background, means3D, colors, opacity, scales, rotations, scale_modifier, cov3D_precomp, viewmatrix,
                         projmatrix, tan_fovx, tan_fovy, image_height, image_width, sh, degree, campos, prefiltered, debug):
    rr = ref_cuda.RefRasterizer()
    n, color, depth, alpha, radii = rr.forward(background.contiguous(), means3D.contiguous(), _opt(colors),
                                               opacity.contiguous(), _opt(scales), _opt(rotations), float(scale_modifier),
                                               _opt(cov3D_precomp), viewmatrix.contiguous(), projmatrix.contiguous(),
                                               float(tan_fovx), float(tan_fovy), int(image_height), int(image_width),
                                               _opt(sh), int(degree), campos.contiguous(), bool(prefiltered), bool(debug))
    hid = _next[0]
    _next[0] += 1
    _handles[hid] = rr
    if len(_handles) > 64:  # forwards whose backward never came (inference)
        for k in sorted(_handles)[:-32]:
            del _handles[k]
    tag = torch.tensor([hid], dtype=torch.int64, device=means3D.device).view(torch.uint8)
    e = torch.empty(0, dtype=torch.uint8, device=means3D.device)
    return n, color, depth, alpha, radii, tag, e, e


def _rasterize_gaussians_backward(background, means3D, radii, colors, scales, rotations, scale_modifier, cov3D_precomp,
                                  viewmatrix, projmatrix, tan_fovx, tan_fovy, dL_dout_color, dL_dout_depth, dL_dout_alpha,
                                  sh, degree, campos, geomBuffer, R, binningBuffer, imageBuffer, alphas, debug):
    hid = int(geomBuffer.view(torch.int64)[0])
    rr = _handles.pop(hid)
    g = rr.backward(background.contiguous(), means3D.contiguous(), radii, _opt(colors), _opt(scales), _opt(rotations),
                    float(scale_modifier), _opt(cov3D_precomp), viewmatrix.contiguous(), projmatrix.contiguous(),
                    float(tan_fovx), float(tan_fovy), dL_dout_color, dL_dout_depth, dL_dout_alpha, _opt(sh), int(degree),
                    campos.contiguous(), alphas.contiguous(), bool(debug))
    return (g["dL_dmeans2D"], g["dL_dcolors"], g["dL_dopacity"], g["dL_dmeans3D"], g["dL_dcov3D"], g["dL_dsh"],
            g["dL_dscales"], g["dL_drotations"])


def _mark_visible(means3D, viewmatrix, projmatrix):
    return ref_cuda.mark_visible(means3D.contiguous(), viewmatrix.contiguous(), projmatrix.contiguous())


def install():
    """Make `import diff_gaussian_rasterization` resolve to the reference's front-end over oracle/_ref.  Returns the
    package.  (Undo with sys.modules.pop or by installing another implementation.)"""
    if not ref_cuda.available():
        raise FileNotFoundError(ref_cuda.LIB_PATH)
    pkg_dir = os.path.join(ref_py.REF_PY, "ref_dgr")
    if not os.path.exists(os.path.join(pkg_dir, "diff_gaussian_rasterization", "__init__.py")):
        raise FileNotFoundError(pkg_dir + ": reference rasteriser front-end not installed (oracle/install_ref_py.sh)")
    for name in ("diff_gaussian_rasterization", "diff_gaussian_rasterization._C"):
        sys.modules.pop(name, None)
    c = types.ModuleType("diff_gaussian_rasterization._C")
    c.rasterize_gaussians = _rasterize_gaussians
    c.rasterize_gaussians_backward = _rasterize_gaussians_backward
    c.mark_visible = _mark_visible
    sys.modules["diff_gaussian_rasterization._C"] = c
    spec = importlib.util.spec_from_file_location(
        "diff_gaussian_rasterization", os.path.join(pkg_dir, "diff_gaussian_rasterization", "__init__.py"),
        submodule_search_locations=[os.path.join(pkg_dir, "diff_gaussian_rasterization")])
    pkg = importlib.util.module_from_spec(spec)
    sys.modules["diff_gaussian_rasterization"] = pkg
    pkg._C = c
    spec.loader.exec_module(pkg)
    return pkg
