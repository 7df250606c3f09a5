"""ORACLE -- TEST INFRASTRUCTURE ONLY.

numpy/ctypes front-end of the CPU restatement of the reference rasteriser and
distCUDA2 (oracle/raster_oracle.c, oracle/knn_oracle.c).  Only tests/,
__graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this package; nothing under mygauhuman_b200/ does.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "liboracle.so")
_lib = None


def build():
    subprocess.run(["make", "-C", _HERE, "-s"], check=True)


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            build()
        _lib = C.CDLL(LIB_PATH)
        _lib.ora_preprocess.restype = C.c_int
        _lib.ora_higher_msb.restype = C.c_uint32
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _f32(a):
    if a is None:
        return None
    a = np.ascontiguousarray(np.asarray(a, dtype=np.float32))
    return a if a.size else None


class ForwardState(dict):
    """Everything the reference keeps in geom/binning/img state, as numpy arrays."""
    __getattr__ = dict.__getitem__


def raster_forward(bg, means3D, colors_precomp, opacities, scales, rotations, scale_modifier, cov3D_precomp, view,
                   proj, tan_fovx, tan_fovy, H, W, shs, degree, campos):
    """Rasterizer::forward (rasterizer_impl.cu:198-341).  Returns ForwardState."""
    l = lib()
    means3D = _f32(means3D); P = 0 if means3D is None else means3D.shape[0]
    bg = _f32(bg); view = _f32(view); proj = _f32(proj); campos = _f32(campos)
    colors_precomp = _f32(colors_precomp); opacities = _f32(opacities); scales = _f32(scales)
    rotations = _f32(rotations); cov3D_precomp = _f32(cov3D_precomp); shs = _f32(shs)
    M = 0 if shs is None else shs.shape[1]
    st = ForwardState(P=P, H=H, W=W, M=M, D=degree)
    X = H * W
    tiles = ((W + 15) // 16) * ((H + 15) // 16)
    st["radii"] = np.zeros(P, np.int32); st["depths"] = np.zeros(P, np.float32)
    st["means2D"] = np.zeros((P, 2), np.float32); st["cov3D"] = np.zeros((P, 6), np.float32)
    st["conic_opacity"] = np.zeros((P, 4), np.float32); st["rgb"] = np.zeros((P, 3), np.float32)
    st["clamped"] = np.zeros((P, 3), np.uint8); st["tiles_touched"] = np.zeros(P, np.uint32)
    st["point_offsets"] = np.zeros(P, np.uint32)
    st["color"] = np.zeros((3, H, W), np.float32); st["depth"] = np.zeros((1, H, W), np.float32)
    st["alpha"] = np.zeros((1, H, W), np.float32); st["final_T"] = np.zeros((H, W), np.float32)
    st["n_contrib"] = np.zeros((H, W), np.uint32); st["ranges"] = np.zeros((tiles, 2), np.uint32)
    if P == 0:
        st["num_rendered"] = 0
        for k in ("keys_unsorted", "point_list_keys"):
            st[k] = np.zeros(0, np.uint64)
        for k in ("vals_unsorted", "point_list"):
            st[k] = np.zeros(0, np.uint32)
        return st
    N = l.ora_preprocess(C.c_int(P), C.c_int(degree), C.c_int(M), C.c_int(W), C.c_int(H), _p(means3D), _p(shs),
                         _p(colors_precomp), _p(opacities), _p(scales), C.c_float(scale_modifier), _p(rotations),
                         _p(cov3D_precomp), _p(view), _p(proj), _p(campos), C.c_float(tan_fovx), C.c_float(tan_fovy),
                         _p(st.radii), _p(st.depths), _p(st.means2D), _p(st.cov3D), _p(st.conic_opacity), _p(st.rgb),
                         _p(st.clamped), _p(st.tiles_touched), _p(st.point_offsets))
    st["num_rendered"] = N
    st["keys_unsorted"] = np.zeros(N, np.uint64); st["vals_unsorted"] = np.zeros(N, np.uint32)
    st["point_list_keys"] = np.zeros(N, np.uint64); st["point_list"] = np.zeros(N, np.uint32)
    l.ora_bin(C.c_int(P), C.c_int(W), C.c_int(H), C.c_int(N), _p(st.radii), _p(st.means2D), _p(st.depths),
              _p(st.point_offsets), _p(st.keys_unsorted), _p(st.vals_unsorted), _p(st.point_list_keys),
              _p(st.point_list), _p(st.ranges))
    feats = colors_precomp if colors_precomp is not None else st.rgb
    st["features"] = feats
    l.ora_blend_forward(C.c_int(W), C.c_int(H), _p(st.ranges), _p(st.point_list), _p(st.means2D), _p(feats),
                        _p(st.depths), _p(st.conic_opacity), _p(bg), _p(st.color), _p(st.depth), _p(st.alpha),
                        _p(st.final_T), _p(st.n_contrib))
    return st


def raster_backward(st, bg, means3D, colors_precomp, scales, rotations, scale_modifier, cov3D_precomp, view, proj,
                    tan_fovx, tan_fovy, dL_dcolor, dL_ddepth, dL_dalpha, shs, degree, campos):
    """Rasterizer::backward (rasterizer_impl.cu:345-446).  Returns a dict of the 8
    gradients of rasterize_points.cu:206 (+ dL_dconic), fp64-accumulated."""
    l = lib()
    P, H, W, M = st.P, st.H, st.W, st.M
    means3D = _f32(means3D); bg = _f32(bg); view = _f32(view); proj = _f32(proj); campos = _f32(campos)
    colors_precomp = _f32(colors_precomp); scales = _f32(scales); rotations = _f32(rotations)
    cov3D_precomp = _f32(cov3D_precomp); shs = _f32(shs)
    dL_dcolor = _f32(dL_dcolor); dL_ddepth = _f32(dL_ddepth); dL_dalpha = _f32(dL_dalpha)
    feats = colors_precomp if colors_precomp is not None else st.rgb
    acc = np.zeros((P, 9), np.float64)
    l.ora_blend_backward(C.c_int(W), C.c_int(H), _p(st.ranges), _p(st.point_list), _p(bg), _p(st.means2D),
                         _p(st.conic_opacity), _p(feats), _p(st.depths), _p(st.final_T), _p(st.n_contrib),
                         _p(dL_dcolor), _p(dL_ddepth), _p(dL_dalpha), _p(acc))
    g = {}
    g["dL_dcolors"] = acc[:, 0:3].astype(np.float32)
    g["dL_dmeans2D"] = np.concatenate([acc[:, 3:5], np.zeros((P, 1))], 1).astype(np.float32)
    dconic = np.zeros((P, 4), np.float32); dconic[:, 0] = acc[:, 5]; dconic[:, 1] = acc[:, 6]; dconic[:, 3] = acc[:, 7]
    g["dL_dconic"] = dconic
    g["dL_dopacity"] = acc[:, 8:9].astype(np.float32)
    g["dL_dmeans3D"] = np.zeros((P, 3), np.float32); g["dL_dcov3D"] = np.zeros((P, 6), np.float32)
    g["dL_dsh"] = np.zeros((P, M, 3), np.float32); g["dL_dscales"] = np.zeros((P, 3), np.float32)
    g["dL_drotations"] = np.zeros((P, 4), np.float32)
    cov = cov3D_precomp if cov3D_precomp is not None else st.cov3D
    dcol = np.ascontiguousarray(g["dL_dcolors"]); dm2 = np.ascontiguousarray(g["dL_dmeans2D"])
    l.ora_preprocess_backward(C.c_int(P), C.c_int(degree), C.c_int(M), C.c_int(W), C.c_int(H), _p(means3D),
                              _p(st.radii), _p(shs), _p(st.clamped), _p(scales), _p(rotations),
                              C.c_float(scale_modifier), _p(cov), _p(view), _p(proj), C.c_float(tan_fovx),
                              C.c_float(tan_fovy), _p(campos), _p(dm2), _p(dconic), _p(g["dL_dmeans3D"]), _p(dcol),
                              _p(g["dL_dcov3D"]), _p(g["dL_dsh"]), _p(g["dL_dscales"]), _p(g["dL_drotations"]))
    return g


def mark_visible(means3D, view, proj):
    means3D = _f32(means3D); P = means3D.shape[0]
    out = np.zeros(P, np.uint8)
    lib().ora_mark_visible(C.c_int(P), _p(means3D), _p(_f32(view)), _p(_f32(proj)), _p(out))
    return out.astype(bool)


def dist2_knn3(points):
    """distCUDA2 (simple-knn spatial.cu:15-26), brute force."""
    pts = _f32(points); P = pts.shape[0]
    out = np.zeros(P, np.float32)
    lib().ora_dist2_knn3(C.c_int(P), _p(pts), _p(out))
    return out


def higher_msb(n):
    return int(lib().ora_higher_msb(C.c_uint32(n)))
