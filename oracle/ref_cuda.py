"""ORACLE -- TEST INFRASTRUCTURE ONLY.

ctypes front-end of oracle/_ref/libref_gauhuman.so: the UNMODIFIED reference
CUDA rasteriser + simple-knn, compiled by oracle/build_ref.sh from
/root/reference (in the dev container; the prebuilt .so travels to the GPU
box).  Used by the -m gpu parity tests (bit-exact integer state), by
oracle/gen_golden_raster.py, and by bench.py's "reference CUDA build on the
same box" arm.  Takes/returns torch CUDA tensors.
"""
import ctypes as C
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "_ref", "libref_gauhuman.so")
_lib = None


def available():
    return os.path.exists(LIB_PATH)


def lib():
    global _lib
    if _lib is None:
        l = C.CDLL(LIB_PATH)
        l.ref_create.restype = C.c_void_p
        l.ref_destroy.argtypes = [C.c_void_p]
        l.ref_forward.restype = C.c_int
        l.ref_backward.restype = C.c_int
        l.ref_state_bytes.restype = C.c_size_t
        l.ref_sort_pairs.restype = C.c_size_t
        _lib = l
    return _lib


def _p(t):
    if t is None or t.numel() == 0:
        return None
    assert t.is_cuda and t.is_contiguous()
    return C.c_void_p(t.data_ptr())


class RefRasterizer:
    """One handle = one set of the reference's geom/binning/img buffers."""

    def __init__(self):
        self.h = C.c_void_p(lib().ref_create())

    def __del__(self):
        try:
            lib().ref_destroy(self.h)
        except Exception:
            pass

    def forward(self, bg, means3D, colors, opacity, scales, rotations, scale_modifier, cov3D, view, proj, tan_fovx,
                tan_fovy, H, W, sh, degree, campos, prefiltered=False, debug=False):
        P = means3D.shape[0]
        dev = means3D.device
        M = sh.shape[1] if sh is not None and sh.numel() else 0
        f = dict(dtype=torch.float32, device=dev)
        color = torch.zeros(3, H, W, **f); depth = torch.zeros(1, H, W, **f); alpha = torch.zeros(1, H, W, **f)
        radii = torch.zeros(P, dtype=torch.int32, device=dev)
        self.shape = (P, degree, M, H, W)
        n = lib().ref_forward(self.h, P, degree, M, _p(bg), W, H, _p(means3D), _p(sh), _p(colors), _p(opacity),
                              _p(scales), C.c_float(scale_modifier), _p(rotations), _p(cov3D), _p(view), _p(proj),
                              _p(campos), C.c_float(tan_fovx), C.c_float(tan_fovy), int(prefiltered), _p(color),
                              _p(depth), _p(alpha), _p(radii), int(debug))
        if n < 0:
            raise RuntimeError("reference forward failed")
        self.num_rendered = n
        return n, color, depth, alpha, radii

    def backward(self, bg, means3D, radii, colors, scales, rotations, scale_modifier, cov3D, view, proj, tan_fovx,
                 tan_fovy, dcolor, ddepth, dalpha, sh, degree, campos, alphas, debug=False):
        P, _, M, H, W = self.shape
        dev = means3D.device
        z = lambda *s: torch.zeros(*s, dtype=torch.float32, device=dev)
        g = {"dL_dmeans2D": z(P, 3), "dL_dconic": z(P, 2, 2), "dL_dopacity": z(P, 1), "dL_dcolors": z(P, 3),
             "dL_dmeans3D": z(P, 3), "dL_dcov3D": z(P, 6), "dL_dsh": z(P, M, 3), "dL_dscales": z(P, 3),
             "dL_drotations": z(P, 4)}
        r = lib().ref_backward(self.h, P, degree, M, self.num_rendered, _p(bg), W, H, _p(means3D), _p(sh), _p(colors),
                               _p(alphas), _p(scales), C.c_float(scale_modifier), _p(rotations), _p(cov3D), _p(view),
                               _p(proj), _p(campos), C.c_float(tan_fovx), C.c_float(tan_fovy), _p(radii),
                               _p(dcolor.contiguous()), _p(ddepth.contiguous()), _p(dalpha.contiguous()),
                               _p(g["dL_dmeans2D"]), _p(g["dL_dconic"]), _p(g["dL_dopacity"]), _p(g["dL_dcolors"]),
                               _p(g["dL_dmeans3D"]), _p(g["dL_dcov3D"]), _p(g["dL_dsh"]) if M else None,
                               _p(g["dL_dscales"]), _p(g["dL_drotations"]), int(debug))
        if r != 0:
            raise RuntimeError("reference backward failed")
        return g

    def state(self):
        """Internal state as host numpy arrays (names follow rasterizer_impl.h:30-65)."""
        import numpy as np
        P, _, M, H, W = self.shape
        N = self.num_rendered
        tiles = ((W + 15) // 16) * ((H + 15) // 16)
        ptrs = (C.c_void_p * 16)()
        lib().ref_sync()
        lib().ref_state_ptrs(self.h, ptrs)
        spec = [("depths", np.float32, (P,)), ("clamped", np.uint8, (P, 3)), ("internal_radii", np.int32, (P,)),
                ("means2D", np.float32, (P, 2)), ("cov3D", np.float32, (P, 6)), ("conic_opacity", np.float32, (P, 4)),
                ("rgb", np.float32, (P, 3)), ("point_offsets", np.uint32, (P,)), ("tiles_touched", np.uint32, (P,)),
                ("point_list", np.uint32, (N,)), ("point_list_unsorted", np.uint32, (N,)),
                ("point_list_keys", np.uint64, (N,)), ("point_list_keys_unsorted", np.uint64, (N,)),
                ("final_T", np.float32, (H, W)), ("n_contrib", np.uint32, (H, W)), ("ranges", np.uint32, (tiles, 2))]
        out = {}
        for i, (name, dt, shape) in enumerate(spec):
            a = np.zeros(shape, dt)
            if a.size and ptrs[i]:
                lib().ref_d2h(a.ctypes.data_as(C.c_void_p), C.c_void_p(ptrs[i]), C.c_size_t(a.nbytes))
            out[name] = a
        return out


def mark_visible(means3D, view, proj):
    P = means3D.shape[0]
    present = torch.zeros(P, dtype=torch.bool, device=means3D.device)
    lib().ref_mark_visible(P, _p(means3D), _p(view), _p(proj), _p(present))
    return present


def dist_cuda2(points):
    P = points.shape[0]
    out = torch.zeros(P, dtype=torch.float32, device=points.device)
    lib().ref_dist_cuda2(P, _p(points.contiguous()), _p(out))
    lib().ref_sync()
    return out


def cub_sort_pairs(keys_in, vals_in, bits):
    """The reference's cub::DeviceRadixSort::SortPairs call; returns (keys, vals, run) where run() re-issues it."""
    N = keys_in.numel()
    keys_out = torch.empty_like(keys_in); vals_out = torch.empty_like(vals_in)
    nbytes = C.c_size_t(0)
    l = lib()
    tb = l.ref_sort_pairs(None, C.c_size_t(0), N, bits, _p(keys_in), _p(keys_out), _p(vals_in), _p(vals_out))
    # query path returns through the by-value arg in C++; ask CUB again with a generous bound instead
    temp = torch.empty(max(int(tb), 16 * N + (1 << 20)), dtype=torch.uint8, device=keys_in.device)

    def run():
        l.ref_sort_pairs(_p(temp), C.c_size_t(temp.numel()), N, bits, _p(keys_in), _p(keys_out), _p(vals_in),
                         _p(vals_out))
    run()
    return keys_out, vals_out, run
