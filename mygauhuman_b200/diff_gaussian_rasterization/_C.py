"""Drop-in for the reference's compiled extension ``diff_gaussian_rasterization._C``
(/root/reference/submodules/diff-gaussian-rasterization/ext.cpp:15-19,
rasterize_points.cu:36-228): same three functions, same positional arguments,
same return tuples -- implemented on the C ABI of libghr.so through ctypes.

Differences that are deliberate and invisible to callers:
* kernels run on torch's *current* stream (the reference uses the legacy
  default stream);
* the three opaque uint8 state tensors use libghr's layout (see
  ``state_views``), not the reference's ``fromChunk`` layout;
* ``prefiltered=True`` with a culled point raises instead of ``__trap()``-ing
  the context (only checked when ``debug`` is set, as it needs a sync).
"""
import ctypes as C

import torch

from .. import _lib
from .._lib import RasterParams, check, lib, state_layout

# Default mode ("optimistic launch"): the binning buffer of a call is sized from the previous call with the same
# (P, W, H) plus 25 % headroom and the whole forward is enqueued at once; the host then waits only for the FIRST kernel
# (an event right after preprocess), reads the exact num_rendered from pinned memory while the GPU is already sorting
# and blending, and re-runs the render stage with an exact-size buffer in the rare case the guess was too small.
# Results are always exact; unlike the reference's blocking read (rasterizer_impl.cu:283) the GPU never idles.
_cap_hint = {}      # (device index, P, W, H) -> capacity guess


def _round_cap(n):
    return (int(n) + 63) // 64 * 64


def _capacity_of(binning, P, W, H):
    """The capacity a binning buffer was laid out for, recovered from its size (ghr_binning_capacity: the size is
    strictly increasing in the capacity).  The reference's backward gets num_rendered only; here the buffer may be
    larger than that (optimistic launch, sync-free mode) and its array offsets depend on the capacity, so nothing but
    the saved tensor itself is trusted: it survives any number of outstanding forwards, hooks that move saved tensors
    (save_on_cpu) and allocator address reuse."""
    cap = int(lib().ghr_binning_capacity(int(P), int(W), int(H), binning.numel()))
    if cap < 0:
        raise _lib.GhrError(_lib.GHR_ERR_BUFFER, "binningBuffer does not have the size of any binning layout")
    return cap


# Sync-free mode: when set to an int, forward() does not read num_rendered back
# to the host; the binning buffer is sized for that many (Gaussian, tile) pairs
# and the returned "num_rendered" is that capacity.  Overflow is latched on the
# device (see ``read_status``).  None = the reference's exact-size behaviour
# (one 4-byte host read after the first kernel).
_capacity = None
_pinned = {}
_last_geom = None   # geometry state of the most recent forward (graphed.py reads num_rendered from it on the device)


def set_sync_free_capacity(capacity):
    global _capacity
    _capacity = None if capacity is None else int(capacity)


def _ptr(t, align=4):
    if t is None or t.numel() == 0:
        return None, t
    if not t.is_cuda:
        raise RuntimeError("libghr: tensor arguments must live on a CUDA device")
    if t.dtype != torch.float32:
        t = t.float()
    t = t.contiguous()
    if t.data_ptr() % align:
        t = t.clone()
    return t.data_ptr(), t


def _stream(dev=None):
    return _lib.raw_stream(torch._C._cuda_getDevice() if dev is None or dev.index is None else dev.index)


def _params(keep, background, means3D, colors, opacity, scales, rotations, scale_modifier, cov3D_precomp, viewmatrix,
            projmatrix, tan_fovx, tan_fovy, H, W, sh, degree, campos, prefiltered, debug):
    p = RasterParams()
    p.P = means3D.size(0)
    p.D = int(degree)
    p.M = sh.size(1) if sh.numel() != 0 else 0
    p.W, p.H = int(W), int(H)
    p.scale_modifier = float(scale_modifier)
    p.tan_fovx, p.tan_fovy = float(tan_fovx), float(tan_fovy)
    p.prefiltered, p.debug = int(bool(prefiltered)), int(bool(debug))
    for name, t, al in (("background", background, 4), ("means3D", means3D, 4), ("shs", sh, 16),
                        ("colors_precomp", colors, 4), ("opacities", opacity, 4), ("scales", scales, 4),
                        ("rotations", rotations, 16), ("cov3D_precomp", cov3D_precomp, 8),
                        ("viewmatrix", viewmatrix, 4), ("projmatrix", projmatrix, 4), ("campos", campos, 4)):
        ptr, t = _ptr(t, al)
        keep.append(t)
        setattr(p, name, ptr)
    return p


def _pinned_counter(device):
    buf = _pinned.get(device)
    if buf is None:
        buf = torch.zeros(1, dtype=torch.int32).pin_memory()
        _pinned[device] = buf
    return buf


def launch_with_capacity(dev, P, W, H, preprocess, render):
    """The binning-capacity policy of a forward: preprocess(counter_ptr | None) enqueues the projection kernel (which
    writes num_rendered to the pinned counter when given one), render(capacity) -> binning buffer enqueues the rest.
    Sync-free mode: fixed capacity, no host read.  Otherwise: optimistic launch (see the comment at _cap_hint).
    Returns (num_rendered, binning)."""
    if _capacity is not None:  # explicit sync-free mode
        preprocess(None)
        capacity = num_rendered = _round_cap(_capacity)
        return num_rendered, render(capacity)
    counter = _pinned_counter(dev)
    preprocess(counter.data_ptr())
    key = (dev.index, P, W, H)
    guess = _cap_hint.get(key)
    if guess is None:
        torch.cuda.current_stream().synchronize()  # first call: the reference's blocking read
        num_rendered = int(counter[0])
        capacity = _round_cap(num_rendered)
        binning = render(capacity)
    else:
        ev = torch.cuda.Event()
        ev.record()
        capacity = guess
        binning = render(capacity)       # GPU keeps working while the host waits for preprocess only
        ev.synchronize()
        num_rendered = int(counter[0])
        if num_rendered > capacity:      # rare: guess too small -> kernels exited early; redo exactly
            capacity = _round_cap(num_rendered)
            binning = render(capacity)
    _cap_hint[key] = _round_cap(num_rendered + num_rendered // 4 + 1024)
    return num_rendered, binning


def rasterize_gaussians(background, means3D, colors, opacity, scales, rotations, scale_modifier, cov3D_precomp,
                        viewmatrix, projmatrix, tan_fovx, tan_fovy, image_height, image_width, sh, degree, campos,
                        prefiltered, debug, _attrs=None):
    """RasterizeGaussiansCUDA, rasterize_points.cu:36-120.  (`_attrs`: rasterize_gaussians_multi's packed attributes.)"""
    if means3D.dim() != 2 or means3D.size(1) != 3:
        raise RuntimeError("means3D must have dimensions (num_points, 3)")
    P, H, W = means3D.size(0), int(image_height), int(image_width)
    dev = means3D.device
    if not means3D.is_cuda:
        raise RuntimeError("libghr: tensor arguments must live on a CUDA device (there is no CPU fallback)")
    f32 = dict(dtype=torch.float32, device=dev)
    if P == 0:
        u8 = dict(dtype=torch.uint8, device=dev)
        return (0, torch.zeros(3, H, W, **f32), torch.zeros(1, H, W, **f32), torch.zeros(1, H, W, **f32),
                torch.zeros(0, dtype=torch.int32, device=dev), torch.empty(0, **u8), torch.empty(0, **u8),
                torch.empty(0, **u8))
    with _lib.on_device(dev):
        keep = []
        p = _params(keep, background, means3D, colors, opacity, scales, rotations, scale_modifier, cov3D_precomp,
                    viewmatrix, projmatrix, tan_fovx, tan_fovy, H, W, sh, degree, campos, prefiltered, debug)
        L0 = state_layout(P, 0, W, H)
        geom = torch.empty(L0.geom_bytes, dtype=torch.uint8, device=dev)
        img = torch.empty(L0.img_bytes, dtype=torch.uint8, device=dev)
        radii = torch.empty(P, dtype=torch.int32, device=dev)
        out_color = torch.empty(3, H, W, **f32)
        out_depth = torch.empty(1, H, W, **f32)
        out_alpha = torch.empty(1, H, W, **f32)
        l = lib()
        s = _stream(dev)

        def render(capacity):
            L = state_layout(P, capacity, W, H)
            binning = torch.empty(L.binning_bytes, dtype=torch.uint8, device=dev)
            if _attrs is not None:
                _attrs.fused = bool(l.ghr_attributes_fused(capacity))
                if _attrs.fused:
                    check(l.ghr_forward_render_multi(C.byref(p), geom.data_ptr(), geom.numel(), binning.data_ptr(),
                                                     binning.numel(), capacity, img.data_ptr(), img.numel(),
                                                     out_color.data_ptr(), out_depth.data_ptr(), out_alpha.data_ptr(),
                                                     _attrs.n, _attrs.packed.data_ptr(), _attrs.images.data_ptr(), s),
                          "ghr_forward_render_multi")
                    return binning
            check(l.ghr_forward_render(C.byref(p), geom.data_ptr(), geom.numel(), binning.data_ptr(), binning.numel(),
                                       capacity, img.data_ptr(), img.numel(), out_color.data_ptr(), out_depth.data_ptr(),
                                       out_alpha.data_ptr(), s), "ghr_forward_render")
            return binning

        def preprocess(counter_ptr):
            check(l.ghr_forward_preprocess(C.byref(p), radii.data_ptr(), geom.data_ptr(), geom.numel(), counter_ptr, s),
                  "ghr_forward_preprocess")

        num_rendered, binning = launch_with_capacity(dev, P, W, H, preprocess, render)
        if debug:
            read_status(geom, raise_on_error=True)
    global _last_geom
    _last_geom = geom
    return num_rendered, out_color, out_depth, out_alpha, radii, geom, binning, img


def rasterize_gaussians_backward(background, means3D, radii, colors, scales, rotations, scale_modifier, cov3D_precomp,
                                 viewmatrix, projmatrix, tan_fovx, tan_fovy, dL_dout_color, dL_dout_depth,
                                 dL_dout_alpha, sh, degree, campos, geomBuffer, R, binningBuffer, imageBuffer, alphas,
                                 debug):
    """RasterizeGaussiansBackwardCUDA, rasterize_points.cu:122-207."""
    P = means3D.size(0)
    H, W = dL_dout_color.size(1), dL_dout_color.size(2)
    M = sh.size(1) if sh.numel() != 0 else 0
    dev = means3D.device
    if P == 0:
        z = lambda *s: torch.zeros(*s, dtype=torch.float32, device=dev)
        return z(0, 3), z(0, 3), z(0, 1), z(0, 3), z(0, 6), z(0, M, 3), z(0, 3), z(0, 4)
    capacity = _capacity_of(binningBuffer, P, W, H)
    with _lib.on_device(dev):
        keep = []
        # opacities are not an input of the reference backward (they live in the saved geometry state)
        p = _params(keep, background, means3D, colors, means3D, scales, rotations, scale_modifier, cov3D_precomp,
                    viewmatrix, projmatrix, tan_fovx, tan_fovy, H, W, sh, degree, campos, False, debug)
        # one allocation for the accumulator and every gradient; only the accumulator is zeroed
        sizes = [12, 4, 3, 3, 1, 3, 6, 3, 3 * M]  # accum, drot, dmeans2D, dcolors, dopacity, dmeans3D, dcov3D, dscales, dsh
        flat = torch.empty(P * sum(sizes), dtype=torch.float32, device=dev)
        views, o = [], 0
        for n in sizes:
            views.append(flat[o:o + P * n])
            o += P * n
        accum, drot, dmeans2D, dcolors, dopacity, dmeans3D, dcov3D, dscales, dsh = views
        accum.zero_()
        g = [t.contiguous() if t.dtype == torch.float32 else t.float().contiguous()
             for t in (dL_dout_color, dL_dout_depth, dL_dout_alpha)]
        radii = radii.contiguous()
        check(lib().ghr_backward(C.byref(p), radii.data_ptr(), geomBuffer.data_ptr(), geomBuffer.numel(),
                                 binningBuffer.data_ptr(), binningBuffer.numel(), capacity, imageBuffer.data_ptr(),
                                 imageBuffer.numel(), g[0].data_ptr(), g[1].data_ptr(), g[2].data_ptr(),
                                 accum.data_ptr(), dmeans2D.data_ptr(), dcolors.data_ptr(), dopacity.data_ptr(),
                                 dmeans3D.data_ptr(), dcov3D.data_ptr(), dsh.data_ptr() if M else None,
                                 dscales.data_ptr(), drot.data_ptr(), _stream(dev)), "ghr_backward")
    return (dmeans2D.view(P, 3), dcolors.view(P, 3), dopacity.view(P, 1), dmeans3D.view(P, 3), dcov3D.view(P, 6),
            dsh.view(P, M, 3), dscales.view(P, 3), drot.view(P, 4))


# ---- multi-attribute rendering: one projection / sort / tile binning for all attributes --------------------------
# (reference gaussian_renderer/__init__.py:203-272 runs the whole rasteriser once per attribute; see include/ghr.h)

_FUSED_MAX = 6  # attributes blended inside the main kernel (in pairs); any further ones are re-blended one by one


class _Attrs:
    def __init__(self, n, packed, images):
        self.n, self.packed, self.images, self.fused = n, packed, images, False


def _rec_colors(P, W, H, geom):
    L = state_layout(P, 0, W, H)
    return geom[L.g_rec:L.g_rec + P * 48].view(torch.float32).view(P, 3, 4)[:, 2, 0:3]


def _ptr_array(tensors):
    return (C.c_void_p * len(tensors))(*[None if t is None else t.data_ptr() for t in tensors])


def rasterize_gaussians_multi(background, means3D, colors, opacity, scales, rotations, scale_modifier, cov3D_precomp,
                              viewmatrix, projmatrix, tan_fovx, tan_fovy, image_height, image_width, sh, degree,
                              campos, prefiltered, debug, extra_colors):
    """rasterize_gaussians for the first attribute (SH or `colors`) plus one image per [P,3] tensor of `extra_colors`,
    all over the same sorted lists: up to six of them inside the same blend kernel, the rest re-blended.  Returns the
    tuple of rasterize_gaussians plus (extra images, colours of the first attribute or None, packed attributes or None)."""
    P, H, W = means3D.size(0), int(image_height), int(image_width)
    dev = means3D.device
    extra = []
    for c in extra_colors:
        if c.dim() != 2 or c.size(0) != P or c.size(1) != 3:
            raise RuntimeError("every extra attribute must have dimensions (num_points, 3)")
        extra.append(_ptr(c)[1] if P else c)
    nf = min(_FUSED_MAX, len(extra) - len(extra) % 2) if P else 0
    att = None
    if nf:
        with _lib.on_device(dev):
            l = lib()
            packed = torch.empty(P, l.ghr_attribute_stride(nf), dtype=torch.float32, device=dev)
            check(l.ghr_pack_attributes(P, nf, _ptr_array(extra[:nf]), packed.data_ptr(), _stream(dev)),
                  "ghr_pack_attributes")
            att = _Attrs(nf, packed, torch.empty(3 * nf, H, W, dtype=torch.float32, device=dev))
    out = rasterize_gaussians(background, means3D, colors, opacity, scales, rotations, scale_modifier, cov3D_precomp,
                              viewmatrix, projmatrix, tan_fovx, tan_fovy, image_height, image_width, sh, degree,
                              campos, prefiltered, debug, _attrs=att)
    num_rendered, out_color, out_depth, out_alpha, radii, geom, binning, img = out
    if P == 0:
        return out + ([torch.zeros(3, H, W, dtype=torch.float32, device=dev) for _ in extra], None, None)
    fused = att is not None and att.fused
    images = [att.images[3 * k:3 * k + 3] for k in range(nf)] if fused else []
    rest = extra[nf:] if fused else extra
    first_colors = _rec_colors(P, W, H, geom).clone() if rest else None
    capacity = _capacity_of(binning, P, W, H)
    if rest:
        with _lib.on_device(dev):
            keep = []
            p = _params(keep, background, means3D, colors, opacity, scales, rotations, scale_modifier, cov3D_precomp,
                        viewmatrix, projmatrix, tan_fovx, tan_fovy, H, W, sh, degree, campos, prefiltered, debug)
            scratch = torch.empty(2, H, W, dtype=torch.float32, device=dev)  # depth / alpha come out the same every time
            l, s = lib(), _stream(dev)
            for c in rest:
                image = torch.empty(3, H, W, dtype=torch.float32, device=dev)
                check(l.ghr_forward_reblend(C.byref(p), c.data_ptr(), geom.data_ptr(), geom.numel(), binning.data_ptr(),
                                            binning.numel(), capacity, img.data_ptr(), img.numel(), image.data_ptr(),
                                            scratch[0].data_ptr(), scratch[1].data_ptr(), s), "ghr_forward_reblend")
                images.append(image)
            _rec_colors(P, W, H, geom).copy_(first_colors)  # the records again hold the first attribute
    return out + (images, first_colors, att.packed if fused else None)


def rasterize_gaussians_multi_backward(background, means3D, radii, colors, scales, rotations, scale_modifier,
                                       cov3D_precomp, viewmatrix, projmatrix, tan_fovx, tan_fovy, dL_dout_color,
                                       dL_dout_depth, dL_dout_alpha, sh, degree, campos, geomBuffer, R, binningBuffer,
                                       imageBuffer, alphas, debug, first_colors, packed, extra_colors, dL_dextra):
    """Backward of rasterize_gaussians_multi.  `dL_dextra[k]` may be None (attribute k did not reach the loss).
    Returns the tuple of rasterize_gaussians_backward plus the list of dL_dextra_colors."""
    P = means3D.size(0)
    H, W = dL_dout_color.size(1), dL_dout_color.size(2)
    M = sh.size(1) if sh.numel() != 0 else 0
    dev = means3D.device
    if P == 0:
        z = lambda *s: torch.zeros(*s, dtype=torch.float32, device=dev)
        return (z(0, 3), z(0, 3), z(0, 1), z(0, 3), z(0, 6), z(0, M, 3), z(0, 3), z(0, 4), [z(0, 3) for _ in extra_colors])
    capacity = _capacity_of(binningBuffer, P, W, H)
    nf = 0 if packed is None else min(_FUSED_MAX, len(extra_colors) - len(extra_colors) % 2)
    with _lib.on_device(dev):
        keep = []
        p = _params(keep, background, means3D, colors, means3D, scales, rotations, scale_modifier, cov3D_precomp,
                    viewmatrix, projmatrix, tan_fovx, tan_fovy, H, W, sh, degree, campos, False, debug)
        sizes = [12, 4, 3, 3, 1, 3, 6, 3, 3 * M]
        flat = torch.empty(P * sum(sizes), dtype=torch.float32, device=dev)
        views, o = [], 0
        for n in sizes:
            views.append(flat[o:o + P * n])
            o += P * n
        accum, drot, dmeans2D, dcolors, dopacity, dmeans3D, dcov3D, dscales, dsh = views
        accum.zero_()
        accum2 = accum.view(P, 12)
        f32c = lambda t: t.contiguous() if t.dtype == torch.float32 else t.float().contiguous()
        radii = radii.contiguous()
        l, s = lib(), _stream(dev)
        geo = (geomBuffer.data_ptr(), geomBuffer.numel(), binningBuffer.data_ptr(), binningBuffer.numel(), capacity,
               imageBuffer.data_ptr(), imageBuffer.numel())
        dextra = [None] * len(extra_colors)
        zeros_plane = None
        # attributes that were re-blended: one blend backward each into the shared accumulator
        for k in range(nf, len(extra_colors)):
            if dL_dextra[k] is None:
                continue
            if zeros_plane is None:
                zeros_plane = torch.zeros(H * W, dtype=torch.float32, device=dev)
            cp, c = _ptr(extra_colors[k])
            g = f32c(dL_dextra[k])
            check(l.ghr_backward_blend(C.byref(p), cp, *geo, g.data_ptr(), None, None, zeros_plane.data_ptr(),
                                       accum.data_ptr(), s), "ghr_backward_blend")
            dextra[k] = accum2[:, 0:3].clone()
            accum2[:, 0:3].zero_()
        g = [f32c(t) for t in (dL_dout_color, dL_dout_depth, dL_dout_alpha)]
        if zeros_plane is not None:
            _rec_colors(P, W, H, geomBuffer).copy_(first_colors)  # ghr_backward_blend left another attribute there
        grads = (dmeans2D.data_ptr(), dcolors.data_ptr(), dopacity.data_ptr(), dmeans3D.data_ptr(), dcov3D.data_ptr(),
                 dsh.data_ptr() if M else None, dscales.data_ptr(), drot.data_ptr())
        if nf and any(dL_dextra[k] is not None for k in range(nf)):
            planes = [f32c(dL_dextra[k]) if dL_dextra[k] is not None
                      else torch.zeros(3, H, W, dtype=torch.float32, device=dev) for k in range(nf)]
            xdpix = torch.cat(planes)
            packed_grad = torch.empty_like(packed)
            check(l.ghr_backward_multi(C.byref(p), radii.data_ptr(), *geo, g[0].data_ptr(), g[1].data_ptr(),
                                       g[2].data_ptr(), nf, packed.data_ptr(), xdpix.data_ptr(), packed_grad.data_ptr(),
                                       accum.data_ptr(), *grads, s), "ghr_backward_multi")
            outs = [torch.empty(P, 3, dtype=torch.float32, device=dev) if dL_dextra[k] is not None else None
                    for k in range(nf)]
            check(l.ghr_unpack_attribute_grads(P, nf, packed_grad.data_ptr(), _ptr_array(outs), s),
                  "ghr_unpack_attribute_grads")
            dextra[:nf] = outs
        else:
            check(l.ghr_backward(C.byref(p), radii.data_ptr(), *geo, g[0].data_ptr(), g[1].data_ptr(), g[2].data_ptr(),
                                 accum.data_ptr(), *grads, s), "ghr_backward")
    return (dmeans2D.view(P, 3), dcolors.view(P, 3), dopacity.view(P, 1), dmeans3D.view(P, 3), dcov3D.view(P, 6),
            dsh.view(P, M, 3), dscales.view(P, 3), drot.view(P, 4), dextra)


def mark_visible(means3D, viewmatrix, projmatrix):
    """markVisible, rasterize_points.cu:209-228."""
    P = means3D.size(0)
    present = torch.zeros(P, dtype=torch.bool, device=means3D.device)
    if P:
        with _lib.on_device(means3D.device):
            keep = []
            ptrs = []
            for t in (means3D, viewmatrix, projmatrix):
                ptr, t = _ptr(t)
                keep.append(t)
                ptrs.append(ptr)
            check(lib().ghr_mark_visible(P, ptrs[0], ptrs[1], ptrs[2], present.data_ptr(), _stream(means3D.device)),
                  "ghr_mark_visible")
    return present


# ---- inspection helpers (parity tests, debugging) ---------------------------

def read_status(geomBuffer, raise_on_error=False):
    """(num_rendered, error) from the geometry header; synchronises."""
    hdr = geomBuffer[:8].view(torch.int32).cpu()
    n, err = int(hdr[0]), int(hdr[1])
    if raise_on_error and err:
        raise _lib.GhrError(-err, "device status")
    return n, err


def state_views(P, capacity, W, H, geomBuffer, binningBuffer, imageBuffer):
    """Typed views of the internal state the reference keeps in its three buffers.
    `capacity` is the num_rendered the forward returned; the capacity the binning buffer was actually laid
    out for (>= num_rendered) is looked up."""
    capacity = _capacity_of(binningBuffer, P, W, H) if binningBuffer.numel() else capacity
    L = state_layout(P, capacity, W, H)
    X = W * H
    tiles = ((W + 15) // 16) * ((H + 15) // 16)
    g, b, im = geomBuffer, binningBuffer, imageBuffer

    def view(buf, off, count, dtype, itemsize):
        return buf[off:off + count * itemsize].view(dtype)

    rec = view(g, L.g_rec, P * 12, torch.float32, 4).view(P, 3, 4)
    out = {
        "means2D": rec[:, 0, 0:2], "depths": rec[:, 0, 2], "conic_opacity": rec[:, 1, :], "rgb": rec[:, 2, 0:3],
        "tiles_touched": view(g, L.g_tiles_touched, P, torch.int32, 4),
        "point_offsets": view(g, L.g_point_offsets, P, torch.int32, 4),
        "cov3D": view(g, L.g_cov3D, P * 6, torch.float32, 4).view(P, 6),
        "clamped": view(g, L.g_clamped, P, torch.uint8, 1),
        "final_T": view(im, L.i_final_T, X, torch.float32, 4).view(H, W),
        "n_contrib": view(im, L.i_n_contrib, X, torch.int32, 4).view(H, W),
        "ranges": view(im, L.i_ranges, tiles * 2, torch.int32, 4).view(tiles, 2),
    }
    if capacity > 0:
        hdr = view(b, L.b_header, 4, torch.int32, 4)
        sel = int(hdr[0])
        n = min(read_status(g)[0], capacity)
        ka = view(b, L.b_keys_a, capacity, torch.int64, 8)
        kb = view(b, L.b_keys_b, capacity, torch.int64, 8)
        va = view(b, L.b_vals_a, capacity, torch.int32, 4)
        vb = view(b, L.b_vals_b, capacity, torch.int32, 4)
        out["point_list_keys"] = (kb if sel else ka)[:n]
        ids = (vb if sel else va)[:n]
        id_mask = int(lib().ghr_point_list_id_mask(int(P), int(capacity)))
        if id_mask != 0xFFFFFFFF:  # large frames: the upper 8 bits of a list value are the pair's block mask
            ids = ids & id_mask
        out["point_list"] = ids
        out["sorted_sel"] = sel
        out["pass_trivial"] = view(b, L.b_header + 16, 8, torch.int32, 4)
    return out
