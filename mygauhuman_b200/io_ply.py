"""Gaussian checkpoints in the reference's PLY layout, packed / unpacked on the GPU.

``save_ply`` / ``load_ply`` write and read exactly the file GaussianModel.save_ply / load_ply do
(scene/gaussian_model.py:309-407: one 'vertex' element of float properties x y z nx ny nz ar ag ab roughness f_dc_*
f_rest_* opacity scale_* rot_*, binary little endian), so models interchange with the reference.  The interleaved rows
are built on the device by ``ghr_ply_pack_rows`` (one kernel + one D2H copy) instead of nine ``.cpu()`` copies, a numpy
concatenate and ``list(map(tuple, ...))``; loading uploads the rows once and scatters them with ``ghr_ply_unpack_rows``.

No CPU fallback: tensors must be CUDA fp32.
"""
import ctypes as C
import os

import numpy as np
import torch

from . import _lib


def attribute_names(M):
    """GaussianModel.construct_list_of_attributes (scene/gaussian_model.py:309-326); M = SH coefficients incl. DC."""
    l = ['x', 'y', 'z', 'nx', 'ny', 'nz', 'ar', 'ag', 'ab', 'roughness']
    l += ['f_dc_{}'.format(i) for i in range(3)]
    l += ['f_rest_{}'.format(i) for i in range(3 * (M - 1))]
    l.append('opacity')
    l += ['scale_{}'.format(i) for i in range(3)]
    l += ['rot_{}'.format(i) for i in range(4)]
    return l


def _dev_f32(name, t, shape):
    if t is None:
        return None
    if not t.is_cuda:
        raise RuntimeError(f"io_ply: {name} must live on a CUDA device (there is no CPU fallback)")
    if tuple(t.shape) != shape:
        raise ValueError(f"io_ply: {name} has shape {tuple(t.shape)}, expected {shape}")
    return t.detach().float().contiguous()


def pack_rows(xyz, features_dc, features_rest, opacity, scaling, rotation, normal=None, albedo=None, roughness=None):
    """-> [P, 18+3M] fp32 CUDA tensor holding the file's vertex rows.  Shapes as the reference's parameters:
    features_dc [P,1,3], features_rest [P,M-1,3], opacity [P,1], roughness [P,1]."""
    P, M = xyz.shape[0], features_rest.shape[1] + 1
    t = dict(xyz=_dev_f32("xyz", xyz, (P, 3)), normal=_dev_f32("normal", normal, (P, 3)),
             albedo=_dev_f32("albedo", albedo, (P, 3)), roughness=_dev_f32("roughness", roughness, (P, 1)),
             f_dc=_dev_f32("features_dc", features_dc, (P, 1, 3)), f_rest=_dev_f32("features_rest", features_rest, (P, M - 1, 3)),
             opacity=_dev_f32("opacity", opacity, (P, 1)), scaling=_dev_f32("scaling", scaling, (P, 3)),
             rotation=_dev_f32("rotation", rotation, (P, 4)))
    l = _lib.lib()
    cols = l.ghr_ply_row_floats(M)
    if cols == 0:
        raise ValueError("io_ply: 1..16 SH coefficients per channel")
    rows = torch.empty(P, cols, device=xyz.device)
    ptr = lambda k: None if t[k] is None or t[k].numel() == 0 else t[k].data_ptr()
    with _lib.on_device(xyz.device):
        _lib.check(l.ghr_ply_pack_rows(P, M, ptr("xyz"), ptr("normal"), ptr("albedo"), ptr("roughness"), ptr("f_dc"),
                                       ptr("f_rest"), ptr("opacity"), ptr("scaling"), ptr("rotation"),
                                       rows.data_ptr() if P else None, _lib.raw_stream(xyz.device.index)),
                   "ghr_ply_pack_rows")
    return rows


def save_ply(path, xyz, features_dc, features_rest, opacity, scaling, rotation, normal=None, albedo=None, roughness=None):
    """GaussianModel.save_ply (scene/gaussian_model.py:329-348): same bytes as plyfile writes for that element."""
    rows = pack_rows(xyz, features_dc, features_rest, opacity, scaling, rotation, normal, albedo, roughness)
    host = rows.cpu().numpy().astype('<f4', copy=False)
    names = attribute_names(features_rest.shape[1] + 1)
    header = ['ply', 'format binary_little_endian 1.0', 'element vertex {}'.format(host.shape[0])]
    header += ['property float {}'.format(n) for n in names] + ['end_header']
    d = os.path.dirname(path)
    if d:
        os.makedirs(d, exist_ok=True)  # mkdir_p(os.path.dirname(path)), :330
    with open(path, 'wb') as f:
        f.write(('\n'.join(header) + '\n').encode('ascii'))
        f.write(host.tobytes())


def _read_vertex_rows(path):
    with open(path, 'rb') as f:
        data = f.read()
    if data[:4] != b'ply\n':
        raise ValueError(f"{path}: not a PLY file")
    end = data.find(b'end_header\n')
    if end < 0:
        raise ValueError(f"{path}: PLY header has no end_header")
    end += len(b'end_header\n')
    names, P, fmt, in_vertex = [], None, None, False
    for line in data[:end].decode('ascii').split('\n'):
        t = line.split()
        if not t or t[0] in ('ply', 'comment', 'obj_info', 'end_header'):
            continue
        if t[0] == 'format':
            fmt = t[1]
        elif t[0] == 'element':
            if P is not None and not in_vertex:
                continue
            if P is not None:
                raise ValueError(f"{path}: only a single 'vertex' element is supported (what save_ply writes)")
            if t[1] != 'vertex':
                raise ValueError(f"{path}: first element is '{t[1]}', expected 'vertex'")
            P, in_vertex = int(t[2]), True
        elif t[0] == 'property':
            if t[1] not in ('float', 'float32'):
                raise ValueError(f"{path}: property {t[-1]} is {t[1]}; the checkpoint format holds float32 only")
            names.append(t[2])
    if fmt != 'binary_little_endian':
        raise ValueError(f"{path}: format {fmt}; the checkpoint format is binary_little_endian")
    if P is None:
        raise ValueError(f"{path}: no vertex element")
    need = P * len(names) * 4
    if len(data) - end < need:
        raise ValueError(f"{path}: truncated ({len(data) - end} data bytes, {need} expected)")
    rows = np.frombuffer(data, dtype='<f4', count=P * len(names), offset=end).reshape(P, len(names))
    return rows, names


def load_ply(path, device="cuda", max_sh_degree=3):
    """GaussianModel.load_ply (scene/gaussian_model.py:355-407) -> dict of fp32 CUDA tensors shaped like the reference's
    parameters: xyz [P,3], features_dc [P,1,3], features_rest [P,M-1,3], opacity [P,1], scaling [P,3], rotation [P,4],
    normal [P,3], albedo [P,3], roughness [P,1].  Properties may come in any order; a missing one raises KeyError (the
    reference's plydata.elements[0][name] does), a wrong f_rest count AssertionError (:372)."""
    rows, names = _read_vertex_rows(path)
    M = (max_sh_degree + 1) ** 2
    canon = attribute_names(M)
    extra = [n for n in names if n.startswith("f_rest_")]
    assert len(extra) == 3 * (max_sh_degree + 1) ** 2 - 3
    where = {n: i for i, n in enumerate(names)}
    columns = (C.c_int32 * len(canon))()
    for c, n in enumerate(canon):
        if n not in where:
            raise KeyError(n)
        columns[c] = where[n]
    dev = torch.device(device)
    if dev.type == "cuda" and dev.index is None:  # "cuda" = the current device
        dev = torch.device("cuda", torch.cuda.current_device())
    if dev.type != "cuda":
        raise RuntimeError("io_ply: load_ply unpacks on a CUDA device (there is no CPU fallback)")
    P = rows.shape[0]
    d_rows = torch.from_numpy(np.array(rows, dtype=np.float32)).to(dev)  # (copy: frombuffer views are read-only)
    z = lambda *s: torch.empty(*s, device=dev)
    out = dict(xyz=z(P, 3), normal=z(P, 3), albedo=z(P, 3), roughness=z(P, 1), features_dc=z(P, 1, 3),
               features_rest=z(P, M - 1, 3), opacity=z(P, 1), scaling=z(P, 3), rotation=z(P, 4))
    with _lib.on_device(dev):
        _lib.check(_lib.lib().ghr_ply_unpack_rows(
            P, M, d_rows.data_ptr() if P else None, len(names), columns, out["xyz"].data_ptr(), out["normal"].data_ptr(),
            out["albedo"].data_ptr(), out["roughness"].data_ptr(), out["features_dc"].data_ptr(),
            out["features_rest"].data_ptr(), out["opacity"].data_ptr(), out["scaling"].data_ptr(),
            out["rotation"].data_ptr(), _lib.raw_stream(dev.index or 0)), "ghr_ply_unpack_rows")
    return out
