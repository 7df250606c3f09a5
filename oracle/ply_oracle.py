"""ORACLE -- TEST INFRASTRUCTURE ONLY (never imported by the product path).

CPU restatement (numpy) of the reference's Gaussian checkpoint format:

* ``attribute_names`` / ``pack_rows`` / ``unpack_rows``: GaussianModel.construct_list_of_attributes, save_ply and
  load_ply (/root/reference/scene/gaussian_model.py:309-326, 329-348, 355-407) -- which tensor element lands in which
  float of a vertex row (f_dc / f_rest are written channel-major: ``.transpose(1, 2).flatten(start_dim=1)``).
* ``write_ply`` / ``read_ply``: the container the reference gets from ``plyfile`` (third-party, NOT vendored under
  /root/reference and not installed in this image; the reference pins no version).  Its published format for
  ``PlyData([PlyElement.describe(structured_f4_array, 'vertex')]).write(path)`` is restated: an ASCII header
  ``ply / format binary_little_endian 1.0 / element vertex N / property float <name> ... / end_header`` followed by
  the N rows as packed little-endian float32.

PINNED (data layout): tests/golden/ply_p37.{ply,npz} were produced by the reference's own save_ply / load_ply
statements (oracle/gen_golden_ply.py imports the reference module; only the plyfile container calls are stubbed with
the functions below).  UNPINNED (container header): plyfile itself is absent, so the header text rests on its
published format, not on bytes it produced here.
"""
import numpy as np


def attribute_names(M):
    """construct_list_of_attributes (:309-326) for M SH coefficients per channel (incl. DC)."""
    l = ['x', 'y', 'z', 'nx', 'ny', 'nz', 'ar', 'ag', 'ab', 'roughness']
    l += [f'f_dc_{i}' for i in range(3)]
    l += [f'f_rest_{i}' for i in range(3 * (M - 1))]
    l.append('opacity')
    l += [f'scale_{i}' for i in range(3)]
    l += [f'rot_{i}' for i in range(4)]
    return l


def pack_rows(xyz, normal, albedo, roughness, f_dc, f_rest, opacity, scaling, rotation):
    """save_ply (:329-345): [P, 18+3M] float32 rows.  f_dc [P,1,3], f_rest [P,M-1,3]."""
    P = xyz.shape[0]
    dc = np.transpose(f_dc, (0, 2, 1)).reshape(P, -1)
    rest = np.transpose(f_rest, (0, 2, 1)).reshape(P, -1)
    return np.concatenate((xyz, normal, albedo, roughness, dc, rest, opacity, scaling, rotation), axis=1).astype(np.float32)


def unpack_rows(rows, names, max_sh_degree=3):
    """load_ply (:355-407) on a [P, len(names)] float array with the file's property names."""
    col = {n: rows[:, i] for i, n in enumerate(names)}
    P = rows.shape[0]
    M = (max_sh_degree + 1) ** 2
    xyz = np.stack((col['x'], col['y'], col['z']), axis=1)
    f_dc = np.stack((col['f_dc_0'], col['f_dc_1'], col['f_dc_2']), axis=1)[:, :, None]        # [P,3,1]
    extra = sorted((n for n in names if n.startswith('f_rest_')), key=lambda x: int(x.split('_')[-1]))
    assert len(extra) == 3 * M - 3
    f_rest = np.stack([col[n] for n in extra], axis=1).reshape(P, 3, M - 1)                  # [P,3,M-1]
    scales = np.stack([col[n] for n in sorted((n for n in names if n.startswith('scale_')), key=lambda x: int(x.split('_')[-1]))], 1)
    rots = np.stack([col[n] for n in sorted((n for n in names if n.startswith('rot')), key=lambda x: int(x.split('_')[-1]))], 1)
    f32 = lambda a: np.ascontiguousarray(a, dtype=np.float32)
    return dict(xyz=f32(xyz), features_dc=f32(np.transpose(f_dc, (0, 2, 1))), features_rest=f32(np.transpose(f_rest, (0, 2, 1))),
                opacity=f32(col['opacity'][:, None]), scaling=f32(scales), rotation=f32(rots),
                normal=f32(np.stack((col['nx'], col['ny'], col['nz']), 1)),
                albedo=f32(np.stack((col['ar'], col['ag'], col['ab']), 1)), roughness=f32(col['roughness'][:, None]))


def header_bytes(names, P):
    lines = ['ply', 'format binary_little_endian 1.0', f'element vertex {P}']
    lines += [f'property float {n}' for n in names]
    lines.append('end_header')
    return ('\n'.join(lines) + '\n').encode('ascii')


def write_ply(path, rows, names):
    rows = np.ascontiguousarray(rows, dtype='<f4')
    assert rows.shape[1] == len(names)
    with open(path, 'wb') as f:
        f.write(header_bytes(names, rows.shape[0]))
        f.write(rows.tobytes())


def read_ply(path):
    """-> (rows [P, n] float32, names).  binary_little_endian float properties of one 'vertex' element only."""
    with open(path, 'rb') as f:
        data = f.read()
    end = data.index(b'end_header\n') + len(b'end_header\n')
    names, P = [], None
    for line in data[:end].decode('ascii').split('\n'):
        t = line.split()
        if t[:1] == ['format']:
            assert t[1] == 'binary_little_endian'
        elif t[:2] == ['element', 'vertex']:
            P = int(t[2])
        elif t[:1] == ['property']:
            assert t[1] in ('float', 'float32')
            names.append(t[2])
    rows = np.frombuffer(data, dtype='<f4', count=P * len(names), offset=end).reshape(P, len(names))
    return rows.astype(np.float32), names
