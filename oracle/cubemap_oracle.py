"""ORACLE -- TEST INFRASTRUCTURE ONLY.  PARITY UNPINNED.

numpy restatement of the cube-map lookup bake_set asks of nvdiffrast,
    dr.texture(tex[1,6,res,res,C], uv[1,H,W,3], filter_mode="nearest", boundary_mode="cube")   (baking.py:290-297).
nvdiffrast (NVlabs/nvdiffrast, the reference pins commit c5caf7b in its README) is an un-vendored third-party dependency
that is absent from /root/reference and from this image, and the reference holds no test or golden vector for this call,
so this restatement cannot be checked against the real thing here: it follows nvdiffrast's published texture.cu
(indexCubeMap + nearest texel) and agrees with the OpenGL 4.6 cube-map face table (8.19), on which nvdiffrast's is
modelled.  The occlusion statements of bake_set around the call (:231,305-307) are restated from the reference itself.
"""
import numpy as np


def cube_face_uv(dirs):
    """dirs [...,3] -> (face [...], u [...], v [...]) with u, v in [0,1]; face -1 where no major axis exists."""
    d = np.asarray(dirs, np.float32).reshape(-1, 3)
    x, y, z = d[:, 0], d[:, 1], d[:, 2]
    ax, ay, az = np.abs(x), np.abs(y), np.abs(z)
    face = np.zeros(d.shape[0], np.int64)
    c = x.copy(); u = z.copy(); v = y.copy()          # +-x faces: (z, y)
    selz = az > np.maximum(ax, ay)
    sely = ~selz & (ay > ax)
    face[selz] = 4; c[selz] = z[selz]; u[selz] = x[selz]; v[selz] = y[selz]
    face[sely] = 2; c[sely] = y[sely]; u[sely] = x[sely]; v[sely] = z[sely]
    face = face + (c < 0)
    with np.errstate(divide="ignore", invalid="ignore"):
        m = np.float32(0.5) / np.abs(c)
        mu = np.where((face == 0) | (face == 5), -m, m)
        mv = np.where(face != 2, -m, m)
        s = u * mu + np.float32(0.5)
        t = v * mv + np.float32(0.5)
    bad = ~np.isfinite(s) | ~np.isfinite(t)
    s = np.clip(s, 0.0, 1.0); t = np.clip(t, 0.0, 1.0)
    face = np.where(bad, -1, face)
    shape = np.asarray(dirs).shape[:-1]
    return face.reshape(shape), s.reshape(shape), t.reshape(shape)


def texture_nearest_cube(tex, dirs):
    """tex [n,6,res,res,C], dirs [...,3] -> [n,...,C]."""
    tex = np.asarray(tex, np.float32)
    res = tex.shape[2]
    face, s, t = cube_face_uv(dirs)
    iu = np.clip(np.floor(s * res).astype(np.int64), 0, res - 1)
    iv = np.clip(np.floor(t * res).astype(np.int64), 0, res - 1)
    out = tex[:, np.maximum(face, 0), iv, iu, :]
    out[:, face < 0, :] = 0.0
    return out


def envmap_dirs(res=(16, 32)):
    """baking.py:35-58 in fp32 numpy."""
    H, W = res
    gy, gx = np.meshgrid(np.linspace(0.0, 1.0, H, dtype=np.float32), np.linspace(-1.0, 1.0, W, dtype=np.float32), indexing="ij")
    st, ct = np.sin(gy * np.float32(np.pi)), np.cos(gy * np.float32(np.pi))
    sp, cp = np.sin(gx * np.float32(np.pi)), np.cos(gx * np.float32(np.pi))
    return np.stack((st * sp, ct, -st * cp), -1).astype(np.float32)


def occlusion(cubemaps, cell_of, normals, dirs):
    """bake_set :231,299-307: cubemaps [C,6,res,res,1], cell_of [P], normals [P,3], dirs [H,W,3] -> [P,H,W,1]."""
    env = texture_nearest_cube(cubemaps, dirs)                      # [C,H,W,1]
    dot_map = (dirs[None] * np.asarray(normals, np.float32)[:, None, None, :]).sum(-1, keepdims=True) > 0
    return dot_map * (1.0 - env[np.asarray(cell_of)])
