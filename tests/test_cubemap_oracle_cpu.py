"""CPU checks of the cube-map lookup restatement (oracle/cubemap_oracle.py): face selection and in-face orientation
against the OpenGL cube-map table, texel addressing, bake_set's occlusion statements.  (nvdiffrast itself is absent:
parity unpinned, see the oracle's header.)"""
import numpy as np

from oracle import cubemap_oracle as CO


def test_faces_follow_the_opengl_table():
    # major axis -> face; (sc, tc) per face: +x (-z,-y)  -x (+z,-y)  +y (+x,+z)  -y (+x,-z)  +z (+x,-y)  -z (-x,-y)
    axes = np.eye(3, dtype=np.float32)
    for k, sign, face in ((0, 1, 0), (0, -1, 1), (1, 1, 2), (1, -1, 3), (2, 1, 4), (2, -1, 5)):
        f, s, t = CO.cube_face_uv(sign * axes[k][None])
        assert f[0] == face and abs(s[0] - 0.5) < 1e-6 and abs(t[0] - 0.5) < 1e-6
    table = {0: ((0, 0, -1), (0, -1, 0)), 1: ((0, 0, 1), (0, -1, 0)), 2: ((1, 0, 0), (0, 0, 1)), 3: ((1, 0, 0), (0, 0, -1)),
             4: ((1, 0, 0), (0, -1, 0)), 5: ((-1, 0, 0), (0, -1, 0))}
    major = {0: (1, 0, 0), 1: (-1, 0, 0), 2: (0, 1, 0), 3: (0, -1, 0), 4: (0, 0, 1), 5: (0, 0, -1)}
    for face, (sc, tc) in table.items():
        d = np.asarray(major[face], np.float32) + 0.3 * np.asarray(sc, np.float32) + 0.2 * np.asarray(tc, np.float32)
        f, s, t = CO.cube_face_uv(d[None])
        assert f[0] == face
        assert abs(s[0] - (0.5 + 0.15)) < 1e-6 and abs(t[0] - (0.5 + 0.1)) < 1e-6, (face, s, t)


def test_texel_addressing_and_degenerate_directions():
    res = 4
    tex = np.arange(6 * res * res, dtype=np.float32).reshape(1, 6, res, res, 1)
    d = np.asarray([[1.0, 0.99, 0.99], [1.0, -0.99, -0.99], [0.0, 0.0, 0.0], [np.nan, 1.0, 0.0]], np.float32)
    out = CO.texture_nearest_cube(tex, d)[0, :, 0]
    # +x face: u = -z, v = -y: (0.99,0.99) -> s,t ~ 0.005 -> texel (0,0); (-0.99,-0.99) -> texel (3,3)
    assert out[0] == tex[0, 0, 0, 0, 0] and out[1] == tex[0, 0, 3, 3, 0]
    assert out[2] == 0.0 and out[3] == 0.0


def test_occlusion_statements():
    rng = np.random.default_rng(0)
    C, P, res = 3, 20, 8
    cubes = rng.uniform(0, 1, size=(C, 6, res, res, 1)).astype(np.float32)
    cell_of = rng.integers(0, C, size=P)
    normals = rng.normal(size=(P, 3)).astype(np.float32)
    dirs = CO.envmap_dirs((16, 32))
    occ = CO.occlusion(cubes, cell_of, normals, dirs)
    assert occ.shape == (P, 16, 32, 1)
    back = (dirs[None] * normals[:, None, None, :]).sum(-1) <= 0
    assert np.all(occ[back] == 0.0) and occ.min() >= 0.0 and occ.max() <= 1.0
    assert np.abs(np.linalg.norm(dirs, axis=-1) - 1.0).max() < 1e-5
