"""-m gpu parity of the view-batched forward (ghr_forward_views / mygauhuman_b200.baking):
every view of a batch must equal the single-view operator (`_C.rasterize_gaussians`, itself bit-exact against the
reference CUDA build) called the way baking.py:237-282 calls it -- on boolean-mask gathers of the Gaussian tensors --
and the C oracle.  Integer outputs (radii, pair counts) exact; images bit-identical to the single-view path (asserted
as ==) and <= 1e-5 abs against the oracle."""
import numpy as np
import pytest
import torch

from tests import common as Cm

pytestmark = pytest.mark.gpu
E = Cm.empty


def single_view(tc, cam, keep=None, H=32, W=32, tanfov=1.0, mode="sh_scale_rot"):
    """One reference-style call (baking.py:259-282) on the kept Gaussians."""
    from mygauhuman_b200.diff_gaussian_rasterization import _C
    k = (lambda t: t) if keep is None else (lambda t: t[keep])
    view, proj, campos = cam[0:16].view(4, 4), cam[16:32].view(4, 4), cam[32:35]
    if mode == "sh_scale_rot":
        args = (tc["bg"], k(tc["means3D"]), E(), k(tc["opacity"]), k(tc["scales"]), k(tc["rotations"]), 1.0, E(), view,
                proj, tanfov, tanfov, H, W, k(tc["sh"]), tc["degree"], campos, False, False)
    else:
        args = (tc["bg"], k(tc["means3D"]), k(tc["colors"]), k(tc["opacity"]), E(), E(), 1.0, k(tc["cov3D"]), view, proj,
                tanfov, tanfov, H, W, E(), 0, campos, False, False)
    return _C.rasterize_gaussians(*args)


def cube_case(P, seed, mode):
    case = Cm.raster_case(P, 32, 32, seed=seed, mode=mode, scale=0.02)
    tc = Cm.to_torch(case, "cuda:0")
    return case, tc


@pytest.mark.parametrize("mode", ["sh_scale_rot", "precomp"])
def test_cube_faces_with_cell_masks_equal_single_view_calls(mode):
    from mygauhuman_b200 import baking as B
    case, tc = cube_case(4000, 11, mode)
    centers, _, cell_of, _ = B.pc_to_grid(tc["means3D"], 4)
    C = centers.shape[0]
    cams = B.cube_cameras(centers)
    skip = torch.arange(C, device="cuda", dtype=torch.int32).repeat_interleave(6)
    kw = (dict(scales=tc["scales"], rotations=tc["rotations"], shs=tc["sh"], sh_degree=tc["degree"])
          if mode == "sh_scale_rot" else dict(cov3D_precomp=tc["cov3D"], colors_precomp=tc["colors"]))
    out = B.rasterize_views(cams, tc["means3D"], tc["opacity"], bg=tc["bg"], labels=cell_of.int(), skip_labels=skip,
                            want_color=True, want_depth=True, want_radii=True, **kw)
    alpha_only = B.rasterize_views(cams, tc["means3D"], tc["opacity"], labels=cell_of.int(), skip_labels=skip,
                                   **{k: v for k, v in kw.items() if k in ("scales", "rotations", "cov3D_precomp")})
    assert torch.equal(alpha_only["alpha"], out["alpha"])  # the alpha-only kernel variant (what bake_set needs)
    total = 0
    checked_nonempty = 0
    for v in range(cams.shape[0]):
        keep = cell_of != int(skip[v])
        n, color, depth, alpha, radii, *_ = single_view(tc, cams[v], keep, mode=mode)
        total += n
        assert torch.equal(out["alpha"][v], alpha), v
        assert torch.equal(out["color"][v], color), v
        assert torch.equal(out["depth"][v], depth), v
        assert torch.equal(out["radii"][v][keep], radii), v
        assert int(out["radii"][v][~keep].abs().max()) == 0 if (~keep).any() else True
        checked_nonempty += int(alpha.max() > 0)
    assert out["num_rendered"] == total
    assert checked_nonempty >= cams.shape[0] // 4   # outward-looking faces of boundary cells see nothing


def test_views_against_the_c_oracle_and_odd_image_sizes():
    """General cameras (not cube faces), image size not a multiple of the tile, no masks."""
    from mygauhuman_b200 import baking as B, synthetic as S
    W, H = 40, 24
    case = Cm.raster_case(2500, W, H, seed=3, mode="sh_scale_rot")
    cams, cases = [], []
    for i in range(5):
        cam = S.make_camera(W, H, angle=0.3 + 1.1 * i, radius=3.0, seed=i)
        c = dict(case, view=cam["world_view_transform"].numpy().copy(), proj=cam["full_proj_transform"].numpy().copy(),
                 campos=cam["camera_center"].numpy().copy(), tanfovx=cam["tanfovx"], tanfovy=cam["tanfovy"])
        cases.append(c)
        cams.append(B.pack_cameras(cam["world_view_transform"][None], cam["full_proj_transform"][None],
                                   cam["camera_center"][None]))
    tc = Cm.to_torch(case, "cuda:0")
    out = B.rasterize_views(torch.cat(cams).cuda(), tc["means3D"], tc["opacity"], scales=tc["scales"],
                            rotations=tc["rotations"], shs=tc["sh"], sh_degree=tc["degree"], image_height=H, image_width=W,
                            tanfovx=cases[0]["tanfovx"], tanfovy=cases[0]["tanfovy"], bg=tc["bg"], want_color=True,
                            want_depth=True, want_radii=True)
    total = 0
    for v, c in enumerate(cases):
        st = Cm.oracle_forward(c)
        total += st.num_rendered
        assert np.array_equal(out["radii"][v].cpu().numpy(), st.radii)
        assert np.abs(out["color"][v].cpu().numpy() - st.color).max() <= 1e-5
        assert np.abs(out["alpha"][v].cpu().numpy() - st.alpha).max() <= 1e-5
        assert np.abs(out["depth"][v].cpu().numpy() - st.depth).max() <= 1e-5
        tcv = Cm.to_torch(c, "cuda:0")
        n, color, depth, alpha, radii, *_ = Cm.ours_forward(tcv)
        assert torch.equal(out["color"][v], color) and torch.equal(out["alpha"][v], alpha)
    assert out["num_rendered"] == total


def test_batches_capacity_retry_and_empty_inputs():
    from mygauhuman_b200 import baking as B
    case, tc = cube_case(1500, 5, "sh_scale_rot")
    pos = tc["means3D"][::500][:3].contiguous()
    cams = B.cube_cameras(pos)
    kw = dict(scales=tc["scales"], rotations=tc["rotations"])
    ref = B.rasterize_views(cams, tc["means3D"], tc["opacity"], **kw)
    B._state["pairs_per_slot"] = 1e-9          # forces the first launch of every batch to overflow its capacity
    small = B.rasterize_views(cams, tc["means3D"], tc["opacity"], views_per_batch=4, **kw)  # 18 views: 4+4+4+4+2
    assert torch.equal(small["alpha"], ref["alpha"]) and small["num_rendered"] == ref["num_rendered"]
    assert ref["alpha"].max() > 0
    # no views / no Gaussians
    none = B.rasterize_views(cams[:0], tc["means3D"], tc["opacity"], **kw)
    assert none["alpha"].shape == (0, 1, 32, 32)
    z = torch.zeros(0, 3, device="cuda")
    empty = B.rasterize_views(cams[:2], z, torch.zeros(0, 1, device="cuda"), scales=z, rotations=torch.zeros(0, 4, device="cuda"),
                              want_color=True, shs=torch.zeros(0, 16, 3, device="cuda"), bg=tc["bg"])
    assert float(empty["alpha"].abs().max()) == 0.0 and float(empty["color"].abs().max()) == 0.0
    with pytest.raises(Exception):
        B.rasterize_views(cams, tc["means3D"], tc["opacity"])                      # neither scale/rot nor cov3D
    with pytest.raises(RuntimeError):
        B.rasterize_views(cams.cpu(), tc["means3D"].cpu(), tc["opacity"].cpu(), **{k: v.cpu() for k, v in kw.items()})


def test_bake_opacity_cubemaps_matches_the_reference_loop():
    """bake_set's cube-map loop (baking.py:204-284) on an SMPL-sized cloud, a sample of cells checked per face."""
    from mygauhuman_b200 import baking as B
    case, tc = cube_case(6890, 7, "sh_scale_rot")
    cubes, centers, cell_of = B.bake_opacity_cubemaps(tc["means3D"], tc["opacity"], tc["scales"], tc["rotations"], grid_res=10)
    C = centers.shape[0]
    assert cubes.shape == (C, 6, 32, 32, 1) and C > 50
    cams = B.cube_cameras(centers)
    for cell in range(0, C, max(1, C // 12)):
        keep = cell_of != cell
        for f in range(6):
            n, color, depth, alpha, radii, *_ = single_view(tc, cams[cell * 6 + f], keep)
            assert torch.equal(cubes[cell, f], alpha.permute(1, 2, 0)), (cell, f)   # opacity_map.permute(1, 2, 0), :283
    assert float(cubes.max()) > 0.5
