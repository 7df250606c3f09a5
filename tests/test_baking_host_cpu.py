"""Host logic of mygauhuman_b200.baking on the CPU: the cube-face cameras and the occupancy grid against the
statements of the reference's baking.py (restated inline, :27-32, :104-134, :196-202, :248-258)."""
import math

import torch

from mygauhuman_b200 import baking as B


def reference_face_camera(rotation, position, projection_matrix):
    c2w = rotation.clone()
    c2w[:3, 3] = position
    w2c = torch.inverse(c2w)
    T = w2c[:3, 3]
    R = w2c[:3, :3].T
    Rt = torch.zeros(4, 4)                      # getWorld2ViewTorch
    Rt[:3, :3] = R[:3, :3].T
    Rt[:3, 3] = T
    Rt[3, 3] = 1.0
    world_view_transform = Rt.transpose(0, 1)
    full_proj_transform = world_view_transform.unsqueeze(0).bmm(projection_matrix.unsqueeze(0)).squeeze(0)
    camera_center = world_view_transform.inverse()[3, :3]
    return world_view_transform, full_proj_transform, camera_center


def test_cube_cameras_follow_the_reference_construction():
    g = torch.Generator().manual_seed(0)
    pos = torch.randn(5, 3, generator=g)
    cams = B.cube_cameras(pos)
    assert cams.shape == (30, 36)
    proj = B.projection_matrix(0.01, 5.0, math.pi * 0.5, math.pi * 0.5).transpose(0, 1)
    rots = B.cube_face_rotations()
    for c in range(5):
        for f in range(6):
            wvt, full, center = reference_face_camera(rots[f], pos[c], proj)
            row = cams[c * 6 + f]
            torch.testing.assert_close(row[0:16].view(4, 4), wvt, rtol=0, atol=1e-6)
            torch.testing.assert_close(row[16:32].view(4, 4), full, rtol=1e-6, atol=1e-6)
            torch.testing.assert_close(row[32:35], center, rtol=0, atol=1e-6)
            torch.testing.assert_close(row[32:35], pos[c], rtol=0, atol=1e-6)
    # 90-degree faces: tan(fov/2) = 1 -> projection [0,0] = [1,1] = 1
    assert abs(float(proj[0, 0]) - 1.0) < 1e-6 and abs(float(proj[1, 1]) - 1.0) < 1e-6


def test_pc_to_grid_labels_every_point_with_its_cell():
    g = torch.Generator().manual_seed(1)
    pc = torch.rand(2000, 3, generator=g) * torch.tensor([0.6, 1.7, 0.3])
    centers, sizes, cell_of, uniq = B.pc_to_grid(pc, 10)
    assert cell_of.shape == (2000,) and int(cell_of.max()) == centers.shape[0] - 1
    assert torch.all((pc - centers[cell_of]).abs() <= sizes[None] * 0.5 + 1e-5)
    assert torch.unique(uniq, dim=0).shape[0] == uniq.shape[0]
    assert torch.all((uniq >= 0) & (uniq <= 9))
