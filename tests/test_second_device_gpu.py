"""-m gpu, needs 2 GPUs: the default-device arguments (``SMPLDeformer(smpl)``: device="cuda") on a
process whose CURRENT device is not 0 (a torchrun rank).  ADVICE r1: an index-less device used to be mapped to GPU 0, so
kernels were launched on GPU 0 against GPU-1 pointers.  Everything must land on, and run on, the current device, and
give the numbers GPU 0 gives."""
import pytest
import torch

from mygauhuman_b200 import synthetic as S

pytestmark = pytest.mark.gpu


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs")
def test_default_device_arguments_follow_the_current_device(smpl_model):
    from mygauhuman_b200 import diff_gaussian_rasterization as dgr
    from mygauhuman_b200.fused import render_deformed
    from mygauhuman_b200.lbs import SMPLDeformer
    big = S.big_pose_params()
    tv = S.big_pose_vertices(smpl_model, big)
    P, W, H = 3000, 96, 80
    g = S.make_gaussians(P, tv.numpy(), seed=1, sh_degree=3)
    cam = S.make_camera(W, H, angle=0.4, radius=3.0, seed=1, focal=537.0 * W / 512)
    pose = S.make_pose_params(1)
    out = {}
    prev = torch.cuda.current_device()
    try:
        for idx in (0, 1):
            torch.cuda.set_device(idx)
            dev = torch.device("cuda")                      # index-less on purpose
            deformer = SMPLDeformer(smpl_model)             # default device argument
            assert deformer.device == torch.device("cuda", idx)
            L = {k: g[k].to(dev) for k in ("xyz", "log_scale", "rotation", "features", "opacity_logit")}
            rs = dgr.GaussianRasterizationSettings(
                image_height=H, image_width=W, tanfovx=cam["tanfovx"], tanfovy=cam["tanfovy"], bg=torch.zeros(3, device=dev),
                scale_modifier=1.0, viewmatrix=cam["world_view_transform"].to(dev),
                projmatrix=cam["full_proj_transform"].to(dev), sh_degree=3, campos=cam["camera_center"].to(dev),
                prefiltered=False, debug=False)
            with torch.no_grad():
                color, radii, depth, alpha = render_deformed(
                    deformer, L["xyz"], {k: v.to(dev) for k, v in pose.items()}, {k: v.to(dev) for k, v in big.items()},
                    tv.to(dev), rs, log_scaling=L["log_scale"], rotation=L["rotation"], features=L["features"],
                    opacity_logit=L["opacity_logit"])
                d = deformer.deform_gaussians(L["xyz"], {k: v.to(dev) for k, v in pose.items()},
                                              {k: v.to(dev) for k, v in big.items()}, tv.to(dev))
            torch.cuda.synchronize()
            assert color.device.index == idx and d["means3D"].device.index == idx
            out[idx] = (color.cpu(), radii.cpu(), d["means3D"].cpu())
        assert torch.equal(out[0][1], out[1][1]) and torch.equal(out[0][0], out[1][0]) and torch.equal(out[0][2], out[1][2])
        assert int((out[1][1] > 0).sum()) > P // 2
    finally:
        torch.cuda.set_device(prev)
