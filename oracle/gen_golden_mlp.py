"""ORACLE -- TEST INFRASTRUCTURE ONLY.  Golden vectors of the reference's LBSOffsetDecoder
(nets/mlp_delta_weight_lbs.py:5-33), produced by importing the reference's OWN module (baseline/_ref, see
oracle/ref_py.py) in the dev container and running it on the CPU in fp32 (no TF32 there): parameters, 700 positions in
the SMPL body's range, and its output [1, 24, P].  Writes tests/golden/lbs_mlp.npz.

    python oracle/gen_golden_mlp.py
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import ref_py  # noqa: E402


def main():
    ns = ref_py.load()
    torch.manual_seed(11)
    net = ns.nets_lbs.LBSOffsetDecoder(total_bones=24)
    with torch.no_grad():  # trained networks have larger weights than the default initialisation
        for p in net.parameters():
            p.mul_(2.0)
    gen = torch.Generator().manual_seed(12)
    pts = (torch.rand(1, 700, 3, generator=gen) * 2.0 - 1.0) * torch.tensor([0.9, 1.1, 0.4])
    with torch.no_grad():
        out = net(pts)
    data = {"pts": pts.numpy(), "out": out.numpy()}
    for k, v in net.state_dict().items():
        data["param." + k] = v.numpy()
    path = os.path.join(ROOT, "tests", "golden", "lbs_mlp.npz")
    np.savez_compressed(path, **data)
    print(path, out.shape, float(out.abs().max()), float(out.std()))




def main_refiner():
    """The same for BodyPoseRefiner (nets/mlp_delta_body_pose.py) in the configuration scene/gaussian_model.py:94-98 builds
    (width 128, depth 2), last layer scaled up so that the corrections are not ~identity.  Writes tests/golden/pose_refiner.npz."""
    ns = ref_py.load()
    torch.manual_seed(21)
    net = ns.nets_pose.BodyPoseRefiner(total_bones=24, embedding_size=69, mlp_width=128, mlp_depth=2)
    with torch.no_grad():
        net.block_mlps[-1].weight.mul_(2e4)
        net.block_mlps[-1].bias.uniform_(-0.1, 0.1)
    pose = torch.randn(1, 69, generator=torch.Generator().manual_seed(22)) * 0.4
    with torch.no_grad():
        Rs = net(pose)["Rs"]
    data = {"pose": pose.numpy(), "Rs": Rs.numpy()}
    for k, v in net.state_dict().items():
        data["param." + k] = v.numpy()
    path = os.path.join(ROOT, "tests", "golden", "pose_refiner.npz")
    np.savez_compressed(path, **data)
    print(path, Rs.shape, float((Rs - torch.eye(3)).abs().max()))


if __name__ == "__main__":
    main()
    main_refiner()
