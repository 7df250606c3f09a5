"""CPU: the host-side mirror of the reference's LBSOffsetDecoder (mygauhuman_b200.nets) against the golden vector produced
by the reference's own module (oracle/gen_golden_mlp.py): same state_dict keys and shapes, and its statement-by-statement
restatement (`torch_forward`, used for the backward) reproduces the reference output.  The tensor-core forward itself needs
a GPU (tests/test_mlp_gpu.py); without one the module must fail loudly, not fall back."""
import os

import numpy as np
import pytest
import torch

GOLDEN = os.path.join(os.path.dirname(__file__), "golden", "lbs_mlp.npz")


def test_state_dict_layout_and_restatement_match_the_reference_golden():
    from mygauhuman_b200.nets import LBSOffsetDecoder, xyz_embedder
    g = np.load(GOLDEN)
    m = LBSOffsetDecoder()
    want_keys = {k[6:] for k in g.files if k.startswith("param.")}
    assert set(m.state_dict().keys()) == want_keys
    for k, v in m.state_dict().items():
        assert tuple(v.shape) == g["param." + k].shape, k
    m.load_state_dict({k: torch.from_numpy(g["param." + k]) for k in want_keys})
    pts = torch.from_numpy(g["pts"])
    assert xyz_embedder(pts).shape[-1] == 63
    with torch.no_grad():
        out = m.torch_forward(pts)
    assert out.shape == (1, 24, pts.shape[1])
    assert float((out - torch.from_numpy(g["out"])).abs().max()) <= 1e-6


def test_forward_refuses_cpu_tensors():
    from mygauhuman_b200.nets import LBSOffsetDecoder
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        LBSOffsetDecoder()(torch.zeros(1, 4, 3))


def test_pose_refiner_mirror_matches_the_reference_golden():
    from mygauhuman_b200.nets import BodyPoseRefiner
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "pose_refiner.npz"))
    m = BodyPoseRefiner(mlp_width=128, mlp_depth=2)
    want_keys = {k[6:] for k in g.files if k.startswith("param.")}
    assert set(m.state_dict().keys()) == want_keys
    # the reference's initialisation: corrections start at (almost) the identity
    with torch.no_grad():
        Rs0 = m.torch_forward(torch.from_numpy(g["pose"]))
    assert float((Rs0 - torch.eye(3)).abs().max()) < 5e-3
    m.load_state_dict({k: torch.from_numpy(g["param." + k]) for k in want_keys})
    with torch.no_grad():
        Rs = m.torch_forward(torch.from_numpy(g["pose"]))
    assert float((Rs - torch.from_numpy(g["Rs"])).abs().max()) <= 1e-6
