"""-m gpu parity of the checkpoint rows: save_ply must write the reference's file byte for byte, load_ply
must return the reference's tensors bit for bit (golden files from the reference's own statements), at any size."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden")
FIELDS = ("xyz", "normal", "albedo", "roughness", "features_dc", "features_rest", "opacity", "scaling", "rotation")


def params(P, seed, M=16):
    g = torch.Generator().manual_seed(seed)
    mk = lambda *s: torch.randn(*s, generator=g).cuda()
    return dict(xyz=mk(P, 3), normal=mk(P, 3), albedo=mk(P, 3), roughness=mk(P, 1), features_dc=mk(P, 1, 3),
                features_rest=mk(P, M - 1, 3), opacity=mk(P, 1), scaling=mk(P, 3), rotation=mk(P, 4))


def test_save_ply_writes_the_reference_bytes(tmp_path):
    from mygauhuman_b200 import io_ply
    z = np.load(os.path.join(GOLD, "ply_p37.npz"))
    t = {k: torch.from_numpy(z["in_" + k]).cuda() for k in FIELDS}
    p = tmp_path / "sub" / "point_cloud.ply"
    io_ply.save_ply(str(p), **t)
    assert p.read_bytes() == open(os.path.join(GOLD, "ply_p37.ply"), "rb").read()


def test_load_ply_returns_the_reference_tensors():
    from mygauhuman_b200 import io_ply
    z = np.load(os.path.join(GOLD, "ply_p37.npz"))
    out = io_ply.load_ply(os.path.join(GOLD, "ply_p37.ply"))
    for k in FIELDS:
        assert out[k].is_cuda and np.array_equal(out[k].cpu().numpy(), z["out_" + k]), k


@pytest.mark.parametrize("P,M", [(0, 16), (1, 16), (50_000, 16), (1000, 4), (777, 1)])
def test_round_trip_and_oracle_rows(tmp_path, P, M):
    from mygauhuman_b200 import io_ply
    from oracle import ply_oracle as PO
    t = params(P, P + M, M)
    rows = io_ply.pack_rows(**t)
    want = PO.pack_rows(*(t[k].cpu().numpy() for k in FIELDS)) if P else np.zeros((0, 18 + 3 * M), np.float32)
    assert np.array_equal(rows.cpu().numpy(), want)
    p = tmp_path / "m.ply"
    io_ply.save_ply(str(p), **t)
    deg = {16: 3, 4: 1, 1: 0}[M]
    back = io_ply.load_ply(str(p), max_sh_degree=deg)
    for k in FIELDS:
        assert torch.equal(back[k], t[k]), k


def test_optional_fields_property_order_and_errors(tmp_path):
    from mygauhuman_b200 import io_ply
    from oracle import ply_oracle as PO
    t = params(300, 9)
    # absent normal / albedo / roughness are written as zeros
    rows = io_ply.pack_rows(t["xyz"], t["features_dc"], t["features_rest"], t["opacity"], t["scaling"], t["rotation"])
    assert float(rows[:, 3:10].abs().max()) == 0.0 and torch.equal(rows[:, 0:3], t["xyz"])
    # a file with shuffled property order and an extra column loads to the same tensors
    full = io_ply.pack_rows(**t).cpu().numpy()
    names = PO.attribute_names(16)
    rng = np.random.default_rng(0)
    perm = rng.permutation(len(names))
    shuffled = np.concatenate([full[:, perm], np.full((300, 1), 7.0, np.float32)], axis=1)
    p = tmp_path / "shuffled.ply"
    PO.write_ply(str(p), shuffled, [names[i] for i in perm] + ["extra"])
    back = io_ply.load_ply(str(p))
    for k in FIELDS:
        assert torch.equal(back[k], t[k]), k
    # the reference raises KeyError through plyfile on a missing property, AssertionError on a wrong f_rest count
    keep = [i for i, n in enumerate(names) if n != "roughness"]
    PO.write_ply(str(p), full[:, keep], [names[i] for i in keep])
    with pytest.raises(KeyError):
        io_ply.load_ply(str(p))
    with pytest.raises(AssertionError):
        io_ply.load_ply(str(tmp_path / "shuffled.ply") if False else os.path.join(GOLD, "ply_p37.ply"), max_sh_degree=2)
    with pytest.raises(RuntimeError):
        io_ply.pack_rows(**{k: v.cpu() for k, v in t.items()})
