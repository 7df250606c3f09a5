"""The PLY oracle (oracle/ply_oracle.py) against golden files produced by the reference's own save_ply / load_ply
statements (oracle/gen_golden_ply.py), and the host-side header parsing of mygauhuman_b200.io_ply (no GPU)."""
import os

import numpy as np
import pytest

from oracle import ply_oracle as PO

GOLD = os.path.join(os.path.dirname(__file__), "golden")
FIELDS = ("xyz", "normal", "albedo", "roughness", "features_dc", "features_rest", "opacity", "scaling", "rotation")


def test_oracle_writer_reproduces_the_reference_file(tmp_path):
    z = np.load(os.path.join(GOLD, "ply_p37.npz"))
    rows = PO.pack_rows(*(z["in_" + k] for k in FIELDS))
    assert rows.shape == (37, 66)
    p = tmp_path / "o.ply"
    PO.write_ply(str(p), rows, PO.attribute_names(16))
    assert p.read_bytes() == open(os.path.join(GOLD, "ply_p37.ply"), "rb").read()


def test_oracle_reader_reproduces_the_reference_load():
    z = np.load(os.path.join(GOLD, "ply_p37.npz"))
    rows, names = PO.read_ply(os.path.join(GOLD, "ply_p37.ply"))
    assert names == PO.attribute_names(16)
    out = PO.unpack_rows(rows, names)
    for k in FIELDS:
        assert np.array_equal(out[k], z["out_" + k]), k


def test_attribute_names_and_header_parser_agree_with_the_oracle(tmp_path):
    from mygauhuman_b200 import io_ply
    assert io_ply.attribute_names(16) == PO.attribute_names(16)
    rows, names = io_ply._read_vertex_rows(os.path.join(GOLD, "ply_p37.ply"))
    ref_rows, ref_names = PO.read_ply(os.path.join(GOLD, "ply_p37.ply"))
    assert names == ref_names and np.array_equal(rows, ref_rows)
    bad = tmp_path / "bad.ply"
    bad.write_bytes(b"ply\nformat ascii 1.0\nelement vertex 0\nproperty float x\nend_header\n")
    with pytest.raises(ValueError):
        io_ply._read_vertex_rows(str(bad))
    bad.write_bytes(b"ply\nformat binary_little_endian 1.0\nelement vertex 2\nproperty float x\nend_header\n\0\0\0\0")
    with pytest.raises(ValueError):
        io_ply._read_vertex_rows(str(bad))   # truncated
    bad.write_bytes(b"ply\nformat binary_little_endian 1.0\nelement vertex 1\nproperty double x\nend_header\n" + b"\0" * 8)
    with pytest.raises(ValueError):
        io_ply._read_vertex_rows(str(bad))
