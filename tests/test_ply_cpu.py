"""CPU: the PLY oracle against the reference's own fixtures (known values of GS-IRON/base/gaussian.py:4-67), and the
host half of the C ABI (gsr_ply_parse_header / gsr_ply_write_header need no GPU) against the oracle."""
import os

import numpy as np
import pytest

from gsr_b200 import ply
from oracle import ply_oracle as po

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _raw(name):
    with open(os.path.join(GOLD, name), "rb") as f:
        return f.read()


def test_oracle_reads_the_reference_fixture_known_answers():
    p = po.load_ply(_raw("single_gaussian.ply"))
    f = np.float32
    assert p["xyz"].shape == (1, 3) and np.array_equal(p["xyz"][0], f([0.0270041, 0.571603, 0.508573]))
    assert np.array_equal(p["features_dc"][0, 0], f([3.790736, 0.610315, 1.0232366]))
    # f_rest is channel-major in the file: f_rest_0 -> (coef 1, R), f_rest_15 -> (coef 1, G), f_rest_30 -> (coef 1, B)
    assert np.array_equal(p["features_rest"][0, 0], f([0.102161, 0.0931693, 0.0943818]))
    assert np.array_equal(p["features_rest"][0, 14], f([-0.00659384, -0.00739053, -0.0224955]))
    assert p["opacity"][0, 0] == f(18.81027) and np.array_equal(p["scaling"][0], f([-0.8, -0.8, -0.8]))
    assert np.array_equal(p["rotation"][0], f([0.510566, 0.502488, 0.00997789, -0.140544]))
    a = po.activate(p)
    assert abs(a["scales"][0, 0] - np.exp(-0.8)) < 1e-7 and abs(np.linalg.norm(a["rotations"][0]) - 1) < 1e-6
    assert po.load_ply(_raw("two_gaussians.ply"))["xyz"].shape == (2, 3)


def test_oracle_save_load_round_trip():
    rng = np.random.default_rng(0)
    P = 37
    p = dict(xyz=rng.standard_normal((P, 3)), features_dc=rng.standard_normal((P, 1, 3)), features_rest=rng.standard_normal((P, 15, 3)),
             opacity=rng.standard_normal((P, 1)), scaling=rng.standard_normal((P, 3)), rotation=rng.standard_normal((P, 4)))
    p = {k: v.astype(np.float32) for k, v in p.items()}
    q = po.load_ply(po.save_ply(p))
    for k in p:
        assert np.array_equal(p[k], q[k]), k


@pytest.mark.parametrize("name", ["single_gaussian.ply", "two_gaussians.ply"])
def test_header_parser_matches_oracle_layout(name):
    raw = _raw(name)
    lay = ply.parse_header(raw)
    v = po.read_vertices(raw)
    assert lay.num_vertices == v.shape[0] and lay.record_bytes == v.dtype.itemsize
    assert lay.header_bytes == raw.index(b"end_header\n") + 11 and lay.n_rest == 45
    want = ["x", "y", "z", "nx", "ny", "nz"] + ["f_dc_%d" % i for i in range(3)] + ["f_rest_%d" % i for i in range(45)] + \
           ["opacity"] + ["scale_%d" % i for i in range(3)] + ["rot_%d" % i for i in range(4)]
    for i, n in enumerate(want):
        assert lay.offset[i] == v.dtype.fields[n][1], n


def test_written_header_equals_reference_save_ply_header():
    p = {k: np.zeros(s, np.float32) for k, s in dict(xyz=(5, 3), features_dc=(5, 1, 3), features_rest=(5, 15, 3), opacity=(5, 1),
                                                       scaling=(5, 3), rotation=(5, 4)).items()}
    raw = po.save_ply(p)
    assert ply.write_header(5, 3) == raw[:raw.index(b"end_header\n") + 11]
    lay = ply.parse_header(ply.write_header(5, 3))
    assert lay.record_bytes == 62 * 4 and [lay.offset[i] for i in range(62)] == [4 * i for i in range(62)]
    assert ply.parse_header(ply.write_header(9, 1)).n_rest == 9


@pytest.mark.parametrize("bad,msg", [
    (b"plx\n", "magic"), (b"ply\nformat ascii 1.0\nelement vertex 1\nproperty float x\nend_header\n", "binary_little_endian"),
    (b"ply\nformat binary_little_endian 1.0\nelement vertex 1\nproperty float x\nend_header\n", "required attribute"),
    (b"ply\nformat binary_little_endian 1.0\nelement vertex 1\nproperty list uchar int idx\nend_header\n", "list property"),
    (b"ply\nformat binary_little_endian 1.0\nelement vertex 1\nproperty float x\n", "end_header"),
    (b"ply\nformat binary_little_endian 1.0\nelement vertex 1\nproperty double x\nend_header\n", "property float"),
])
def test_header_errors(bad, msg):
    with pytest.raises(RuntimeError, match=msg):
        ply.parse_header(bad)


def test_mixed_type_properties_are_skipped_by_size():
    head = b"ply\nformat binary_little_endian 1.0\ncomment made by a test\nelement vertex 3\nproperty uchar red\nproperty float x\n" \
           b"property float y\nproperty float z\nproperty double weight\n" + \
           b"".join(b"property float %s\n" % n for n in [b"f_dc_0", b"f_dc_1", b"f_dc_2", b"opacity", b"scale_0", b"scale_1", b"scale_2",
                                                        b"rot_0", b"rot_1", b"rot_2", b"rot_3"]) + b"element face 0\nproperty list uchar int vertex_indices\nend_header\n"
    lay = ply.parse_header(head)
    assert lay.offset[0] == 1 and lay.offset[2] == 9 and lay.offset[6] == 21 and lay.record_bytes == 1 + 12 + 8 + 11 * 4 and lay.n_rest == 0
