"""Reference point_cloud.ply layout (scene/gaussian_model.py:317-386): property order, channel-major SH /
feature columns, round trip through trase_b200.ply_io.  CPU only."""
import numpy as np
import torch

from trase_b200 import ply_io
from trase_b200.render import GaussianParams


def _pc(P=37):
    g = torch.Generator().manual_seed(2)
    r = lambda *s: torch.randn(*s, generator=g)
    return GaussianParams(r(P, 3), r(P, 1, 3), r(P, 15, 3), r(P, 3), r(P, 4), r(P, 1), r(P, 1, 32))


def test_property_order_is_the_references():
    names = ply_io.attribute_names(15, 32, with_cls=True)
    assert names[:9] == ["x", "y", "z", "nx", "ny", "nz", "f_dc_0", "f_dc_1", "f_dc_2"]
    assert names[9] == "f_rest_0" and names[9 + 44] == "f_rest_44" and names[54] == "opacity"
    assert names[55:62] == ["scale_0", "scale_1", "scale_2", "rot_0", "rot_1", "rot_2", "rot_3"]
    assert names[62] == "gaussian_feats_0" and names[93] == "gaussian_feats_31" and names[-1] == "cls" and len(names) == 95


def test_round_trip_and_channel_major_columns(tmp_path):
    pc = _pc()
    cls = torch.arange(37, dtype=torch.float32).reshape(37, 1)
    path = str(tmp_path / "point_cloud" / "iteration_7" / "point_cloud.ply")
    ply_io.save_ply(path, pc, cls=cls)
    raw = ply_io.read_vertex_table(path)
    # the reference flattens [P,15,3] -> transpose -> [P,3,15]: f_rest_(c*15 + k) holds coefficient k of channel c
    assert np.array_equal(raw["f_rest_17"], pc._features_rest.detach()[:, 2, 1].numpy())
    assert np.array_equal(raw["f_dc_2"], pc._features_dc.detach()[:, 0, 2].numpy())
    assert np.array_equal(raw["gaussian_feats_5"], pc._gaussian_features.detach()[:, 0, 5].numpy())
    assert not raw["nx"].any()
    back, cls2 = ply_io.load_ply(path, device="cpu")
    for name in ("_xyz", "_features_dc", "_features_rest", "_scaling", "_rotation", "_opacity", "_gaussian_features"):
        assert torch.equal(getattr(back, name).detach(), getattr(pc, name).detach()), name
    assert torch.equal(cls2, cls)
    # header is plain ASCII, binary little endian, one vertex element
    head = open(path, "rb").read(200).decode("ascii", "ignore")
    assert head.startswith("ply\nformat binary_little_endian 1.0\nelement vertex 37\nproperty float x\n")
