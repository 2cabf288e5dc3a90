"""PLY + semantics loader / merger, CPU only.  The reference loader needs the
`plyfile` package (absent here), so the pin is the file format itself: a PLY written in the reference's
property order (semantic_3dgs_wrapper.py:60-98) with hand-built header bytes must load into the tensor
layouts of SURVEY.md Appendix A, and merging must follow gaussian_merger.py:232-274."""
import json
import os

import numpy as np
import pytest
import torch

from gsworld_b200 import fixtures as fx
from gsworld_b200.gaussian_merger import GaussianModelMerger, load_ply, read_ply_vertices, save_ply


def _scene(n, seed):
    g = torch.Generator().manual_seed(seed)
    r = lambda *s: torch.randn(*s, generator=g)
    return fx.GaussianScene(r(n, 3), r(n, 1, 3), r(n, 15, 3), r(n, 1, 1), r(n, 3), r(n, 4),
                            torch.randint(-1, 5, (n, 1), generator=g).float())


def test_hand_written_ply_loads_into_reference_layouts(tmp_path):
    # 2 vertices, values = property index + 100 * vertex, so every property's destination is checkable
    names = (["x", "y", "z", "nx", "ny", "nz"] + [f"f_dc_{k}" for k in range(3)] + [f"f_rest_{k}" for k in range(45)]
             + ["opacity"] + [f"scale_{k}" for k in range(3)] + [f"rot_{k}" for k in range(4)])
    vals = np.array([[k + 100 * v for k in range(len(names))] for v in range(2)], dtype="<f4")
    p = tmp_path / "a.ply"
    with open(p, "wb") as f:
        f.write(b"ply\nformat binary_little_endian 1.0\nelement vertex 2\n")
        for n in names:
            f.write(f"property float {n}\n".encode())
        f.write(b"end_header\n")
        f.write(vals.tobytes())
    m = load_ply(str(p), device="cpu")
    idx = {n: k for k, n in enumerate(names)}
    assert m._xyz.shape == (2, 3) and m._xyz[1].tolist() == [100.0, 101.0, 102.0]
    assert m._features_dc.shape == (2, 1, 3) and m._features_dc[0, 0].tolist() == [idx["f_dc_0"], idx["f_dc_1"], idx["f_dc_2"]]
    # f_rest_i with i = c*15 + k  ->  _features_rest[:, k, c]
    assert m._features_rest.shape == (2, 15, 3)
    for c in range(3):
        for k in (0, 7, 14):
            assert float(m._features_rest[0, k, c]) == idx[f"f_rest_{c * 15 + k}"]
    assert m._opacity.shape == (2, 1, 1) and float(m._opacity[1]) == idx["opacity"] + 100
    assert m._scaling[0].tolist() == [idx["scale_0"], idx["scale_1"], idx["scale_2"]]
    assert m._rotation[0].tolist() == [idx[f"rot_{k}"] for k in range(4)]
    assert m._semantics.shape == (2, 1) and float(m._semantics.abs().sum()) == 0.0  # absent -> zeros
    assert m.get_features.shape == (2, 16, 3)


def test_save_load_round_trip_and_header(tmp_path):
    s = _scene(37, 1)
    p = str(tmp_path / "b.ply")
    save_ply(s, p)
    v = read_ply_vertices(p)
    assert v.dtype.names[:6] == ("x", "y", "z", "nx", "ny", "nz") and v.dtype.names[-1] == "semantics"
    assert len(v.dtype.names) == 6 + 3 + 45 + 1 + 3 + 4 + 1
    m = load_ply(p, device="cpu")
    for a in ("_xyz", "_features_dc", "_features_rest", "_opacity", "_scaling", "_rotation", "_semantics"):
        assert torch.equal(getattr(m, a), getattr(s, a)), a


def test_merger_follows_config_order_and_label_rules(tmp_path):
    a, b, c = _scene(20, 2), _scene(7, 3), _scene(5, 4)
    os.makedirs(tmp_path / "scene"), os.makedirs(tmp_path / "objs")
    save_ply(a, str(tmp_path / "scene" / "robot.ply"), with_semantics=False)
    save_ply(b, str(tmp_path / "objs" / "mustard.ply"), with_semantics=False)
    save_ply(c, str(tmp_path / "objs" / "plain.ply"))
    labels = np.arange(20, dtype=np.int64) % 14 - 1
    np.save(tmp_path / "scene" / "robot_semantics_gs.npy", labels)
    cfg = {"models": [
        {"data_path": "./scene/robot.ply", "semantic_labels": "./scene/robot_semantics_gs.npy", "transformation": []},
        {"data_path": "./objs/mustard.ply", "semantic_labels": 100, "transformation": [[2, 0, 0, 0]] * 4},
        {"data_path": "./objs/plain.ply"}]}
    cfg_path = tmp_path / "cfg.json"
    cfg_path.write_text(json.dumps(cfg))
    mg = GaussianModelMerger(device="cpu", asset_dir=str(tmp_path))
    assert mg.load_models_from_config(str(cfg_path)) == [0, 1, 2]
    m = mg.merge_models()
    assert m.num_points == 32
    assert torch.equal(m._xyz, torch.cat([a._xyz, b._xyz, c._xyz]))          # "transformation" is never applied
    assert m._semantics.shape == (32, 1) and m._semantics.dtype == torch.float32
    assert torch.equal(m._semantics[:20, 0].long(), torch.from_numpy(labels))
    assert (m._semantics[20:27] == 100).all()
    assert torch.equal(m._semantics[27:], c._semantics)                       # no label given: the PLY's own column
    with pytest.raises(FileNotFoundError):
        GaussianModelMerger(device="cpu", asset_dir=str(tmp_path)).load_models_from_config(str(tmp_path / "nope.json"))
    with pytest.raises(ValueError):
        (tmp_path / "bad.json").write_text("{}")
        GaussianModelMerger(device="cpu", asset_dir=str(tmp_path)).load_models_from_config(str(tmp_path / "bad.json"))
