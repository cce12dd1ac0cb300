"""MapObjects.txt wire format and the offline extraction driver (System_util.cc:126-152,
extract_map_objects.py:44-61, extract_map_objects_multi.py:78-98)."""
import os

import numpy as np
import pytest

from qsp_slam_b200 import map_io


def _objects(n=3, seed=0):
    rng = np.random.default_rng(seed)
    out = []
    for i in range(n):
        pose = np.eye(4, dtype=np.float32)
        pose[:3, :4] = rng.standard_normal((3, 4)).astype(np.float32) * (10.0 ** rng.integers(-2, 3))
        out.append({"id": 7 * i + 2, "label": i % 2, "pose": pose, "code": (rng.standard_normal(64) * 0.3).astype(np.float32)})
    return out


def test_format_matches_cpp_writer():
    # what `f << fixed << setprecision(9)` and Eigen's aligned row-vector printing produce
    pose = np.eye(4, dtype=np.float32)
    pose[0, 3] = -12.5
    code = np.zeros(64, np.float32)
    code[0], code[1] = -1.25, 10.5
    txt = map_io.format_map_object(5, 1, pose, code).split("\n")
    assert txt[0] == "5 1"
    assert txt[1].startswith("1.000000000 0.000000000 0.000000000 -12.500000000 0.000000000 1.000000000 ")
    assert txt[1].endswith(" ")                                     # every pose value is followed by a blank
    items = txt[2].split(" ")
    assert txt[2].startswith("-1.250000000 10.500000000  0.000000000")   # right-aligned to the widest coefficient
    assert len([t for t in items if t]) == 64


def test_round_trip_and_reference_reader_rules(tmp_path):
    objs = _objects(4)
    path = str(tmp_path / "MapObjects.txt")
    assert map_io.write_map_objects(path, reversed(objs)) == 4       # written sorted by id
    got = map_io.read_map_objects(path)
    assert [g["id"] for g in got] == sorted(o["id"] for o in objs)
    for g, o in zip(got, sorted(objs, key=lambda o: o["id"])):
        assert g["label"] == o["label"]
        assert g["pose"].shape == (4, 4) and g["pose"].dtype == np.float64
        np.testing.assert_array_equal(g["pose"][3], [0, 0, 0, 1])
        np.testing.assert_allclose(g["pose"][:3], o["pose"][:3], atol=5.1e-10 + 1e-7 * np.abs(o["pose"][:3]).max())
        assert g["code"].dtype == np.float32 and g["code"].shape == (64,)
        np.testing.assert_allclose(g["code"], o["code"], atol=1e-7)
    # the single-class reader's rule: int() of the whole first line works when there is no label
    path2 = str(tmp_path / "single.txt")
    map_io.write_map_objects(path2, [dict(o, label=None) for o in objs])
    with open(path2) as f:
        lines = f.readlines()
    assert len(lines) == 12 and int(lines[3]) == sorted(o["id"] for o in objs)[1]
    assert [g["label"] for g in map_io.read_map_objects(path2)] == [None] * 4


def test_truncated_file_drops_partial_record(tmp_path):
    path = str(tmp_path / "MapObjects.txt")
    map_io.write_map_objects(path, _objects(2))
    with open(path) as f:
        lines = f.readlines()
    with open(path, "w") as f:
        f.writelines(lines[:5])                                      # N = len(lines) // 3
    assert len(map_io.read_map_objects(path)) == 1
    open(path, "w").close()
    assert map_io.read_map_objects(path) == []


def test_ply_layout(tmp_path):
    v = np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0], [0, 0, 1]], np.float32)
    f = np.array([[0, 1, 2], [0, 2, 3]], np.int32)
    p = str(tmp_path / "m.ply")
    map_io.write_mesh_to_ply(v, f, p)
    raw = open(p, "rb").read()
    head = raw[:raw.index(b"end_header\n")].decode().splitlines()
    assert head[:2] == ["ply", "format binary_little_endian 1.0"]
    assert "element vertex 4" in head and "element face 2" in head
    assert "property list uchar int vertex_indices" in head
    assert len(raw) == raw.index(b"end_header\n") + 11 + 4 * 12 + 2 * 13
    v2, f2 = map_io.read_ply(p)
    np.testing.assert_array_equal(v2, v)
    np.testing.assert_array_equal(f2, f)


@pytest.mark.gpu
def test_extract_map_objects_driver(tmp_path):
    from qsp_slam_b200 import synth
    from qsp_slam_b200.decoder import DeviceDecoder
    from qsp_slam_b200.reconstruct.optimizer import MeshExtractor
    dev = DeviceDecoder(synth.make_decoder_state(0), 0)
    ex = MeshExtractor(dev, 64, 32)
    objs = _objects(3, seed=4)
    for o in objs:
        o["code"] = (o["code"] * 0.3).astype(np.float32)
    map_io.write_map_objects(str(tmp_path / "MapObjects.txt"), objs)
    done = map_io.extract_map_objects(str(tmp_path), {0: ex, 1: ex})
    assert sorted(done) == sorted(o["id"] for o in objs)
    for o in map_io.read_map_objects(str(tmp_path / "MapObjects.txt")):
        pose = np.load(os.path.join(tmp_path, "objects", "%d.npy" % o["id"]))
        np.testing.assert_array_equal(pose, o["pose"])
        v, f = map_io.read_ply(os.path.join(tmp_path, "objects", "%d.ply" % o["id"]))
        one = ex.extract_mesh_from_code(o["code"])
        assert v.shape == one.vertices.shape and f.shape == one.faces.shape
        np.testing.assert_allclose(v, one.vertices, atol=1e-4)
        assert f.size == 0 or (f.min() >= 0 and f.max() < v.shape[0])


def test_multi_class_settings_yaml(tmp_path):
    """extract_map_objects_multi.py:28-33: OpenCV FileStorage YAML, first line skipped."""
    from qsp_slam_b200 import extract_map_objects_multi as M
    y = tmp_path / "settings.yaml"
    y.write_text("%YAML:1.0\nYoloClasses: [56, 62]\nDecoderPaths: ['weights/chairs_64', 'weights/monitors_64']\nCamera.fx: 535.4\n")
    d = M.read_yaml_file(str(y))
    assert d["YoloClasses"] == [56, 62] and d["DecoderPaths"][1] == "weights/monitors_64"
    a = M.config_parser().parse_args(["-y", str(y), "-m", "map"])
    assert a.voxels_dim == 128 and a.map_dir == "map"


@pytest.mark.gpu
def test_extract_map_objects_multi_driver(tmp_path):
    """Two classes with two different decoders (experiment directories in the reference's checkpoint format): every
    object is meshed by the decoder of its class."""
    import json
    import torch
    from qsp_slam_b200 import synth, extract_map_objects_multi as M
    from qsp_slam_b200.deep_sdf.workspace import config_decoder
    from qsp_slam_b200.reconstruct.optimizer import MeshExtractor
    paths = []
    for cls, seed in ((56, 0), (62, 1)):
        d = tmp_path / f"dec{cls}"
        (d / "ModelParameters").mkdir(parents=True)
        (d / "specs.json").write_text(json.dumps({"CodeLength": 64, "NetworkSpecs": {"dims": [512] * 8, "latent_in": [4],
                                                  "weight_norm": True, "xyz_in_all": False}}))
        st = {"module." + k: torch.from_numpy(np.asarray(v)) for k, v in synth.make_decoder_state(seed).items()}
        torch.save({"model_state_dict": st, "epoch": 1}, str(d / "ModelParameters" / "latest.pth"))
        paths.append(str(d))
    (tmp_path / "settings.yaml").write_text("%YAML:1.0\nYoloClasses: [56, 62]\nDecoderPaths: ['" + paths[0] + "', '" + paths[1] + "']\n")
    objs = _objects(4, seed=9)
    for i, o in enumerate(objs):
        o["code"] = (o["code"] * 0.3).astype(np.float32)
        o["label"] = 56 if i % 2 == 0 else 62
    map_dir = tmp_path / "map"
    map_dir.mkdir()
    map_io.write_map_objects(str(map_dir / "MapObjects.txt"), objs)
    done = M.main(["-y", str(tmp_path / "settings.yaml"), "-m", str(map_dir), "-n", "24"])
    assert sorted(done) == sorted(o["id"] for o in objs)
    ex = {56: MeshExtractor(config_decoder(paths[0]), 64, 24), 62: MeshExtractor(config_decoder(paths[1]), 64, 24)}
    for o in map_io.read_map_objects(str(map_dir / "MapObjects.txt")):
        v, f = map_io.read_ply(os.path.join(map_dir, "objects", "%d.ply" % o["id"]))
        own = ex[o["label"]].extract_mesh_from_code(o["code"])
        other = ex[56 if o["label"] == 62 else 62].extract_mesh_from_code(o["code"])
        assert v.shape == own.vertices.shape
        np.testing.assert_allclose(v, own.vertices, atol=1e-4)
        assert v.shape != other.vertices.shape or not np.allclose(v, other.vertices, atol=1e-4)
