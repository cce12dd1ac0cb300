"""On-disk map format ``MapObjects.txt`` and the offline extraction driver.

Format (writer: src/System_util.cc:126-152; readers: extract_map_objects.py:44-61,
extract_map_objects_multi.py:78-98; assets/保存格式.md): three text lines per object,
  1. ``<mnId> <label>``            (the single-class reader parses only an integer id on this line)
  2. 12 floats, the top 3x4 of the object->world Sim(3) matrix, row-major, ``fixed`` with 9 decimals,
     each followed by one space
  3. the shape code (64 floats, ``fixed``, 9 decimals) as Eigen prints a row vector: coefficients
     right-aligned to a common width and separated by one space (hence runs of blanks, which the
     readers skip)
"""
from __future__ import annotations

import os
import struct
from typing import Dict, Iterable, List, Optional, Sequence

import numpy as np


def format_map_object(obj_id: int, label: Optional[int], pose, code) -> str:
    """The three lines the C++ writer emits for one object (System_util.cc:145-151)."""
    pose = np.asarray(pose, np.float32)
    code = np.asarray(code, np.float32).reshape(-1)
    head = f"{int(obj_id)}" if label is None else f"{int(obj_id)} {int(label)}"
    line_pose = "".join(f"{float(pose[r, c]):.9f} " for r in range(3) for c in range(4))
    items = [f"{float(x):.9f}" for x in code]
    width = max(len(s) for s in items) if items else 0
    line_code = " ".join(s.rjust(width) for s in items)          # Eigen's default IOFormat aligns columns
    return f"{head}\n{line_pose}\n{line_code}\n"


def write_map_objects(path: str, objects: Iterable[dict]) -> int:
    """objects: dicts with ``id``, optional ``label``, ``pose`` (>=3x4) and ``code``; written sorted by id
    like the C++ side (System_util.cc:131).  Returns the number of objects written."""
    objs = sorted(objects, key=lambda o: int(o["id"]))
    with open(path, "w") as f:
        for o in objs:
            f.write(format_map_object(o["id"], o.get("label"), o["pose"], o["code"]))
    return len(objs)


def read_map_objects(path: str) -> List[dict]:
    """Parses ``MapObjects.txt`` the way both reference readers do: ``N = len(lines) // 3`` records, the first
    integer of line 1 is the id, an optional second one the class id; pose is padded to 4x4 with
    [0 0 0 1] (float64, as ``np.asarray([float(x) ...])`` gives), the code is float32."""
    with open(path) as f:
        lines = f.readlines()
    out = []
    for i in range(len(lines) // 3):
        info = lines[3 * i].strip().split(" ")
        info = [t for t in info if t]
        obj_id = int(info[0])
        label = int(info[1]) if len(info) > 1 else None
        pose = np.asarray([float(x) for x in lines[3 * i + 1].strip().split(" ")]).reshape(3, 4)
        pose = np.concatenate([pose, np.array([0.0, 0.0, 0.0, 1.0]).reshape(1, 4)], axis=0)
        code = np.asarray([float(t) for t in lines[3 * i + 2].strip().split(" ") if len(t) > 0]).astype(np.float32)
        out.append({"id": obj_id, "label": label, "pose": pose, "code": code})
    return out


def write_mesh_to_ply(v, f, ply_filename_out: str) -> None:
    """Binary little-endian PLY with the element/property layout ``plyfile`` produces for the reference
    (reconstruct/utils.py:144-164): vertex (float x, y, z), face (list uchar int vertex_indices)."""
    v = np.ascontiguousarray(v, "<f4").reshape(-1, 3)
    f = np.ascontiguousarray(f, "<i4").reshape(-1, 3)
    header = ("ply\nformat binary_little_endian 1.0\n"
              f"element vertex {v.shape[0]}\nproperty float x\nproperty float y\nproperty float z\n"
              f"element face {f.shape[0]}\nproperty list uchar int vertex_indices\nend_header\n")
    faces = np.empty(f.shape[0], dtype=[("n", "u1"), ("idx", "<i4", (3,))])
    faces["n"] = 3
    faces["idx"] = f
    with open(ply_filename_out, "wb") as out:
        out.write(header.encode("ascii"))
        out.write(v.tobytes())
        out.write(faces.tobytes())


def read_ply(path: str):
    """Reads back what ``write_mesh_to_ply`` wrote (tests)."""
    with open(path, "rb") as fh:
        data = fh.read()
    end = data.index(b"end_header\n") + len(b"end_header\n")
    nv = nf = 0
    for line in data[:end].decode("ascii").splitlines():
        if line.startswith("element vertex"):
            nv = int(line.split()[-1])
        if line.startswith("element face"):
            nf = int(line.split()[-1])
    v = np.frombuffer(data, "<f4", nv * 3, end).reshape(nv, 3)
    faces = np.frombuffer(data, np.dtype([("n", "u1"), ("idx", "<i4", (3,))]), nf, end + nv * 12)
    return v.copy(), faces["idx"].copy()


def extract_map_objects(map_dir: str, extractors, default_class: int = 0) -> List[int]:
    """The body of extract_map_objects.py:40-61 / extract_map_objects_multi.py:74-98: for every object of
    ``<map_dir>/MapObjects.txt`` save ``objects/<id>.npy`` (4x4 pose) and ``objects/<id>.ply`` (mesh of its code).
    ``extractors``: one MeshExtractor, or a dict class id -> MeshExtractor.  One native call per object
    (qsp_extract_mesh: first-pass grid decode, exact re-evaluation of the cells near the level set, isosurface
    kernels; only vertices and faces come back).  The many-grids-in-one-launch form of BASELINE config 4 is
    ``MeshExtractor.decode_grid(codes)`` / bench.py's ``mesh_grid_128x64``, which decodes full grids and is not needed for the meshes."""
    save_dir = os.path.join(map_dir, "objects")
    os.makedirs(save_dir, exist_ok=True)
    objs = read_map_objects(os.path.join(map_dir, "MapObjects.txt"))
    by_class: Dict[int, List[dict]] = {}
    for o in objs:
        by_class.setdefault(o["label"] if (o["label"] is not None and isinstance(extractors, dict)) else default_class, []).append(o)
    done = []
    for cls, group in by_class.items():
        ex = extractors[cls] if isinstance(extractors, dict) else extractors
        batch = max(1, int(os.environ.get("QSP_MESH_BATCH", "8")))
        for b0 in range(0, len(group), batch):
            part = group[b0:b0 + batch]
            meshes = ex.extract_meshes_from_codes(np.stack([o["code"][:ex.code_len] for o in part]))
            for o, mesh in zip(part, meshes):
                np.save(os.path.join(save_dir, "%d.npy" % o["id"]), o["pose"])
                write_mesh_to_ply(mesh.vertices, mesh.faces, os.path.join(save_dir, "%d.ply" % o["id"]))
                done.append(o["id"])
    return done
