"""TEST INFRASTRUCTURE ONLY - CPU marching tetrahedra (Kuhn subdivision), the checker for qsp_mesh_*.

The reference extracts meshes with scikit-image's Lewiner marching cubes (reconstruct/utils.py:131),
which is absent here and whose case tables are not restated; parity for the mesh step is therefore
UNPINNED against the reference.  What is checked instead: the CUDA kernel against this independent
loop implementation (same vertices, same triangles up to ordering) and level-set properties.
"""
from __future__ import annotations

import numpy as np

TETS = [(0, 1, 3, 7), (0, 1, 5, 7), (0, 2, 3, 7), (0, 2, 6, 7), (0, 4, 5, 7), (0, 4, 6, 7)]


def corner(b):
    return np.array([b & 1, (b >> 1) & 1, (b >> 2) & 1])


def marching_tets(sdf: np.ndarray, level: float, spacing: float, origin=(-1.0, -1.0, -1.0)):
    """Returns (vertices (V,3) float64 in output coordinates, triangles as a list of 3 vertex ids)."""
    D = sdf.shape[0]
    s = sdf.astype(np.float32) - np.float32(level)
    verts, index, tris = [], {}, []
    org = np.asarray(origin, np.float64)

    def vid(pa, pb):
        key = (tuple(pa), tuple(pb))
        if key not in index:
            sa, sb = np.float32(s[tuple(pa)]), np.float32(s[tuple(pb)])
            w = np.float32(sa / (sa - sb))
            p = (pa + np.float64(w) * (pb - pa)) * spacing + org
            index[key] = len(verts)
            verts.append(p)
        return index[key]

    for x in range(D - 1):
        for y in range(D - 1):
            for z in range(D - 1):
                base = np.array([x, y, z])
                vals = [s[tuple(base + corner(b))] for b in range(8)]
                if all(v < 0 for v in vals) or all(v >= 0 for v in vals):
                    continue
                for tet in TETS:
                    ins = [q for q in tet if vals[q] < 0]
                    outs = [q for q in tet if vals[q] >= 0]
                    if not ins or not outs:
                        continue
                    def e(a, b):
                        a, b = min(a, b), max(a, b)
                        return vid(base + corner(a), base + corner(b))
                    if len(ins) == 1:
                        faces = [[e(ins[0], o) for o in outs]]
                    elif len(ins) == 3:
                        faces = [[e(outs[0], i) for i in ins]]
                    else:
                        q = [e(ins[0], outs[0]), e(ins[0], outs[1]), e(ins[1], outs[1]), e(ins[1], outs[0])]
                        faces = [[q[0], q[1], q[2]], [q[0], q[2], q[3]]]
                    cin = (base + np.mean([corner(q) for q in ins], 0)) * spacing + org
                    for f in faces:
                        p0, p1, p2 = (verts[i] for i in f)
                        nrm = np.cross(p1 - p0, p2 - p0)
                        if np.dot(nrm, p0 - cin) < 0:
                            f = [f[0], f[2], f[1]]
                        tris.append(f)
    return np.array(verts).reshape(-1, 3), tris
