"""TEST INFRASTRUCTURE ONLY - CPU isosurface checkers for qsp_mesh_*: marching cubes (default extractor) and
marching tetrahedra (Kuhn subdivision, QSP_FLAG_MESH_TETS).

The reference extracts meshes with scikit-image's Lewiner marching cubes (reconstruct/utils.py:131),
which is absent here and whose case tables are not restated; parity for the TRIANGULATION is therefore
UNPINNED against the reference.  What is pinned: the vertex set of marching cubes - one vertex, by linear
interpolation, on every grid edge whose end points lie on different sides of the level (``sign_change_vertices``,
plain numpy, independent of any table) - which is the vertex set of the reference's mesh up to Lewiner's rare
interior vertices.  What is checked besides: the CUDA kernels against these loop implementations (same vertices,
same triangles up to ordering) and level-set properties (closed 2-manifold, orientation).
"""
from __future__ import annotations

import numpy as np

TETS = [(0, 1, 3, 7), (0, 1, 5, 7), (0, 2, 3, 7), (0, 2, 6, 7), (0, 4, 5, 7), (0, 4, 6, 7)]


def corner(b):
    return np.array([b & 1, (b >> 1) & 1, (b >> 2) & 1])


def sign_change_vertices(sdf: np.ndarray, level: float, spacing: float, origin=(-1.0, -1.0, -1.0)) -> np.ndarray:
    """Positions (n,3) float64 of the marching-cubes vertices: for every pair of axis neighbours a, b with
    (sdf[a] < level) != (sdf[b] < level) the point a + w (b - a), w = (sdf[a]-level) / (sdf[a] - sdf[b]) in fp32 -
    what ``measure.marching_cubes_lewiner`` interpolates (utils.py:131).  Order: by grid vertex, then axis x, y, z."""
    s = sdf.astype(np.float32) - np.float32(level)
    D = s.shape[0]
    inside = s < 0
    out = []
    org = np.asarray(origin, np.float64)
    idx = np.indices((D, D, D)).reshape(3, -1).T
    lin = np.arange(D ** 3)
    per_axis = []
    for ax in range(3):
        ok = idx[:, ax] < D - 1
        a = idx[ok]
        b = a.copy()
        b[:, ax] += 1
        cross = inside[a[:, 0], a[:, 1], a[:, 2]] != inside[b[:, 0], b[:, 1], b[:, 2]]
        a, b = a[cross], b[cross]
        sa, sb = s[a[:, 0], a[:, 1], a[:, 2]], s[b[:, 0], b[:, 1], b[:, 2]]
        w = (sa / (sa - sb)).astype(np.float32).astype(np.float64)
        p = a.astype(np.float64)
        p[:, ax] += w
        per_axis.append((lin[ok][cross], ax, p * spacing + org))
    keys = np.concatenate([k * 3 + ax for k, ax, _ in per_axis])
    pts = np.concatenate([p for _, _, p in per_axis], 0)
    return pts[np.argsort(keys, kind="stable")]


def marching_cubes_ref(sdf: np.ndarray, level: float, spacing: float, origin=(-1.0, -1.0, -1.0)):
    """Loop implementation of marching cubes with the generated case table (csrc/gen_mc_table.py).  Returns
    (vertices (V,3) float64, triangles as lists of 3 vertex ids); vertex order as ``sign_change_vertices``."""
    from qsp_slam_b200.csrc import gen_mc_table as T
    table = T.build_table()
    D = sdf.shape[0]
    s = sdf.astype(np.float32) - np.float32(level)
    inside = s < 0
    org = np.asarray(origin, np.float64)
    verts, index, tris = [], {}, []
    # vertex ids in the order of the kernels: grid vertex major, axis x, y, z
    for x in range(D):
        for y in range(D):
            for z in range(D):
                for ax in range(3):
                    b = [x, y, z]
                    b[ax] += 1
                    if b[ax] < D and inside[x, y, z] != inside[tuple(b)]:
                        sa, sb = np.float32(s[x, y, z]), np.float32(s[tuple(b)])
                        w = np.float64(np.float32(sa / (sa - sb)))
                        p = np.array([x, y, z], np.float64)
                        p[ax] += w
                        index[((x, y, z), tuple(b))] = len(verts)
                        verts.append(p * spacing + org)
    for x in range(D - 1):
        for y in range(D - 1):
            for z in range(D - 1):
                case = 0
                for b in range(8):
                    if inside[x + (b & 1), y + ((b >> 1) & 1), z + ((b >> 2) & 1)]:
                        case |= 1 << b
                for t in table[case]:
                    tri = []
                    for e in t:
                        a, b = T.EDGES[e]
                        pa = (x + (a & 1), y + ((a >> 1) & 1), z + ((a >> 2) & 1))
                        pb = (x + (b & 1), y + ((b >> 1) & 1), z + ((b >> 2) & 1))
                        tri.append(index[(pa, pb)])
                    tris.append(tri)
    return np.array(verts).reshape(-1, 3), tris


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
