"""Generates the marching-cubes case table of mesh_kernels.cuh (``mc_table.inc``).

The table is DERIVED here, not copied from a published one: for each of the 256 inside/outside patterns of a cell's
corners, the sign-changing cube edges are joined into closed loops by walking the cell's six faces, and every loop
is triangulated as a fan.  A face with four crossings (two diagonally opposite inside corners) is resolved by a fixed
rule - each INSIDE corner is cut off on its own - which depends on the face's four corner states only, so the two
cells sharing the face agree and the surface is watertight.  (scikit-image's Lewiner tables, which the reference
calls through ``measure.marching_cubes_lewiner``, resolve these faces - and the interior ambiguities - from the
values; they are not available here, so the triangulation is unpinned.  The VERTEX SET is the same: one vertex on
every grid edge with a sign change.)

Corner b of a cell sits at offset (b & 1, (b >> 1) & 1, (b >> 2) & 1) along the array axes (x slowest); a corner is
inside when sdf < level.  Triangles are oriented with the normal towards increasing SDF.

    python -m qsp_slam_b200.csrc.gen_mc_table        rewrites qsp_slam_b200/csrc/mc_table.inc
"""
from __future__ import annotations

import os

import numpy as np

CORNER = [np.array([b & 1, (b >> 1) & 1, (b >> 2) & 1], float) for b in range(8)]
# the 12 cube edges (a < b, one bit apart); an edge is named by its index here
EDGES = [(a, a | (1 << k)) for a in range(8) for k in range(3) if not a & (1 << k)]
EDGE_ID = {e: i for i, e in enumerate(EDGES)}
# the 6 faces as cyclic corner quadruples: axis k fixed to 0 or 1
FACES = []
for k in range(3):
    u, v = [(1, 2), (0, 2), (0, 1)][k]
    for side in (0, 1):
        base = side << k
        FACES.append([base, base | (1 << u), base | (1 << u) | (1 << v), base | (1 << v)])


def edge_of(a, b):
    return EDGE_ID[(min(a, b), max(a, b))]


def share_face(e0: int, e1: int) -> bool:
    """Do the two cube edges lie on a common face of the cell?"""
    c = [*EDGES[e0], *EDGES[e1]]
    return any(all(((x >> k) & 1) == side for x in c) for k in range(3) for side in (0, 1))


def triangulate(loop):
    """Triangulation of the loop polygon none of whose diagonals lies in a face of the cell: a diagonal across an
    ambiguous face would be met by the neighbouring cell's triangles in the same plane (an edge with four triangles).
    Returns a list of index triples into the loop, or None."""
    n = len(loop)
    if n == 3:
        return [(0, 1, 2)]

    def rec(idx):
        if len(idx) == 3:
            return [tuple(idx)]
        # the side idx[0]-idx[-1] is closed by a triangle with apex idx[k]
        for k in range(1, len(idx) - 1):
            ok = True
            for a, b in ((idx[0], idx[k]), (idx[k], idx[-1])):
                if abs(idx.index(a) - idx.index(b)) not in (1, len(idx) - 1) or len(idx) > 3:
                    # a-b is a diagonal of the ORIGINAL polygon unless the two are neighbours there
                    if (b - a) % n not in (1, n - 1) and share_face(loop[a], loop[b]):
                        ok = False
            if not ok:
                continue
            left = rec(idx[:k + 1]) if k >= 2 else []
            right = rec(idx[k:]) if len(idx) - k >= 3 else []
            if left is None or right is None:
                continue
            return left + [(idx[0], idx[k], idx[-1])] + right
        return None

    for rot in range(n):
        idx = [(rot + i) % n for i in range(n)]
        t = rec(idx)
        if t is not None:
            return t
    return None


def case_triangles(case: int):
    inside = [(case >> b) & 1 for b in range(8)]
    # segments between edge crossings, face by face
    adj = {}
    for f in FACES:
        cross = []                                         # crossings in cyclic order: (edge id, corner index within the face of the inside end)
        for i in range(4):
            a, b = f[i], f[(i + 1) % 4]
            if inside[a] != inside[b]:
                cross.append((edge_of(a, b), i if inside[a] else (i + 1) % 4))
        if len(cross) == 2:
            pairs = [(cross[0][0], cross[1][0])]
        elif len(cross) == 4:
            # two diagonal inside corners: cut each one off with the two crossings adjacent to it
            pairs = []
            for c in set(cc for _, cc in cross):
                es = [e for e, cc in cross if cc == c]
                assert len(es) == 2
                pairs.append((es[0], es[1]))
        else:
            assert len(cross) == 0
            pairs = []
        for p, q in pairs:
            adj.setdefault(p, []).append(q)
            adj.setdefault(q, []).append(p)
    assert all(len(v) == 2 for v in adj.values())
    # closed loops: every crossing has exactly two neighbours (one segment on each of the two faces of its edge)
    loops, seen = [], set()
    for start in sorted(adj):
        if start in seen:
            continue
        loop, prev, cur = [], None, start
        while True:
            loop.append(cur)
            seen.add(cur)
            n0, n1 = adj[cur]
            nxt = n0 if n0 != prev else n1
            prev, cur = cur, nxt
            if cur == start:
                break
        loops.append(loop)
    tris = []
    mid = lambda e: 0.5 * (CORNER[EDGES[e][0]] + CORNER[EDGES[e][1]])
    for loop in loops:
        assert len(loop) >= 3
        pts = [mid(e) for e in loop]
        area = sum(np.cross(pts[i], pts[(i + 1) % len(pts)]) for i in range(len(pts)))
        # the normal has to point from the inside end of the crossed edges to their outside end
        s = 0.0
        for e in loop:
            a, b = EDGES[e]
            d = CORNER[b] - CORNER[a] if inside[a] else CORNER[a] - CORNER[b]
            s += float(np.dot(area, d))
        assert abs(s) > 1e-9
        if s < 0:
            loop = loop[::-1]
        t = triangulate(loop)
        assert t is not None, (case, loop)
        for i, j, k in t:
            tris.append((loop[i], loop[j], loop[k]))
    return tris


def build_table():
    table = [case_triangles(c) for c in range(256)]
    return table


def write_inc(path: str):
    table = build_table()
    mx = max(len(t) for t in table)
    lines = ["// GENERATED by qsp_slam_b200/csrc/gen_mc_table.py - do not edit.",
             "// c_mc_edges[e] = the two corners of cube edge e; c_mc_ntri[case]; c_mc_tri[case][3*t + k] = edge of corner k of triangle t",
             f"#define MC_MAX_TRI {mx}",
             "__constant__ unsigned char c_mc_edges[12][2] = {" + ", ".join("{%d, %d}" % e for e in EDGES) + "};",
             "__constant__ unsigned char c_mc_ntri[256] = {" + ", ".join(str(len(t)) for t in table) + "};",
             f"__constant__ unsigned char c_mc_tri[256][{3 * mx}] = {{"]
    for t in table:
        flat = [e for tri in t for e in tri] + [0] * (3 * (mx - len(t)))
        lines.append("  {" + ", ".join(str(e) for e in flat) + "},")
    lines.append("};")
    with open(path, "w") as f:
        f.write("\n".join(lines) + "\n")
    return mx


if __name__ == "__main__":
    out = os.path.join(os.path.dirname(os.path.abspath(__file__)), "mc_table.inc")
    print("wrote", out, "max triangles per cell:", write_inc(out))
