"""CPU checks of the marching-cubes case table (qsp_slam_b200/csrc/gen_mc_table.py -> mc_table.inc) and of the
checker in oracle/mesh_oracle.py: the committed table is the generator's output, every case's triangles use exactly
the sign-changing edges, and the surface of random fields is a closed, consistently oriented 2-manifold whose vertex
set is the set of sign-changing grid edges."""
import os
import re

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_committed_table_is_the_generators_output(tmp_path):
    from qsp_slam_b200.csrc import gen_mc_table as T
    out = tmp_path / "mc_table.inc"
    mx = T.write_inc(str(out))
    assert mx == 5
    assert out.read_text() == open(os.path.join(ROOT, "qsp_slam_b200", "csrc", "mc_table.inc")).read()


def test_every_case_uses_exactly_its_crossed_edges():
    from qsp_slam_b200.csrc import gen_mc_table as T
    tab = T.build_table()
    assert len(tab) == 256 and tab[0] == [] and tab[255] == []
    for case, tris in enumerate(tab):
        inside = [(case >> b) & 1 for b in range(8)]
        crossed = {i for i, (a, b) in enumerate(T.EDGES) if inside[a] != inside[b]}
        used = {e for t in tris for e in t}
        assert used == crossed, case
        # loops of n vertices give n - 2 triangles each; never a diagonal inside a face of the cell
        for t in tris:
            assert len(set(t)) == 3


def test_random_fields_give_closed_oriented_manifolds():
    from oracle.mesh_oracle import marching_cubes_ref, sign_change_vertices
    rng = np.random.default_rng(3)
    for trial in range(3):
        D = 10
        g = rng.standard_normal((D, D, D)).astype(np.float32)
        g[0], g[-1], g[:, 0], g[:, -1], g[:, :, 0], g[:, :, -1] = 1, 1, 1, 1, 1, 1      # surface stays inside the cube
        v, tris = marching_cubes_ref(g, 0.0, 1.0, (0.0, 0.0, 0.0))
        sv = sign_change_vertices(g, 0.0, 1.0, (0.0, 0.0, 0.0))
        assert v.shape == sv.shape and np.abs(v - sv).max() < 1e-12
        f = np.asarray(tris)
        de = np.concatenate([f[:, [0, 1]], f[:, [1, 2]], f[:, [2, 0]]])
        fwd = set(map(tuple, de.tolist()))
        assert len(fwd) == de.shape[0]                       # no directed edge twice
        assert all((b, a) in fwd for a, b in fwd)            # every edge has exactly one opposite: closed 2-manifold
        # orientation: the signed volume equals the volume enclosed around the inside (negative) region, > 0
        vol = (v[f[:, 0]] * np.cross(v[f[:, 1]], v[f[:, 2]])).sum() / 6.0
        assert vol > 0
