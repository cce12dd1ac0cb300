"""Mesh extraction on the GPU - marching cubes (default) and marching tetrahedra - against the CPU loop
implementations and against level-set properties.  The marching-cubes VERTEX SET is pinned: one vertex per grid edge
with a sign change, computed independently in numpy (what scikit-image's Lewiner routine of the reference
interpolates, utils.py:131).  The triangulation is unpinned (skimage is absent in both environments); see
oracle/mesh_oracle.py."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def same_mesh(v, f, vr, fr, tol=1e-4):
    """Same vertex set (nearest-neighbour match within tol) and same oriented triangles up to order."""
    if v.shape[0] != vr.shape[0] or len(f) != len(fr):
        return False
    d = np.abs(v[:, None, :].astype(np.float64) - vr[None, :, :]).max(-1)
    match = d.argmin(1)
    if d.min(1).max() > tol or len(set(match.tolist())) != v.shape[0]:
        return False

    def canon(tri):
        k = min(range(3), key=lambda i: tri[i])
        return (tri[k], tri[(k + 1) % 3], tri[(k + 2) % 3])

    a = sorted(canon([int(match[i]) for i in t]) for t in f)
    b = sorted(canon([int(i) for i in t]) for t in fr)
    return a == b


def test_sphere_and_random_field_match_cpu_marching_cubes():
    from qsp_slam_b200.mesh import marching_cubes
    from oracle.mesh_oracle import marching_cubes_ref, sign_change_vertices
    D = 11
    lin = np.linspace(-1, 1, D, dtype=np.float32)
    X, Y, Z = np.meshgrid(lin, lin, lin, indexing="ij")
    fields = [np.sqrt(X ** 2 + Y ** 2 + Z ** 2) - 0.63,
              np.random.default_rng(0).standard_normal((D, D, D)).astype(np.float32)]
    for sdf in fields:
        sdf = sdf.astype(np.float32)
        v, f = marching_cubes(sdf, 0.0, 2.0 / (D - 1))
        vr, fr = marching_cubes_ref(sdf, 0.0, 2.0 / (D - 1))
        assert v.dtype == np.float32 and f.dtype == np.int32 and v.shape[1] == 3 and f.shape[1] == 3
        assert v.shape[0] == vr.shape[0] and f.shape[0] == len(fr)
        assert f.min() >= 0 and f.max() < v.shape[0]
        # same vertices in the same order, same triangles in the same order (cell-major, table order)
        assert np.abs(v.astype(np.float64) - vr).max() < 1e-6
        assert np.array_equal(f, np.asarray(fr, np.int32))
        # the vertex set is exactly the set of sign-changing grid edges (independent numpy computation)
        sv = sign_change_vertices(sdf, 0.0, 2.0 / (D - 1))
        assert sv.shape == v.shape and np.abs(v.astype(np.float64) - sv).max() < 1e-6
        assert np.unique(f).size == v.shape[0]            # every vertex is used


def test_marching_cubes_vertex_set_on_a_decoded_grid(decoder_state):
    """On the SDF grid of the decoder itself (D = 48): vertices == the sign-changing grid edges (positions within 1e-6,
    same order), the surface is a closed 2-manifold wherever it does not leave the cube, and marching tetrahedra on
    the same grid produce 2-3x the triangles (the reason for the switch)."""
    from qsp_slam_b200.decoder import DeviceDecoder
    from qsp_slam_b200.reconstruct.optimizer import MeshExtractor
    from qsp_slam_b200.mesh import marching_cubes
    from oracle.mesh_oracle import sign_change_vertices
    D = 48
    mx = MeshExtractor(DeviceDecoder(decoder_state, 0), 64, D)
    code = (np.random.default_rng(1).standard_normal(64) * 0.05).astype(np.float32)
    sdf = mx.decode_grid(code)
    v, f = marching_cubes(sdf, 0.0, 2.0 / (D - 1))
    sv = sign_change_vertices(sdf, 0.0, 2.0 / (D - 1))
    assert v.shape == sv.shape and v.shape[0] > 1000
    assert np.abs(v.astype(np.float64) - sv).max() < 1e-6
    e = np.sort(np.concatenate([f[:, [0, 1]], f[:, [1, 2]], f[:, [2, 0]]]), axis=1)
    _, cnt = np.unique(e, axis=0, return_counts=True)
    assert cnt.max() <= 2                                    # 2-manifold: no edge with more than two triangles
    # edges with a single triangle can only lie on the boundary of the cube (open surface there)
    ue = np.unique(e, axis=0)
    single = ue[cnt == 1]
    if single.size:
        on_boundary = (np.abs(np.abs(v[single.reshape(-1)]) - 1.0) < 1e-6).any(1)
        assert on_boundary.all()
    # consistent orientation: every interior directed edge has its opposite
    de = np.concatenate([f[:, [0, 1]], f[:, [1, 2]], f[:, [2, 0]]])
    fwd = set(map(tuple, de.tolist()))
    assert len(fwd) == de.shape[0]
    assert sum((b, a) in fwd for a, b in fwd) == len(fwd) - single.shape[0]
    m = mx.extract_mesh_from_code(code)
    assert np.array_equal(m.vertices, v) and np.array_equal(m.faces, f)
    vt, ft = marching_cubes(sdf, 0.0, 2.0 / (D - 1), method="tets")
    assert 1.8 * f.shape[0] < ft.shape[0] < 3.5 * f.shape[0] and vt.shape[0] > 2 * v.shape[0]


def test_sphere_and_random_field_match_cpu_marching_tets():
    from qsp_slam_b200.mesh import marching_cubes as _mc
    from oracle.mesh_oracle import marching_tets
    marching_cubes = lambda *a: _mc(*a, method="tets")  # noqa: E731
    D = 11
    lin = np.linspace(-1, 1, D, dtype=np.float32)
    X, Y, Z = np.meshgrid(lin, lin, lin, indexing="ij")
    fields = [np.sqrt(X ** 2 + Y ** 2 + Z ** 2) - 0.63,
              np.random.default_rng(0).standard_normal((D, D, D)).astype(np.float32)]
    for sdf in fields:
        sdf = sdf.astype(np.float32)
        v, f = marching_cubes(sdf, 0.0, 2.0 / (D - 1))
        vr, fr = marching_tets(sdf, 0.0, 2.0 / (D - 1))
        assert v.dtype == np.float32 and f.dtype == np.int32 and v.shape[1] == 3 and f.shape[1] == 3
        assert v.shape[0] == vr.shape[0] and f.shape[0] == len(fr)
        assert f.min() >= 0 and f.max() < v.shape[0]
        assert same_mesh(v, f, vr, fr)


@pytest.mark.parametrize("method", ["cubes", "tets"])
def test_sphere_mesh_properties(method):
    from qsp_slam_b200.mesh import marching_cubes as _mc
    marching_cubes = lambda *a: _mc(*a, method=method)  # noqa: E731
    D = 48
    lin = np.linspace(-1, 1, D, dtype=np.float32)
    X, Y, Z = np.meshgrid(lin, lin, lin, indexing="ij")
    sdf = (np.sqrt(X ** 2 + Y ** 2 + Z ** 2) - 0.7).astype(np.float32)
    v, f = marching_cubes(sdf, 0.0, 2.0 / (D - 1))
    r = np.linalg.norm(v, axis=1)
    assert np.abs(r - 0.7).max() < 2e-3                       # vertices lie on the level set
    # closed surface: every undirected edge is shared by exactly two triangles
    e = np.sort(np.concatenate([f[:, [0, 1]], f[:, [1, 2]], f[:, [2, 0]]]), axis=1)
    _, cnt = np.unique(e, axis=0, return_counts=True)
    assert (cnt == 2).all()
    # outward orientation: normals point away from the centre; Euler characteristic of a sphere
    n = np.cross(v[f[:, 1]] - v[f[:, 0]], v[f[:, 2]] - v[f[:, 0]])
    assert ((n * v[f[:, 0]]).sum(1) > 0).all()
    assert v.shape[0] - len(cnt) + f.shape[0] == 2
    area = 0.5 * np.linalg.norm(n, axis=1).sum()
    assert abs(area - 4 * np.pi * 0.49) < 0.02 * 4 * np.pi * 0.49


def test_extract_mesh_from_code(decoder_state):
    from qsp_slam_b200.decoder import DeviceDecoder
    from qsp_slam_b200.reconstruct.optimizer import MeshExtractor
    from qsp_slam_b200.mesh import marching_cubes
    dev = DeviceDecoder(decoder_state, 0)
    mx = MeshExtractor(dev, 64, 32)
    m = mx.extract_mesh_from_code(np.zeros(64, np.float32))
    assert m.vertices.dtype == np.float32 and m.faces.dtype == np.int32
    assert m.vertices.shape[0] > 100 and m.faces.shape[0] > 100
    assert np.abs(m.vertices).max() <= 1.0 + 1e-6
    # device-resident variant gives the same mesh
    sdf = torch.from_numpy(mx.decode_grid(np.zeros(64, np.float32))).cuda()
    v, f = marching_cubes(sdf, 0.0, 2.0 / 31)
    assert np.array_equal(v, m.vertices) and np.array_equal(f, m.faces)
    # empty result for a level outside the value range
    v0, f0 = marching_cubes(sdf, 5.0, 2.0 / 31)
    assert v0.shape == (0, 3) and f0.shape == (0, 3)


@pytest.mark.parametrize("D", [32, 64, 97])
def test_fused_extraction_is_bit_identical_to_the_exact_grid(decoder_state, D):
    """qsp_extract_mesh decodes only the corners of cells the level set can cross with the fp32-equivalent kernel
    (first-pass fp16 elsewhere).  The isosurface kernels read nothing else, so vertices and faces must equal, bit for
    bit, those of the exact grid - through the same call (QSP_FLAG_EXACT_GRID) and through the unfused
    decode_grid + marching_cubes path of the reference (optimizer.py:292-304).  Codes of growing norm: the
    first-pass error and with it the margin grow."""
    from qsp_slam_b200.decoder import DeviceDecoder
    from qsp_slam_b200.reconstruct.optimizer import MeshExtractor
    from qsp_slam_b200.mesh import extract_mesh
    dev = DeviceDecoder(decoder_state, 0)
    mx = MeshExtractor(dev, 64, D)
    rng = np.random.default_rng(D)
    for sigma in (0.0, 0.1, 0.4):
        code = (rng.standard_normal(64) * sigma).astype(np.float32)
        m = mx.extract_mesh_from_code(code)
        me = mx.extract_mesh_from_code(code, exact_grid=True)
        mr = mx.extract_mesh_from_code_reference_grid(code)
        if sigma <= 0.1:
            assert m.vertices.shape[0] > 50              # (a large random code may have no zero level in the cube)
        for other in (me, mr):
            np.testing.assert_array_equal(m.vertices, other.vertices)
            np.testing.assert_array_equal(m.faces, other.faces)
        # the mixed grid: exact wherever it differs from the first pass, and exact on every crossing cell
        v, f, g = extract_mesh(dev.handle, code, 0.0, 2.0 / (D - 1), return_grid_dim=D)
        exact = mx.decode_grid(code)
        fast = dev.decode_sdf(code, np.ascontiguousarray(np.asarray(mx.voxel_points)[:, :3]), mode="fast").reshape(D, D, D)
        refined = g != fast
        assert np.array_equal(g[refined], exact[refined])
        assert refined.sum() < 0.5 * g.size and (refined.sum() > 0 or m.vertices.shape[0] == 0)
        lo = np.minimum.reduce([exact[a:D - 1 + a, b:D - 1 + b, c:D - 1 + c] for a in (0, 1) for b in (0, 1) for c in (0, 1)])
        hi = np.maximum.reduce([exact[a:D - 1 + a, b:D - 1 + b, c:D - 1 + c] for a in (0, 1) for b in (0, 1) for c in (0, 1)])
        cross = (lo < 0) & (hi >= 0)
        for a in (0, 1):
            for b in (0, 1):
                for c in (0, 1):
                    sub_g = g[a:D - 1 + a, b:D - 1 + b, c:D - 1 + c]
                    sub_e = exact[a:D - 1 + a, b:D - 1 + b, c:D - 1 + c]
                    assert np.array_equal(sub_g[cross], sub_e[cross])


def test_two_extractors_share_one_decoder(decoder_state):
    """One decoder handle, extractors of different voxels_dim (LocalMapping.cc:45,65 builds one per class): each call
    re-sends its grid when the other extractor replaced it."""
    from qsp_slam_b200.decoder import DeviceDecoder
    from qsp_slam_b200.reconstruct.optimizer import MeshExtractor
    dev = DeviceDecoder(decoder_state, 0)
    a, b = MeshExtractor(dev, 64, 24), MeshExtractor(dev, 64, 40)
    z = np.zeros(64, np.float32)
    ma, mb, ma2 = a.extract_mesh_from_code(z), b.extract_mesh_from_code(z), a.extract_mesh_from_code(z)
    assert np.array_equal(ma.vertices, ma2.vertices) and np.array_equal(ma.faces, ma2.faces)
    assert mb.vertices.shape[0] > ma.vertices.shape[0]


def test_convert_sdf_voxels_to_mesh_and_ply(decoder_state, tmp_path):
    """The reference's names for the last two steps (utils.py:120-164): isosurface of a decoded grid, PLY on disk."""
    from qsp_slam_b200.decoder import DeviceDecoder
    from qsp_slam_b200.reconstruct.optimizer import MeshExtractor
    from qsp_slam_b200.reconstruct.utils import convert_sdf_voxels_to_mesh, write_mesh_to_ply
    from qsp_slam_b200.map_io import read_ply
    mx = MeshExtractor(DeviceDecoder(decoder_state, 0), 64, 32)
    z = np.zeros(64, np.float32)
    grid = mx.decode_grid(z)
    v, f = convert_sdf_voxels_to_mesh(torch.from_numpy(grid))
    m = mx.extract_mesh_from_code(z)
    assert np.array_equal(v, m.vertices) and np.array_equal(f, m.faces)
    vc, fc = convert_sdf_voxels_to_mesh(torch.from_numpy(grid).cuda())
    assert np.array_equal(vc, v) and np.array_equal(fc, f)
    write_mesh_to_ply(v, f, str(tmp_path / "m.ply"))
    v2, f2 = read_ply(str(tmp_path / "m.ply"))
    assert np.array_equal(v2, v) and np.array_equal(f2, f)
