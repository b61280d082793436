"""The bundled mesh generator (omega_b200/meshgen.py) held to the criteria the reference applies to the meshes it reads
(test/ocn/HorzMeshTest.cpp:178-650: element counts, every point on one sphere, lon / lat consistent with x, y, z,
bathymetry range, cell / triangle / kite areas summing to the sphere within 5 %, dcEdge and dvEdge equal to the great
circle distances, angleEdge in [-pi, pi], |weightsOnEdge| <= 1, the edge-sign convention) plus the structure a TRiSK
mesh has by construction: Euler's formula, every reference mutual, kites tiling cells and triangles, the antisymmetry
of the tangential-reconstruction weights that makes the Coriolis term energy neutral, and the MPAS culler conventions
of cull_cells.  The TRiSK weights are pinned numerically by tests/test_oracle_kat.py (the reference's plane norms)."""
import numpy as np
import pytest

from omega_b200 import decomp, meshgen
from tests import cases

TOL = 1e-6   # HorzMeshTest.cpp:45


def sphere_distance(lon1, lat1, lon2, lat2):
    """HorzMeshTest.cpp:32-43 (haversine)"""
    a = np.sin((lat2 - lat1) / 2) ** 2 + np.cos(lat1) * np.cos(lat2) * np.sin((lon2 - lon1) / 2) ** 2
    return 2 * np.arctan2(np.sqrt(a), np.sqrt(1 - a))


MESHES = {
    "icos3": lambda: cases.sphere_mesh(3),
    "icos5": lambda: meshgen.icosahedral(5),
    "land4": lambda: cases.land_mesh(4),
    "hept3": lambda: cases.hept_mesh(3),
}


@pytest.fixture(scope="module", params=list(MESHES))
def mesh(request):
    return request.param, MESHES[request.param]()


def test_reference_horzmesh_criteria(mesh):
    name, m = mesh
    R = m.sphere_radius
    for el in ("Cell", "Edge", "Vertex"):
        x, y, z = m["x" + el], m["y" + el], m["z" + el]
        assert np.abs(np.sqrt(x * x + y * y + z * z) - R).max() <= TOL * R            # one sphere (:196-214)
        lon = np.arctan2(y, x); lon = np.where(lon < 0, lon + 2 * np.pi, lon)          # computeLon / computeLat (:20-30)
        lat = np.arcsin(np.clip(z / R, -1, 1))
        dlon = np.abs(lon - m["lon" + el]); dlon = np.minimum(dlon, 2 * np.pi - dlon)
        pole = np.abs(np.abs(lat) - np.pi / 2) < 1e-9                                 # the longitude of a pole is free
        assert dlon[~pole].max() <= TOL and np.abs(lat - m["lat" + el]).max() <= TOL   # (:216-242)
    assert m.bottomDepth.min() > 0 and m.bottomDepth.max() < 11000.0                   # (:375-395)
    sphere = 4 * np.pi * R * R
    ocean = sphere * (1.0 if not name.startswith("land") else m.areaCell.sum() / sphere)
    assert abs(m.areaCell.sum() - ocean) / ocean < 0.05                                # (:398-415)
    if not name.startswith("land"):
        assert abs(m.areaTriangle.sum() - sphere) / sphere < 0.05                      # (:418-434)
        assert abs(m.kiteAreasOnVertex.sum() - sphere) / sphere < 0.05                 # (:437-455)
    c0, c1 = m.cellsOnEdge[:, 0], m.cellsOnEdge[:, 1]
    both = c1 >= 0
    dc = R * sphere_distance(m.lonCell[c0[both]], m.latCell[c0[both]], m.lonCell[c1[both]], m.latCell[c1[both]])
    assert np.abs(dc / m.dcEdge[both] - 1).max() <= TOL                                # (:458-484)
    v0, v1 = m.verticesOnEdge[:, 0], m.verticesOnEdge[:, 1]
    dv = R * sphere_distance(m.lonVertex[v0], m.latVertex[v0], m.lonVertex[v1], m.latVertex[v1])
    assert np.abs(dv / m.dvEdge - 1).max() <= TOL                                      # (:486-515)
    assert np.abs(m.angleEdge).max() <= np.pi                                          # (:517-532)
    assert np.abs(m.weightsOnEdge).max() <= 1.0                                        # (:594-612)
    om = meshgen.OMEGA_EARTH
    for el in ("Cell", "Edge", "Vertex"):
        assert np.abs(m["f" + el] - 2 * om * np.sin(m["lat" + el])).max() <= 1e-12     # (:534-592)


def test_edge_sign_convention_after_decomposition(mesh):
    """HorzMeshTest.cpp:614-686 on the local mesh of one task of three: -1 on the first cell of an edge, +1 on the second;
    +1 at the second vertex, -1 at the first (HorzMesh.cpp:527-588)"""
    from tests.oracle_lib import Oracle
    from omega_b200 import capi
    name, m = mesh
    if m.nCells > 5000:
        pytest.skip("one size is enough")
    L = decomp.build_all(m, 3, 1)[1]
    o = Oracle(L, capi.TendConfig.default(), 0)
    esc, esv, mask = o.derived()
    o.close()
    for e in range(L.NEdgesOwned):
        for side, want in ((0, -1.0), (1, 1.0)):
            c = L.CellsOnEdge[e, side]
            if c < L.NCellsAll:
                j = list(L.EdgesOnCell[c, :L.NEdgesOnCell[c]]).index(e)
                assert esc[c, j] == want
            v = L.VerticesOnEdge[e, side]
            if v < L.NVerticesAll:
                j = list(L.EdgesOnVertex[v]).index(e)
                assert esv[v, j] == want
        coastal_or_cut = L.CellsOnEdge[e, 0] == L.NCellsAll or L.CellsOnEdge[e, 1] == L.NCellsAll
        assert mask[e] == (0.0 if coastal_or_cut else 1.0)                             # HorzMesh.cpp:590-598


def test_structure(mesh):
    name, m = mesh
    nC, nE, nV = m.nCells, m.nEdges, m.nVertices
    if not name.startswith("land"):
        assert nV - nE + nC == 2                                                       # Euler, closed sphere
        assert (m.cellsOnEdge >= 0).all() and (m.cellsOnVertex >= 0).all() and (m.edgesOnVertex >= 0).all()
    # cell <-> edge, cell <-> vertex, vertex <-> edge references are mutual
    for c in range(nC):
        n = m.nEdgesOnCell[c]
        assert (m.edgesOnCell[c, :n] >= 0).all() and (m.edgesOnCell[c, n:] < 0).all()
        for j in range(n):
            e, v, c2 = m.edgesOnCell[c, j], m.verticesOnCell[c, j], m.cellsOnCell[c, j]
            assert c in m.cellsOnEdge[e] and c in m.cellsOnVertex[v]
            other = m.cellsOnEdge[e][m.cellsOnEdge[e] != c]
            assert (other[0] if other.size else -1) == c2                             # the neighbour ACROSS edge j
    for e in range(nE):
        assert m.cellsOnEdge[e, 0] >= 0                                                # the owner rule needs it
        for v in m.verticesOnEdge[e]:
            assert e in m.edgesOnVertex[v]
    # kites tile cells and triangles
    area = np.zeros(nC)
    ok = m.cellsOnVertex >= 0
    np.add.at(area, m.cellsOnVertex[ok], m.kiteAreasOnVertex[ok])
    assert np.abs(area / m.areaCell - 1).max() <= 1e-12
    assert np.abs(m.kiteAreasOnVertex.sum(1) / m.areaTriangle - 1).max() <= 1e-12
    # v0 -> v1 = k x n: the vertices of an edge are ordered to the left of its normal (c0 -> c1)
    p = lambda el, i: np.stack([m["x" + el][i], m["y" + el][i], m["z" + el][i]], -1)
    both = m.cellsOnEdge[:, 1] >= 0
    if not name.startswith("hept"):      # the flipped pairs of the heptagon mesh are not Delaunay: dual edges reversed
        n_vec = p("Cell", m.cellsOnEdge[both, 1]) - p("Cell", m.cellsOnEdge[both, 0])
        t_vec = p("Vertex", m.verticesOnEdge[both, 1]) - p("Vertex", m.verticesOnEdge[both, 0])
        k_vec = p("Edge", np.nonzero(both)[0])
        assert (np.einsum("ij,ij->i", np.cross(k_vec, n_vec), t_vec) > 0).all()


def test_trisk_weights_are_antisymmetric(mesh):
    """w(e, e') dc(e) / dv(e') = -w(e', e) dc(e') / dv(e): what makes the nonlinear Coriolis term energy neutral
    (Thuburn et al. 2009, eq. 24; Ringler et al. 2010) -- holds cell by cell, so coasts and heptagons keep it"""
    name, m = mesh
    table = {}
    for e in range(m.nEdges):
        for j in range(m.nEdgesOnEdge[e]):
            e2 = int(m.edgesOnEdge[e, j])
            table[(e, e2)] = table.get((e, e2), 0.0) + m.weightsOnEdge[e, j] * m.dcEdge[e] / m.dvEdge[e2]
    worst = 0.0
    for (e, e2), w in table.items():
        assert (e2, e) in table
        worst = max(worst, abs(w + table[(e2, e)]))
    assert worst <= 1e-13
    assert (m.edgesOnEdge[np.arange(m.maxEdges2)[None, :] >= m.nEdgesOnEdge[:, None]] < 0).all()


def test_cull_cells_conventions():
    full, m = cases.sphere_mesh(4), cases.land_mesh(4)
    ce = m.culled_from
    assert m.nCells < full.nCells and (m.cellsOnEdge[:, 1] < 0).sum() > 50 and ce["swapped"].sum() > 10
    # geometry of what is left did not move; a swapped edge turned by pi
    assert np.array_equal(m.areaCell, full.areaCell[ce["cells"]]) and np.array_equal(m.dvEdge, full.dvEdge[ce["edges"]])
    d = np.abs(m.angleEdge - full.angleEdge[ce["edges"]])
    assert np.allclose(d[ce["swapped"]], np.pi) and (d[~ce["swapped"]] == 0).all()
    # a coastal edge keeps only the edges of its one cell; an interior edge keeps its weights (sign turned with a swapped partner)
    coast = m.cellsOnEdge[:, 1] < 0
    assert (m.nEdgesOnEdge[coast] == m.nEdgesOnCell[m.cellsOnEdge[coast, 0]] - 1).all()
    assert np.array_equal(m.nEdgesOnEdge[~coast], full.nEdgesOnEdge[ce["edges"][~coast]])
    # cellsOnCell keeps its alignment with edgesOnCell: the -1 sits where the coastal edge is
    for c in np.nonzero((m.cellsOnCell < 0).any(1) & (m.nEdgesOnCell == 6))[0][:50]:
        for j in range(6):
            assert (m.cellsOnCell[c, j] < 0) == bool(coast[m.edgesOnCell[c, j]])
    with pytest.raises(ValueError):
        meshgen.cull_cells(full, np.zeros(full.nCells, bool))


def test_heptagon_mesh_counts():
    m = cases.hept_mesh(3)
    hist = np.bincount(m.nEdgesOnCell, minlength=8)
    assert m.maxEdges == 7 and hist[7] == 12 and hist[5] == 24 and hist[6] == m.nCells - 36
    assert m.nEdgesOnEdge.max() == 12 and m.maxEdges2 == 14


def test_planar_mesh_is_a_torus():
    m = cases.plane_mesh(12, 12, 1000.0)
    assert m.nVertices - m.nEdges + m.nCells == 0 and (m.nEdgesOnCell == 6).all() and (m.nEdgesOnEdge == 10).all()
    assert np.allclose(m.areaCell.sum(), m.Lx * m.Ly) and np.allclose(m.areaTriangle.sum(), m.Lx * m.Ly)
