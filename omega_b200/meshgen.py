"""Synthetic MPAS-convention mesh generator (host-side, numpy).

The reference ships no mesh generator: its tests download three fixtures
(`components/omega/doc/devGuide/QuickStart.md:146-162`).  This module writes
the two synthetic families BASELINE.json names, in the file schema the
reference reads (`components/omega/src/base/Decomp.cpp:301-395`,
`components/omega/src/ocn/HorzMesh.cpp:424-523`):

* doubly-periodic planar hex  (``planar_hex``)
* icosahedral quasi-uniform sphere, 10*4**n+2 cells (``icosahedral``)

Both go through one connectivity builder that turns a closed, consistently
oriented (counter-clockwise) Delaunay triangulation of the cell centres into
the MPAS Voronoi arrays.  Conventions (all 0-based here, -1 = none; the
NetCDF writer converts to the file's 1-based / 0 = none):

* ``cellsOnVertex[v]``      CCW corners of Delaunay triangle v
* ``cellsOnEdge[e]=(c0,c1)`` edge normal points c0 -> c1
* ``verticesOnEdge[e]=(v0,v1)`` v0 -> v1 is k x n (90 deg CCW of the normal)
* ``edgesOnCell / cellsOnCell / verticesOnCell`` CCW rings; edge j separates
  the cell from ``cellsOnCell[:, j]``; vertex j lies between edge j and j+1
* ``edgesOnVertex[v, j]`` edge between ``cellsOnVertex[v, j]`` and ``[v, j+1]``
* ``edgesOnEdge / weightsOnEdge`` TRiSK tangential-reconstruction stencil
  (Thuburn et al. 2009; Ringler et al. 2010): the other edges of c0 walking
  CCW from e, then the other edges of c1.
"""
from __future__ import annotations

import numpy as np

SPHERE_RADIUS = 6371220.0  # components/omega/test/ocn/TendenciesTest.cpp:27
OMEGA_EARTH = 7.29212e-5   # rad/s


class Mesh(dict):
    """dict of numpy arrays + scalar sizes with attribute access."""

    def __getattr__(self, k):
        try:
            return self[k]
        except KeyError as e:  # pragma: no cover
            raise AttributeError(k) from e

    def __setattr__(self, k, v):
        self[k] = v


# --------------------------------------------------------------------------
# connectivity from a closed CCW triangulation
# --------------------------------------------------------------------------
def _connectivity(tri: np.ndarray, n_cells: int, max_edges: int) -> Mesh:
    tri = np.ascontiguousarray(tri, dtype=np.int64)
    n_vert = tri.shape[0]
    m = Mesh()
    m.nCells, m.nVertices, m.maxEdges, m.vertexDegree = n_cells, n_vert, max_edges, 3

    # half-edges: id = 3*t + p goes tri[t,p] -> tri[t,p+1]; triangle t lies to
    # its left
    he_from = tri.reshape(-1)
    he_to = np.roll(tri, -1, axis=1).reshape(-1)
    n_he = he_from.size
    key = he_from * n_cells + he_to
    order = np.argsort(key, kind="stable")
    skey = key[order]
    if np.any(skey[1:] == skey[:-1]):
        raise ValueError("triangulation is not consistently oriented")
    # twin half-edge (to -> from)
    tkey = he_to * n_cells + he_from
    pos = np.searchsorted(skey, tkey)
    if np.any(pos >= n_he) or np.any(skey[np.minimum(pos, n_he - 1)] != tkey):
        raise ValueError("triangulation is not closed")
    twin = order[pos]

    # undirected edges: one per half-edge pair, numbered by ascending
    # (min cell, max cell); c0 alternates so both normal orientations occur
    he = np.arange(n_he)
    rep = he[he_from < he_to]
    rep = rep[np.argsort(he_from[rep] * n_cells + he_to[rep], kind="stable")]
    n_edges = rep.size
    m.nEdges = n_edges
    flip = (np.arange(n_edges) % 2).astype(bool)
    fwd = np.where(flip, twin[rep], rep)  # the half-edge running c0 -> c1
    c0, c1 = he_from[fwd], he_to[fwd]
    m.cellsOnEdge = np.stack([c0, c1], 1).astype(np.int32)
    # triangle left of c0->c1 is v1, right is v0 (v0 -> v1 = k x n)
    m.verticesOnEdge = np.stack([twin[fwd] // 3, fwd // 3], 1).astype(np.int32)
    edge_of_he = np.empty(n_he, np.int64)
    edge_of_he[fwd] = np.arange(n_edges)
    edge_of_he[twin[fwd]] = np.arange(n_edges)

    m.cellsOnVertex = tri.astype(np.int32)
    m.edgesOnVertex = edge_of_he.reshape(n_vert, 3).astype(np.int32)

    # CCW ring around each cell: next(a->b in t) = a->c, c = third corner of t
    p = he % 3
    t = he // 3
    prev_in_tri = 3 * t + (p + 2) % 3          # c -> a
    nxt = twin[prev_in_tri]                    # a -> c
    # start ring at the half-edge towards the smallest neighbour id
    o2 = np.lexsort((he_to, he_from))
    first = np.ones(n_he, bool)
    first[1:] = he_from[o2][1:] != he_from[o2][:-1]
    start = np.full(n_cells, -1, np.int64)
    start[he_from[o2][first]] = o2[first]
    if np.any(start < 0):
        raise ValueError("cell without triangles")
    coc = np.full((n_cells, max_edges), -1, np.int32)
    eoc = np.full((n_cells, max_edges), -1, np.int32)
    voc = np.full((n_cells, max_edges), -1, np.int32)
    neoc = np.zeros(n_cells, np.int32)
    cur = start.copy()
    alive = np.ones(n_cells, bool)
    for j in range(max_edges + 1):
        idx = np.nonzero(alive)[0]
        if idx.size == 0:
            break
        if j == max_edges:
            raise ValueError("cell with more than maxEdges edges")
        h = cur[idx]
        coc[idx, j] = he_to[h]
        eoc[idx, j] = edge_of_he[h]
        voc[idx, j] = h // 3
        neoc[idx] = j + 1
        h2 = nxt[h]
        cur[idx] = h2
        alive[idx] = h2 != start[idx]
    m.cellsOnCell, m.edgesOnCell, m.verticesOnCell, m.nEdgesOnCell = coc, eoc, voc, neoc
    m.maxEdges2 = 2 * max_edges
    return m


def _trisk(m: Mesh) -> None:
    """edgesOnEdge / weightsOnEdge / nEdgesOnEdge from kite areas.

    w(e,e') = side * (1/2 - sum_v R(cell,v)) * n(e',cell) * dv(e') / dc(e),
    R = kite/areaCell summed over the vertices passed walking CCW from e to
    e', side = +1 for c0 / -1 for c1, n(e',cell) = +1 when e' points out of
    the cell.  Reconstructs u.t with t = k x n (the direction the reference's
    TangentialReconOnEdge test compares with,
    components/omega/test/ocn/OceanTestCommon.h:221-226).
    """
    n_edges, me = m.nEdges, m.maxEdges
    eoe = np.full((n_edges, 2 * me), -1, np.int32)
    woe = np.zeros((n_edges, 2 * me))
    neoe = np.zeros(n_edges, np.int32)
    # kite fraction R[cell, j] for verticesOnCell[cell, j]
    voc, eoc, nn = m.verticesOnCell, m.edgesOnCell, m.nEdgesOnCell
    cov = m.cellsOnVertex
    vsafe = np.where(voc >= 0, voc, 0)
    which = (cov[vsafe] == np.arange(m.nCells)[:, None, None]).argmax(-1)
    R = np.take_along_axis(m.kiteAreasOnVertex[vsafe], which[..., None], -1)[..., 0]
    R = np.where(voc >= 0, R, 0.0) / m.areaCell[:, None]
    base = np.zeros(n_edges, np.int64)
    e_ids = np.arange(n_edges)
    for side in (0, 1):
        cell = m.cellsOnEdge[:, side].astype(np.int64)
        there = cell >= 0                       # a coastal edge has one cell only (cull_cells)
        cell = np.where(there, cell, 0)
        n = np.where(there, nn[cell], 1).astype(np.int64)
        pos = (eoc[cell] == e_ids[:, None]).argmax(1)
        sum_r = np.zeros(n_edges)
        sgn = 1.0 if side == 0 else -1.0
        for j in range(1, me):
            act = (j < n) & there
            idx = (pos + j) % n
            vidx = (pos + j - 1) % n
            sum_r = sum_r + R[cell, vidx]
            e2 = eoc[cell, idx].astype(np.int64)
            e2s = np.where(act, e2, 0)
            nsign = np.where(m.cellsOnEdge[e2s, 0] == cell, 1.0, -1.0)
            w = sgn * (0.5 - sum_r) * nsign * m.dvEdge[e2s] / m.dcEdge
            slot = base + j - 1
            a = np.nonzero(act)[0]
            eoe[a, slot[a]] = e2[a]
            woe[a, slot[a]] = w[a]
        base = base + n - 1
    neoe[:] = base
    m.edgesOnEdge, m.weightsOnEdge, m.nEdgesOnEdge = eoe, woe, neoe


# --------------------------------------------------------------------------
# land: remove cells, as the MPAS cell culler does for ocean meshes
# --------------------------------------------------------------------------
def cull_cells(m: Mesh, keep) -> Mesh:
    """The mesh without the cells where `keep` is False (land), in the conventions ocean meshes come in:

    * an edge stays when one of its cells stays, a vertex when one of its cells stays; everything is renumbered in
      the old order;
    * references to removed elements become -1 IN PLACE (0 in the file): cellsOnCell keeps its alignment with
      edgesOnCell, cellsOnVertex with kiteAreasOnVertex (whose values all stay: the geometry did not change);
    * a coastal edge lists its ocean cell FIRST -- the owner rule of the reference's Decomp ("the first cell ID
      associated with an edge ... is assumed to own the edge", src/base/Decomp.h:96-99) needs it; where that swaps
      the cells, the vertices swap and angleEdge turns by pi, so v0 -> v1 = k x n still holds;
    * edgesOnEdge / weightsOnEdge are rebuilt from the kites of the cells that are left (a coastal edge has the
      other edges of its one cell), nEdgesOnEdge counts them.
    The reference reads such a mesh as it reads any (HorzMesh.cpp:262-525); the coast shows up as boundary-slot
    entries in the local connectivity and EdgeMask = 0 on coastal edges (HorzMesh.cpp:590-598)."""
    keep = np.asarray(keep, bool)
    if keep.shape != (m.nCells,) or not keep.any():
        raise ValueError("keep: one flag per cell, at least one cell left")

    def renumber(flags):
        mp = np.full(flags.size + 1, -1, np.int64)      # index -1 ("none") maps to -1
        mp[:-1][flags] = np.arange(int(flags.sum()))
        return mp

    coe, cov = m.cellsOnEdge.astype(np.int64), m.cellsOnVertex.astype(np.int64)
    ekeep = keep[coe].any(1)
    vkeep = keep[cov].any(1)
    cmap, emap, vmap = renumber(keep), renumber(ekeep), renumber(vkeep)
    o = Mesh()
    o.geometry, o.sphere_radius = m.geometry, m.sphere_radius
    for k in ("Lx", "Ly", "dc"):
        if k in m:
            o[k] = m[k]
    o.nCells, o.nEdges, o.nVertices = int(keep.sum()), int(ekeep.sum()), int(vkeep.sum())
    o.maxEdges, o.maxEdges2, o.vertexDegree = m.maxEdges, m.maxEdges2, m.vertexDegree
    i32 = lambda a: a.astype(np.int32)
    o.cellsOnCell, o.edgesOnCell = i32(cmap[m.cellsOnCell[keep]]), i32(emap[m.edgesOnCell[keep]])
    o.verticesOnCell, o.nEdgesOnCell = i32(vmap[m.verticesOnCell[keep]]), m.nEdgesOnCell[keep].copy()
    o.cellsOnVertex, o.edgesOnVertex = i32(cmap[cov[vkeep]]), i32(emap[m.edgesOnVertex[vkeep]])
    ce, ve = cmap[coe[ekeep]], vmap[m.verticesOnEdge[ekeep]]
    swap = ce[:, 0] < 0
    ce[swap], ve[swap] = ce[swap][:, ::-1], ve[swap][:, ::-1]
    o.cellsOnEdge, o.verticesOnEdge = i32(ce), i32(ve)
    ang = m.angleEdge[ekeep].copy()
    ang[swap] = np.where(ang[swap] > 0, ang[swap] - np.pi, ang[swap] + np.pi)
    o.angleEdge = ang
    for nm in ("areaCell", "bottomDepth", "meshDensity", "fCell", "xCell", "yCell", "zCell", "lonCell", "latCell"):
        o[nm] = m[nm][keep].copy()
    for nm in ("dcEdge", "dvEdge", "fEdge", "xEdge", "yEdge", "zEdge", "lonEdge", "latEdge"):
        o[nm] = m[nm][ekeep].copy()
    for nm in ("areaTriangle", "fVertex", "xVertex", "yVertex", "zVertex", "lonVertex", "latVertex", "kiteAreasOnVertex"):
        o[nm] = m[nm][vkeep].copy()
    _trisk(o)
    o.culled_from = dict(cells=np.nonzero(keep)[0], edges=np.nonzero(ekeep)[0], vertices=np.nonzero(vkeep)[0], swapped=swap)
    return o


# --------------------------------------------------------------------------
# planar doubly periodic hex
# --------------------------------------------------------------------------
def planar_hex(nx: int, ny: int, dc: float, f0: float = 1.0e-4,
               bottom_depth: float = 1000.0, x0: float = 0.0, y0: float = 0.0) -> Mesh:
    """nx x ny doubly periodic hexagon mesh, cell spacing dc (ny even).

    Row j is shifted by dc/2 on odd rows; Lx = nx*dc, Ly = ny*dc*sqrt(3)/2
    (the 48x48, Lx=1, Ly=sqrt(3)/2 fixture of
    components/omega/test/ocn/TendencyTermsTest.cpp:40-41 is nx=ny=48, dc=1/48).
    """
    if ny % 2:
        raise ValueError("ny must be even for periodicity")
    n_cells = nx * ny
    jj, ii = np.meshgrid(np.arange(ny), np.arange(nx), indexing="ij")
    cid = lambda i, j: (j % ny) * nx + (i % nx)
    odd = jj % 2
    # upper neighbours of (i,j): NW = (i-1+odd, j+1), NE = (i+odd, j+1)
    a = cid(ii, jj)
    e = cid(ii + 1, jj)
    ne = cid(ii + odd, jj + 1)
    nw = cid(ii - 1 + odd, jj + 1)
    tri = np.concatenate([np.stack([a, e, ne], -1).reshape(-1, 3),
                          np.stack([a, ne, nw], -1).reshape(-1, 3)])
    m = _connectivity(tri, n_cells, 6)
    Lx, Ly = nx * dc, ny * dc * np.sqrt(3.0) / 2.0
    m.geometry, m.Lx, m.Ly, m.dc = "planar", Lx, Ly, dc
    xc = (dc * (ii + 0.5 * odd)).reshape(-1) + x0
    yc = (dc * np.sqrt(3.0) / 2.0 * jj).reshape(-1) + y0

    def mind(d, L):
        return d - L * np.round(d / L)

    c0, c1 = m.cellsOnEdge[:, 0], m.cellsOnEdge[:, 1]
    dx, dy = mind(xc[c1] - xc[c0], Lx), mind(yc[c1] - yc[c0], Ly)
    m.dcEdge = np.hypot(dx, dy)
    m.angleEdge = np.arctan2(dy, dx)
    xe, ye = xc[c0] + 0.5 * dx, yc[c0] + 0.5 * dy
    t = m.cellsOnVertex
    dxb, dyb = mind(xc[t[:, 1]] - xc[t[:, 0]], Lx), mind(yc[t[:, 1]] - yc[t[:, 0]], Ly)
    dxc, dyc = mind(xc[t[:, 2]] - xc[t[:, 0]], Lx), mind(yc[t[:, 2]] - yc[t[:, 0]], Ly)
    xv, yv = xc[t[:, 0]] + (dxb + dxc) / 3.0, yc[t[:, 0]] + (dyb + dyc) / 3.0
    v0, v1 = m.verticesOnEdge[:, 0], m.verticesOnEdge[:, 1]
    m.dvEdge = np.hypot(mind(xv[v1] - xv[v0], Lx), mind(yv[v1] - yv[v0], Ly))
    m.areaTriangle = 0.5 * np.abs(dxb * dyc - dyb * dxc)
    m.kiteAreasOnVertex = np.repeat(m.areaTriangle[:, None] / 3.0, 3, 1)
    m.areaCell = np.full(n_cells, dc * dc * np.sqrt(3.0) / 2.0)
    wrap = lambda v, L: v  # coordinates kept unwrapped-in-row; fields are periodic
    for nm, x, y in (("Cell", xc, yc), ("Edge", xe, ye), ("Vertex", xv, yv)):
        m["x" + nm], m["y" + nm] = wrap(x, Lx), wrap(y, Ly)
        m["z" + nm] = np.zeros_like(x)
        m["lon" + nm], m["lat" + nm] = np.zeros_like(x), np.zeros_like(x)
    m.fCell, m.fEdge, m.fVertex = (np.full(n_cells, f0), np.full(m.nEdges, f0),
                                   np.full(m.nVertices, f0))
    m.bottomDepth = np.full(n_cells, bottom_depth)
    m.meshDensity = np.ones(n_cells)
    m.sphere_radius = 0.0
    _trisk(m)
    return m


# --------------------------------------------------------------------------
# icosahedral sphere
# --------------------------------------------------------------------------
def _icosahedron():
    phi = (1.0 + np.sqrt(5.0)) / 2.0
    v = np.array([[-1, phi, 0], [1, phi, 0], [-1, -phi, 0], [1, -phi, 0],
                  [0, -1, phi], [0, 1, phi], [0, -1, -phi], [0, 1, -phi],
                  [phi, 0, -1], [phi, 0, 1], [-phi, 0, -1], [-phi, 0, 1]], float)
    v /= np.linalg.norm(v, axis=1)[:, None]
    f = np.array([[0, 11, 5], [0, 5, 1], [0, 1, 7], [0, 7, 10], [0, 10, 11],
                  [1, 5, 9], [5, 11, 4], [11, 10, 2], [10, 7, 6], [7, 1, 8],
                  [3, 9, 4], [3, 4, 2], [3, 2, 6], [3, 6, 8], [3, 8, 9],
                  [4, 9, 5], [2, 4, 11], [6, 2, 10], [8, 6, 7], [9, 8, 1]], np.int64)
    # rotate a little so no cell centre sits exactly on a pole / date line
    a, b = 0.3, 0.2
    rz = np.array([[np.cos(a), -np.sin(a), 0], [np.sin(a), np.cos(a), 0], [0, 0, 1]])
    rx = np.array([[1, 0, 0], [0, np.cos(b), -np.sin(b)], [0, np.sin(b), np.cos(b)]])
    v = v @ (rz @ rx).T
    return v, f


def _subdivide(v, f):
    n = v.shape[0]
    e = np.concatenate([f[:, [0, 1]], f[:, [1, 2]], f[:, [2, 0]]])
    lo, hi = e.min(1), e.max(1)
    key = lo * n + hi
    uk, inv = np.unique(key, return_inverse=True)
    mid = v[uk // n] + v[uk % n]
    mid /= np.linalg.norm(mid, axis=1)[:, None]
    nf = f.shape[0]
    m01, m12, m20 = n + inv[:nf], n + inv[nf:2 * nf], n + inv[2 * nf:]
    v2 = np.concatenate([v, mid])
    f2 = np.concatenate([np.stack([f[:, 0], m01, m20], 1),
                         np.stack([f[:, 1], m12, m01], 1),
                         np.stack([f[:, 2], m20, m12], 1),
                         np.stack([m01, m12, m20], 1)])
    return v2, f2


def _tri_area_sphere(a, b, c):
    """Area of unit-sphere triangles (Van Oosterom & Strackee)."""
    num = np.einsum("ij,ij->i", a, np.cross(b, c))
    den = 1.0 + np.einsum("ij,ij->i", a, b) + np.einsum("ij,ij->i", b, c) + np.einsum("ij,ij->i", c, a)
    return 2.0 * np.arctan2(np.abs(num), den)


def _arc(a, b):
    return 2.0 * np.arcsin(np.clip(0.5 * np.linalg.norm(a - b, axis=1), 0, 1))


def icosahedral(level: int, radius: float = SPHERE_RADIUS, bottom_depth: float = 5960.0) -> Mesh:
    """Icosahedral bisection mesh: 10*4**level+2 Voronoi cells (12 pentagons)."""
    v, f = _icosahedron()
    for _ in range(level):
        v, f = _subdivide(v, f)
    # orient every triangle CCW seen from outside
    nrm = np.cross(v[f[:, 1]] - v[f[:, 0]], v[f[:, 2]] - v[f[:, 0]])
    bad = np.einsum("ij,ij->i", nrm, v[f[:, 0]]) < 0
    f[bad] = f[bad][:, [0, 2, 1]]
    return _sphere_mesh(v, f, 6, radius, bottom_depth)


def icosahedral_with_heptagons(level: int, nflips: int = 6, radius: float = SPHERE_RADIUS, bottom_depth: float = 5960.0) -> Mesh:
    """The icosahedral mesh with `nflips` diagonals of its triangulation flipped: each flip turns two hexagons into
    pentagons and two into HEPTAGONS (maxEdges = 7, up to 11 edgesOnEdge), the cell shapes variable-resolution ocean
    meshes have.  The flipped pairs are no longer Delaunay (a vertex lands on a cell centre): geometrically poor,
    every array finite and consistent -- a connectivity test mesh, not one to measure convergence on."""
    v, f = _icosahedron()
    for _ in range(level):
        v, f = _subdivide(v, f)
    nrm = np.cross(v[f[:, 1]] - v[f[:, 0]], v[f[:, 2]] - v[f[:, 0]])
    bad = np.einsum("ij,ij->i", nrm, v[f[:, 0]]) < 0
    f[bad] = f[bad][:, [0, 2, 1]]
    f = f.copy()
    deg = np.bincount(f.reshape(-1), minlength=v.shape[0])
    touched = np.zeros(v.shape[0], bool)
    done = 0
    stride = next(p for p in (211, 223, 227, 229, 233) if f.shape[0] % p)      # spreads the flips over the sphere
    for t in (np.arange(f.shape[0]) * stride) % f.shape[0]:
        if done == nflips:
            break
        a, b, c = f[t]
        # the triangle on the other side of a -> b is (b, a, d)
        cand = np.nonzero((f == a).any(1) & (f == b).any(1))[0]
        other = [x for x in cand if x != t]
        if len(other) == 1:
            t2 = other[0]
            d = [x for x in f[t2] if x != a and x != b][0]
            quad = np.array([a, b, c, d])
            if (deg[quad] == 6).all() and not touched[quad].any():
                f[t], f[t2] = (a, d, c), (d, b, c)      # a, d, b, c is the quadrilateral in CCW order
                touched[quad] = True
                touched[f[(f == a).any(1) | (f == b).any(1) | (f == c).any(1) | (f == d).any(1)].reshape(-1)] = True
                done += 1
    if done < nflips:
        raise ValueError("not enough separate all-hexagon quadrilaterals for %d flips" % nflips)
    return _sphere_mesh(v, f, 7, radius, bottom_depth)


def _sphere_mesh(v, f, max_edges, radius, bottom_depth) -> Mesh:
    """Voronoi-style mesh on the sphere from a closed CCW triangulation f of the unit vectors v (cell centres)."""
    # spatially coherent file order (z-band then longitude) is NOT applied:
    # file order is subdivision order, the device layer does its own SFC sort
    n_cells = v.shape[0]
    m = _connectivity(f, n_cells, max_edges)
    m.geometry, m.sphere_radius = "sphere", radius
    pc = v
    t = m.cellsOnVertex
    cc = np.cross(pc[t[:, 1]] - pc[t[:, 0]], pc[t[:, 2]] - pc[t[:, 0]])
    pv = cc / np.linalg.norm(cc, axis=1)[:, None]        # circumcentres
    c0, c1 = m.cellsOnEdge[:, 0], m.cellsOnEdge[:, 1]
    pe = pc[c0] + pc[c1]
    pe /= np.linalg.norm(pe, axis=1)[:, None]
    v0, v1 = m.verticesOnEdge[:, 0], m.verticesOnEdge[:, 1]
    m.dcEdge = radius * _arc(pc[c0], pc[c1])
    m.dvEdge = radius * _arc(pv[v0], pv[v1])
    # kites: cell corner j of triangle v, with the midpoints of its two
    # Delaunay edges (edgesOnVertex[v, j] joins corner j and j+1)
    eov = m.edgesOnVertex
    kite = np.empty((m.nVertices, 3))
    for j in range(3):
        a = pc[t[:, j]]
        m_next = pe[eov[:, j]]             # between corner j and j+1
        m_prev = pe[eov[:, (j + 2) % 3]]   # between corner j-1 and j
        kite[:, j] = _tri_area_sphere(a, m_next, pv) + _tri_area_sphere(a, pv, m_prev)
    kite *= radius * radius
    m.kiteAreasOnVertex = kite
    m.areaTriangle = kite.sum(1)
    area = np.zeros(n_cells)
    np.add.at(area, t.reshape(-1), kite.reshape(-1))
    m.areaCell = area
    # angle between the edge normal (c0 -> c1 tangent at the edge point) and east
    nvec = pc[c1] - pc[c0]
    nvec -= np.einsum("ij,ij->i", nvec, pe)[:, None] * pe
    nvec /= np.linalg.norm(nvec, axis=1)[:, None]
    lon_e = np.arctan2(pe[:, 1], pe[:, 0])
    lat_e = np.arcsin(np.clip(pe[:, 2], -1, 1))
    east = np.stack([-np.sin(lon_e), np.cos(lon_e), np.zeros_like(lon_e)], 1)
    north = np.stack([-np.sin(lat_e) * np.cos(lon_e), -np.sin(lat_e) * np.sin(lon_e), np.cos(lat_e)], 1)
    m.angleEdge = np.arctan2(np.einsum("ij,ij->i", nvec, north), np.einsum("ij,ij->i", nvec, east))
    for nm, p in (("Cell", pc), ("Edge", pe), ("Vertex", pv)):
        m["x" + nm], m["y" + nm], m["z" + nm] = radius * p[:, 0], radius * p[:, 1], radius * p[:, 2]
        lon = np.arctan2(p[:, 1], p[:, 0])
        m["lon" + nm] = np.where(lon < 0, lon + 2 * np.pi, lon)
        m["lat" + nm] = np.arcsin(np.clip(p[:, 2], -1, 1))
    for nm in ("Cell", "Edge", "Vertex"):
        m["f" + nm] = 2.0 * OMEGA_EARTH * np.sin(m["lat" + nm])
    m.bottomDepth = np.full(n_cells, bottom_depth)
    m.meshDensity = np.ones(n_cells)
    _trisk(m)
    return m


# --------------------------------------------------------------------------
# NetCDF-3 (64-bit offset) writer / reader in the reference's file schema
# --------------------------------------------------------------------------
_INT_VARS = {"cellsOnCell": ("nCells", "maxEdges"), "edgesOnCell": ("nCells", "maxEdges"),
             "verticesOnCell": ("nCells", "maxEdges"), "nEdgesOnCell": ("nCells",),
             "cellsOnEdge": ("nEdges", "TWO"), "verticesOnEdge": ("nEdges", "TWO"),
             "edgesOnEdge": ("nEdges", "maxEdges2"), "nEdgesOnEdge": ("nEdges",),
             "cellsOnVertex": ("nVertices", "vertexDegree"),
             "edgesOnVertex": ("nVertices", "vertexDegree")}
_REAL_VARS = {"areaCell": "nCells", "bottomDepth": "nCells", "meshDensity": "nCells", "fCell": "nCells",
              "areaTriangle": "nVertices", "fVertex": "nVertices",
              "dvEdge": "nEdges", "dcEdge": "nEdges", "angleEdge": "nEdges", "fEdge": "nEdges"}


def write_netcdf(m: Mesh, path: str, state: dict | None = None) -> None:
    """Write mesh (+ optional state dict of (nElem, nVertLevels) arrays)."""
    from scipy.io import netcdf_file
    with netcdf_file(path, "w", version=2) as f:
        for d in ("nCells", "nEdges", "nVertices", "maxEdges", "maxEdges2", "vertexDegree"):
            f.createDimension(d, int(m[d]))
        f.createDimension("TWO", 2)
        f.on_a_sphere = b"YES" if m.geometry == "sphere" else b"NO"
        f.sphere_radius = float(m.sphere_radius)
        for nm, dims in _INT_VARS.items():
            a = m[nm]
            var = f.createVariable(nm, "i4", dims)
            var[:] = a if nm.startswith("nEdges") else a + 1   # 1-based, 0 = none
        for nm, dim in _REAL_VARS.items():
            f.createVariable(nm, "f8", (dim,))[:] = m[nm]
        for el, dim in (("Cell", "nCells"), ("Edge", "nEdges"), ("Vertex", "nVertices")):
            for c in ("x", "y", "z", "lon", "lat"):
                f.createVariable(c + el, "f8", (dim,))[:] = m[c + el]
        f.createVariable("kiteAreasOnVertex", "f8", ("nVertices", "vertexDegree"))[:] = m.kiteAreasOnVertex
        f.createVariable("weightsOnEdge", "f8", ("nEdges", "maxEdges2"))[:] = m.weightsOnEdge
        if state:
            nlev = next(iter(state.values())).shape[1]
            f.createDimension("nVertLevels", nlev)
            for nm, a in state.items():
                dim = "nEdges" if a.shape[0] == m.nEdges else "nCells"
                f.createVariable(nm, "f8", (dim, "nVertLevels"))[:] = a


def read_netcdf(path: str):
    """Inverse of write_netcdf: returns (Mesh, state dict)."""
    from scipy.io import netcdf_file
    m, state = Mesh(), {}
    with netcdf_file(path, "r", mmap=False) as f:
        for d in ("nCells", "nEdges", "nVertices", "maxEdges", "maxEdges2", "vertexDegree"):
            m[d] = int(f.dimensions[d])
        m.geometry = "sphere" if f.on_a_sphere == b"YES" else "planar"
        m.sphere_radius = float(f.sphere_radius)
        for nm, var in f.variables.items():
            a = np.array(var[:])
            if nm in _INT_VARS:
                m[nm] = (a if nm.startswith("nEdges") else a - 1).astype(np.int32)
            elif var.dimensions[-1] == "nVertLevels":
                state[nm] = a.astype(np.float64)
            else:
                m[nm] = a.astype(np.float64)
    return m, state
