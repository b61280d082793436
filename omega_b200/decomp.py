"""Domain decomposition: partition, halo layers, Omega local numbering.

Host-side restatement of the rules in the reference's
``components/omega/src/base/Decomp.cpp`` that define what "bit-exact
indexing" means for this path:

* cells: owned in ascending global ID, then ``HaloWidth`` layers each sorted
  ascending by global ID (Decomp.cpp:979-1100);
* edges / vertices: owner = task of the first valid cell in
  cellsOnEdge / cellsOnVertex (:1493-1502); owned ones numbered in
  first-encounter order walking EdgesOnCell of owned cells, non-owned ones
  around owned cells stored in REVERSE encounter order, then one block per
  cell-halo layer in encounter order (:1544-1609);
* every XxOnYy entry is translated to a local index, anything not local
  (or invalid) becomes the boundary slot N*All (:552-712);
* NEdgesOnCell counts valid entries (EdgesOnCell compacted, :2043-2064);
  NEdgesOnEdge counts entries that are valid or literally 0 (:2186-2200).

METIS is a third-party partitioner the reference calls with default options
(Decomp.cpp:968-971); results on owned elements are partition independent by
design (doc/design/OmegaV0ShallowWater.md:57-59) so any compact k-way
partition serves: ``partition_cells`` is a recursive coordinate bisection.
"""
from __future__ import annotations

import numpy as np

from .meshgen import Mesh

HALO_WIDTH = 3  # configs/Default.yml:15


def partition_cells(mesh: Mesh, ntasks: int) -> np.ndarray:
    """Recursive coordinate bisection of cell centres into ntasks parts."""
    pts = np.stack([mesh.xCell, mesh.yCell, mesh.zCell], 1)
    task = np.zeros(mesh.nCells, np.int32)

    def split(idx, t0, n):
        if n == 1:
            task[idx] = t0
            return
        n_left = n // 2
        p = pts[idx]
        ax = int(np.argmax(p.max(0) - p.min(0)))
        k = int(round(idx.size * n_left / n))
        order = np.argsort(p[:, ax], kind="stable")
        split(idx[order[:k]], t0, n_left)
        split(idx[order[k:]], t0 + n_left, n - n_left)

    split(np.arange(mesh.nCells), 0, ntasks)
    return task


def _encounter_order(seq: np.ndarray) -> np.ndarray:
    """Unique values of seq (>=0) in order of first appearance."""
    seq = seq[seq >= 0]
    uniq, first = np.unique(seq, return_index=True)
    return uniq[np.argsort(first, kind="stable")]


def _order_dual(x_on_cell_local_rows, n_owned_cells, cells_halo, owner_is_mine):
    """Local order of edges (or vertices) from the XxOnCell rows of the local
    cells (already in local cell order).  Returns (ids, n_owned, n_halo[])."""
    hw = len(cells_halo)
    around_owned = _encounter_order(x_on_cell_local_rows[:n_owned_cells].ravel())
    mine = owner_is_mine[around_owned]
    owned = around_owned[mine]
    halo1 = around_owned[~mine][::-1]           # stored from the end inward
    ids = [owned, halo1]
    seen = np.zeros(owner_is_mine.size, bool)
    seen[around_owned] = True
    n_halo = [around_owned.size]
    start = n_owned_cells
    count = around_owned.size
    for h in range(hw):
        end = cells_halo[h]
        enc = _encounter_order(x_on_cell_local_rows[start:end].ravel())
        enc = enc[~seen[enc]]
        seen[enc] = True
        ids.append(enc)
        count += enc.size
        start = end
        if h + 1 < hw:
            n_halo.append(count)
    return np.concatenate(ids), owned.size, np.array(n_halo, np.int32)


def build_local(mesh: Mesh, cell_task: np.ndarray, task: int, nlev: int,
                halo_width: int = HALO_WIDTH) -> Mesh:
    """Local (per-rank) mesh in Omega's local numbering, as HorzMesh holds it."""
    nC, nE, nV = mesh.nCells, mesh.nEdges, mesh.nVertices
    ntasks = int(cell_task.max()) + 1
    coc = mesh.cellsOnCell
    # ---- cells ----------------------------------------------------------
    owned = np.nonzero(cell_task == task)[0]            # ascending global id
    in_list = np.zeros(nC, bool)
    in_list[owned] = True
    layers = [owned]
    n_cells_halo = []
    prev = owned
    for _ in range(halo_width):
        nb = coc[prev].ravel()
        nb = nb[nb >= 0]
        nb = np.unique(nb[(cell_task[nb] != task) & ~in_list[nb]])
        in_list[nb] = True
        layers.append(nb)
        prev = nb
        n_cells_halo.append(sum(l.size for l in layers))
    cell_ids = np.concatenate(layers)
    n_cells_owned, n_cells_all = owned.size, cell_ids.size
    # address of every global cell on its owner
    loc_on_owner = np.zeros(nC, np.int64)
    for t in range(ntasks):
        idx = np.nonzero(cell_task == t)[0]
        loc_on_owner[idx] = np.arange(idx.size)

    # ---- edges / vertices -------------------------------------------------
    def first_valid_owner(cells_on_x):
        c = cells_on_x.copy()
        big = np.where(c >= 0, np.arange(c.shape[1])[None, :], c.shape[1])
        j = big.argmin(1)
        first = c[np.arange(c.shape[0]), j]
        return cell_task[first]

    edge_owner = first_valid_owner(mesh.cellsOnEdge)
    vert_owner = first_valid_owner(mesh.cellsOnVertex)
    edge_ids, n_edges_owned, n_edges_halo = _order_dual(
        mesh.edgesOnCell[cell_ids], n_cells_owned, n_cells_halo, edge_owner == task)
    vert_ids, n_verts_owned, n_verts_halo = _order_dual(
        mesh.verticesOnCell[cell_ids], n_cells_owned, n_cells_halo, vert_owner == task)
    n_edges_all, n_verts_all = edge_ids.size, vert_ids.size

    L = Mesh()
    L.geometry, L.sphere_radius = mesh.geometry, mesh.sphere_radius
    for k in ("Lx", "Ly", "dc"):
        if k in mesh:
            L[k] = mesh[k]
    L.MyTask, L.NumTasks, L.HaloWidth = task, ntasks, halo_width
    L.NVertLayers = nlev
    L.NCellsOwned, L.NCellsAll, L.NCellsSize = n_cells_owned, n_cells_all, n_cells_all + 1
    L.NEdgesOwned, L.NEdgesAll, L.NEdgesSize = n_edges_owned, n_edges_all, n_edges_all + 1
    L.NVerticesOwned, L.NVerticesAll, L.NVerticesSize = n_verts_owned, n_verts_all, n_verts_all + 1
    L.NCellsHalo = np.array(n_cells_halo, np.int32)
    L.NEdgesHalo, L.NVerticesHalo = n_edges_halo, n_verts_halo
    L.MaxEdges, L.MaxEdges2, L.VertexDegree = mesh.maxEdges, mesh.maxEdges2, mesh.vertexDegree
    L.NCellsGlobal, L.NEdgesGlobal, L.NVerticesGlobal = nC, nE, nV

    def ids_padded(ids, nglob):
        return np.concatenate([ids + 1, [nglob + 1]]).astype(np.int32)  # 1-based IDs

    L.CellID, L.EdgeID, L.VertexID = ids_padded(cell_ids, nC), ids_padded(edge_ids, nE), ids_padded(vert_ids, nV)
    # the reference never writes the boundary-slot entry of CellID / CellLoc (Decomp.cpp:1081-1098: Kokkos leaves it
    # zero), unlike those of the edge and vertex arrays; tests/test_decomp_vs_reference.py compares every entry
    L.CellID[-1] = 0

    # owner address (task, local index on owner) -- CellLoc / EdgeLoc / VertexLoc
    def loc_table(ids, owner, n_all, owned_rank_of):
        loc = np.zeros((n_all + 1, 2), np.int32)
        loc[:n_all, 0] = owner[ids]
        loc[:n_all, 1] = owned_rank_of[ids]
        loc[n_all] = (task, n_all)
        return loc

    L.CellLoc = loc_table(cell_ids, cell_task, n_cells_all, loc_on_owner)
    L.CellLoc[-1] = 0
    L._edge_owner, L._vert_owner = edge_owner[edge_ids], vert_owner[vert_ids]

    # global -> local maps; everything else -> boundary slot
    def g2l(ids, nglob, n_all):
        m = np.full(nglob + 1, n_all, np.int32)   # last entry serves index -1
        m[ids] = np.arange(ids.size, dtype=np.int32)
        return m

    c_map, e_map, v_map = g2l(cell_ids, nC, n_cells_all), g2l(edge_ids, nE, n_edges_all), g2l(vert_ids, nV, n_verts_all)

    def translate(arr, rows, gmap, n_all_rows, pad_value):
        out = np.full((n_all_rows + 1, arr.shape[1]), pad_value, np.int32)
        out[:n_all_rows] = gmap[arr[rows]]      # -1 indexes the trailing boundary entry
        return out

    L.CellsOnCell = translate(mesh.cellsOnCell, cell_ids, c_map, n_cells_all, n_cells_all)
    L.EdgesOnCell = translate(mesh.edgesOnCell, cell_ids, e_map, n_cells_all, n_edges_all)
    L.VerticesOnCell = translate(mesh.verticesOnCell, cell_ids, v_map, n_cells_all, n_verts_all)
    L.NEdgesOnCell = np.concatenate([(mesh.edgesOnCell[cell_ids] >= 0).sum(1), [0]]).astype(np.int32)
    L.CellsOnEdge = translate(mesh.cellsOnEdge, edge_ids, c_map, n_edges_all, n_cells_all)
    L.VerticesOnEdge = translate(mesh.verticesOnEdge, edge_ids, v_map, n_edges_all, n_verts_all)
    L.EdgesOnEdge = translate(mesh.edgesOnEdge, edge_ids, e_map, n_edges_all, n_edges_all)
    # Decomp.cpp:2186-2200: a 0 in the file's edgesOnEdge is kept in place as a
    # boundary-slot entry AND counted, so a file padded with zeros (MPAS
    # convention, and what meshgen.write_netcdf emits) yields MaxEdges2 for
    # every local edge; the padded entries carry weight 0.
    L.NEdgesOnEdge = np.concatenate([np.full(n_edges_all, mesh.maxEdges2), [0]]).astype(np.int32)
    L.CellsOnVertex = translate(mesh.cellsOnVertex, vert_ids, c_map, n_verts_all, n_cells_all)
    L.EdgesOnVertex = translate(mesh.edgesOnVertex, vert_ids, e_map, n_verts_all, n_edges_all)

    def pad(a, ids):
        out = np.zeros((ids.size + 1,) + a.shape[1:], a.dtype)
        out[:ids.size] = a[ids]
        return out

    for nm, src in (("AreaCell", "areaCell"), ("BottomDepth", "bottomDepth"), ("XCell", "xCell"),
                    ("YCell", "yCell"), ("ZCell", "zCell"), ("LonCell", "lonCell"), ("LatCell", "latCell"),
                    ("FCell", "fCell")):
        L[nm] = pad(mesh[src], cell_ids)
    for nm, src in (("DcEdge", "dcEdge"), ("DvEdge", "dvEdge"), ("AngleEdge", "angleEdge"),
                    ("FEdge", "fEdge"), ("XEdge", "xEdge"), ("YEdge", "yEdge"), ("ZEdge", "zEdge"),
                    ("LonEdge", "lonEdge"), ("LatEdge", "latEdge"), ("WeightsOnEdge", "weightsOnEdge")):
        L[nm] = pad(mesh[src], edge_ids)
    for nm, src in (("AreaTriangle", "areaTriangle"), ("FVertex", "fVertex"), ("XVertex", "xVertex"),
                    ("YVertex", "yVertex"), ("ZVertex", "zVertex"), ("LonVertex", "lonVertex"),
                    ("LatVertex", "latVertex"), ("KiteAreasOnVertex", "kiteAreasOnVertex")):
        L[nm] = pad(mesh[src], vert_ids)
    L._cell_ids, L._edge_ids, L._vert_ids = cell_ids, edge_ids, vert_ids
    return L


def build_all(mesh: Mesh, ntasks: int, nlev: int, halo_width: int = HALO_WIDTH,
              cell_task: np.ndarray | None = None) -> list[Mesh]:
    """Local meshes of every task + EdgeLoc / VertexLoc owner addresses."""
    if cell_task is None:
        cell_task = partition_cells(mesh, ntasks) if ntasks > 1 else np.zeros(mesh.nCells, np.int32)
    locs = [build_local(mesh, cell_task, t, nlev, halo_width) for t in range(ntasks)]
    # address of each global edge / vertex on its owner (Decomp.cpp:1639-1676)
    for kind, nglob in (("Edge", mesh.nEdges), ("Vertex", mesh.nVertices)):
        addr = np.zeros(nglob, np.int64)
        for L in locs:
            n_own = L["N%ssOwned" % ("Edge" if kind == "Edge" else "Vertice")]
            ids = L["_edge_ids" if kind == "Edge" else "_vert_ids"]
            addr[ids[:n_own]] = np.arange(n_own)
        for L in locs:
            ids = L["_edge_ids" if kind == "Edge" else "_vert_ids"]
            n_all = ids.size
            own = L["_edge_owner" if kind == "Edge" else "_vert_owner"]
            loc = np.zeros((n_all + 1, 2), np.int32)
            loc[:n_all, 0], loc[:n_all, 1] = own, addr[ids]
            loc[n_all] = (L.MyTask, n_all)
            L[kind + "Loc"] = loc
    return locs


# ---------------------------------------------------------------------------
# The compiled Decomp (csrc/decomp.cpp, omega_decomp_* of include/omega_b200.h): ONE rank's partition,
# numbering, owner addresses and halo lists without building anybody else's mesh.  The numpy code above
# stays as its cross-check (tests/test_decomp_lib.py: identical indices) and as the N-rank emulation.
# ---------------------------------------------------------------------------
class GlobalMeshDesc:
    """ctypes view of OmegaGlobalMesh over a meshgen.Mesh (0-based, -1 = none)."""

    def __init__(self, mesh: Mesh):
        import ctypes as C
        from . import capi
        ip, dp = capi.c_int_p, capi.c_dbl_p

        class S(C.Structure):
            _fields_ = ([(n, C.c_int) for n in ("NCells", "NEdges", "NVertices", "MaxEdges", "MaxEdges2", "VertexDegree")]
                        + [(n, ip) for n in ("CellsOnCell", "EdgesOnCell", "VerticesOnCell", "CellsOnEdge", "VerticesOnEdge",
                                             "EdgesOnEdge", "CellsOnVertex", "EdgesOnVertex")]
                        + [(n, dp) for n in ("XCell", "YCell", "ZCell")])
        self.keep = []
        d = S(NCells=mesh.nCells, NEdges=mesh.nEdges, NVertices=mesh.nVertices, MaxEdges=mesh.maxEdges,
              MaxEdges2=mesh.maxEdges2, VertexDegree=mesh.vertexDegree)
        for n, src in (("CellsOnCell", "cellsOnCell"), ("EdgesOnCell", "edgesOnCell"), ("VerticesOnCell", "verticesOnCell"),
                       ("CellsOnEdge", "cellsOnEdge"), ("VerticesOnEdge", "verticesOnEdge"), ("EdgesOnEdge", "edgesOnEdge"),
                       ("CellsOnVertex", "cellsOnVertex"), ("EdgesOnVertex", "edgesOnVertex")):
            a = np.ascontiguousarray(mesh[src], np.int32)
            self.keep.append(a)
            setattr(d, n, a.ctypes.data_as(ip))
        for n, src in (("XCell", "xCell"), ("YCell", "yCell"), ("ZCell", "zCell")):
            a = np.ascontiguousarray(mesh[src], np.float64)
            self.keep.append(a)
            setattr(d, n, a.ctypes.data_as(dp))
        self.c = d


def _lib():
    from . import capi
    return capi.load_library()


def partition_cells_lib(mesh: Mesh, ntasks: int, method: int = 0):
    """(cell_task, edge_cut): 0 = recursive coordinate bisection (== partition_cells), 1 = + boundary refinement."""
    import ctypes as C
    lib, g = _lib(), GlobalMeshDesc(mesh)
    task = np.zeros(mesh.nCells, np.int32)
    cut = C.c_longlong()
    rc = lib.omega_decomp_partition(C.byref(g.c), ntasks, method, task.ctypes.data_as(C.POINTER(C.c_int)), C.byref(cut))
    if rc:
        raise RuntimeError(lib.omega_last_error().decode())
    return task, int(cut.value)


def build_local_lib(mesh: Mesh, cell_task: np.ndarray, task: int, nlev: int, halo_width: int = HALO_WIDTH):
    """(local mesh, halo lists) of ONE rank from the compiled Decomp; same contents as build_local +
    EdgeLoc / VertexLoc + parallel.build_halo_lists()[task]."""
    import ctypes as C
    lib, g = _lib(), GlobalMeshDesc(mesh)
    ip = C.POINTER(C.c_int)
    cell_task = np.ascontiguousarray(cell_task, np.int32)
    ntasks = int(cell_task.max()) + 1
    h = C.c_void_p()
    rc = lib.omega_decomp_create(C.byref(g.c), cell_task.ctypes.data_as(ip), ntasks, task, halo_width, C.byref(h))
    if rc:
        raise RuntimeError(lib.omega_last_error().decode())
    try:
        w = 2 + halo_width
        cnt = np.zeros(3 * w, np.int32)
        lib.omega_decomp_counts(h, cnt.ctypes.data_as(ip))
        cnt = cnt.reshape(3, w)
        ids, loc = [], []
        for k in range(3):
            a = np.zeros(cnt[k, 1], np.int32)
            lib.omega_decomp_ids(h, k, a.ctypes.data_as(ip))
            b = np.zeros((cnt[k, 1] + 1, 2), np.int32)
            lib.omega_decomp_loc(h, k, b.ctypes.data_as(ip))
            ids.append(a); loc.append(b)
        widths = (mesh.maxEdges, mesh.maxEdges, mesh.maxEdges, 2, 2, mesh.maxEdges2, mesh.vertexDegree, mesh.vertexDegree)
        rows = (0, 0, 0, 1, 1, 1, 2, 2)
        conn = []
        for which in range(8):
            a = np.zeros((cnt[rows[which], 1] + 1, widths[which]), np.int32)
            lib.omega_decomp_connectivity(h, which, a.ctypes.data_as(ip))
            conn.append(a)
        nn = C.c_int()
        lib.omega_decomp_halo_neighbors(h, C.byref(nn), None)
        nbrs = np.zeros(max(nn.value, 1), np.int32)
        lib.omega_decomp_halo_neighbors(h, C.byref(nn), nbrs.ctypes.data_as(ip))
        nbrs = [int(x) for x in nbrs[:nn.value]]
        lists = dict(neighbors=nbrs, send={}, recv={})
        for k, kind in enumerate(("Cell", "Edge", "Vertex")):
            so, ro = np.zeros(nn.value + 1, np.int32), np.zeros(nn.value + 1, np.int32)
            lib.omega_decomp_halo_lists(h, k, so.ctypes.data_as(ip), None, ro.ctypes.data_as(ip), None)
            si, ri = np.zeros(max(so[-1], 1), np.int32), np.zeros(max(ro[-1], 1), np.int32)
            lib.omega_decomp_halo_lists(h, k, None, si.ctypes.data_as(ip), None, ri.ctypes.data_as(ip))
            lists["send"][kind] = {t: si[so[i]:so[i + 1]].copy() for i, t in enumerate(nbrs) if so[i + 1] > so[i]}
            lists["recv"][kind] = {t: ri[ro[i]:ro[i + 1]].copy() for i, t in enumerate(nbrs) if ro[i + 1] > ro[i]}
    finally:
        lib.omega_decomp_destroy(h)
    cell_ids, edge_ids, vert_ids = (a.astype(np.int64) for a in ids)
    nC, nE, nV = mesh.nCells, mesh.nEdges, mesh.nVertices
    L = Mesh()
    L.geometry, L.sphere_radius = mesh.geometry, mesh.sphere_radius
    for k in ("Lx", "Ly", "dc"):
        if k in mesh:
            L[k] = mesh[k]
    L.MyTask, L.NumTasks, L.HaloWidth, L.NVertLayers = task, ntasks, halo_width, nlev
    for k, nm in enumerate(("Cells", "Edges", "Vertices")):
        L["N%sOwned" % nm], L["N%sAll" % nm], L["N%sSize" % nm] = int(cnt[k, 0]), int(cnt[k, 1]), int(cnt[k, 1]) + 1
        L["N%sHalo" % nm] = cnt[k, 2:].copy()
    L.MaxEdges, L.MaxEdges2, L.VertexDegree = mesh.maxEdges, mesh.maxEdges2, mesh.vertexDegree
    L.NCellsGlobal, L.NEdgesGlobal, L.NVerticesGlobal = nC, nE, nV
    pad = lambda i, n: np.concatenate([i + 1, [n + 1]]).astype(np.int32)
    L.CellID, L.EdgeID, L.VertexID = pad(cell_ids, nC), pad(edge_ids, nE), pad(vert_ids, nV)
    L.CellID[-1] = 0     # as the reference leaves it (see build_local)
    L.CellLoc, L.EdgeLoc, L.VertexLoc = loc
    (L.CellsOnCell, L.EdgesOnCell, L.VerticesOnCell, L.CellsOnEdge, L.VerticesOnEdge, L.EdgesOnEdge,
     L.CellsOnVertex, L.EdgesOnVertex) = conn
    L.NEdgesOnCell = np.concatenate([(mesh.edgesOnCell[cell_ids] >= 0).sum(1), [0]]).astype(np.int32)
    L.NEdgesOnEdge = np.concatenate([np.full(edge_ids.size, mesh.maxEdges2), [0]]).astype(np.int32)

    def padf(a, i):
        out = np.zeros((i.size + 1,) + a.shape[1:], a.dtype)
        out[:i.size] = a[i]
        return out

    for nm, src in (("AreaCell", "areaCell"), ("BottomDepth", "bottomDepth"), ("XCell", "xCell"), ("YCell", "yCell"),
                    ("ZCell", "zCell"), ("LonCell", "lonCell"), ("LatCell", "latCell"), ("FCell", "fCell")):
        L[nm] = padf(mesh[src], cell_ids)
    for nm, src in (("DcEdge", "dcEdge"), ("DvEdge", "dvEdge"), ("AngleEdge", "angleEdge"), ("FEdge", "fEdge"),
                    ("XEdge", "xEdge"), ("YEdge", "yEdge"), ("ZEdge", "zEdge"), ("LonEdge", "lonEdge"),
                    ("LatEdge", "latEdge"), ("WeightsOnEdge", "weightsOnEdge")):
        L[nm] = padf(mesh[src], edge_ids)
    for nm, src in (("AreaTriangle", "areaTriangle"), ("FVertex", "fVertex"), ("XVertex", "xVertex"), ("YVertex", "yVertex"),
                    ("ZVertex", "zVertex"), ("LonVertex", "lonVertex"), ("LatVertex", "latVertex"),
                    ("KiteAreasOnVertex", "kiteAreasOnVertex")):
        L[nm] = padf(mesh[src], vert_ids)
    L._cell_ids, L._edge_ids, L._vert_ids = cell_ids, edge_ids, vert_ids
    return L, lists
