"""Host-side multi-rank logic on CPU: decomposition, Omega local numbering,
exchange lists, phase-split stepping.  Criteria are the reference's own:
DecompTest (sum of owned IDs, test/base/DecompTest.cpp:126-136), HaloTest
(exact equality of every local element after an exchange of owned IDs,
test/base/HaloTest.cpp:41-99), and the design requirement that owned results
do not depend on the partition (doc/design/OmegaV0ShallowWater.md:57-59)."""
import os

import numpy as np
import pytest

from omega_b200 import analytic, capi, decomp, meshgen, parallel
from tests import cases
from tests.oracle_lib import Oracle


@pytest.fixture(scope="module")
def mesh():
    return cases.sphere_mesh(4)  # 2562 cells


@pytest.mark.parametrize("ntasks", [1, 2, 3, 8])
def test_decomposition_rules(mesh, ntasks):
    locs = decomp.build_all(mesh, ntasks, 4)
    tot = {"Cell": 0, "Edge": 0, "Vertex": 0}
    nglob = {"Cell": mesh.nCells, "Edge": mesh.nEdges, "Vertex": mesh.nVertices}
    for L in locs:
        # cells: owned ascending by global ID, each halo layer ascending
        assert np.all(np.diff(L.CellID[:L.NCellsOwned]) > 0)
        b = [L.NCellsOwned] + list(L.NCellsHalo)
        for lo, hi in zip(b[:-1], b[1:]):
            assert np.all(np.diff(L.CellID[lo:hi]) > 0)
        assert L.NCellsHalo[-1] == L.NCellsAll and L.NCellsSize == L.NCellsAll + 1
        for kind in ("Cell", "Edge", "Vertex"):
            n_owned = parallel._n(L, kind, "Owned")
            tot[kind] += int(L[kind + "ID"][:n_owned].astype(np.int64).sum())
            loc = L[kind + "Loc"]
            assert np.all(loc[:n_owned, 0] == L.MyTask) and np.array_equal(loc[:n_owned, 1], np.arange(n_owned))
        # every edge of every local cell is local; connectivity points into [0, NAll]
        ne = L.NEdgesOnCell[:L.NCellsAll]
        for j in range(L.MaxEdges):
            assert np.all(L.EdgesOnCell[:L.NCellsAll, j][ne > j] < L.NEdgesAll)
        for nm, hi in (("CellsOnEdge", L.NCellsAll), ("EdgesOnEdge", L.NEdgesAll), ("VerticesOnEdge", L.NVerticesAll),
                       ("CellsOnVertex", L.NCellsAll), ("EdgesOnVertex", L.NEdgesAll)):
            assert L[nm].min() >= 0 and L[nm].max() <= hi
        # edge owner = task of the first valid cell of cellsOnEdge (Decomp.cpp:1493-1502)
        ge = L.EdgeID[:L.NEdgesOwned] - 1
        assert np.all(decomp.partition_cells(mesh, ntasks)[mesh.cellsOnEdge[ge, 0]] == L.MyTask) or ntasks == 1
    for kind in tot:  # DecompTest criterion
        assert tot[kind] == nglob[kind] * (nglob[kind] + 1) // 2


def test_single_task_edge_order_is_first_encounter(mesh):
    L = decomp.build_all(mesh, 1, 1)[0]
    seen, order = set(), []
    for c in range(mesh.nCells):
        for e in mesh.edgesOnCell[c]:
            if e >= 0 and e not in seen:
                seen.add(e); order.append(e + 1)
    assert np.array_equal(L.EdgeID[:-1], np.array(order))
    assert np.array_equal(L.CellID[:-1], np.arange(1, mesh.nCells + 1))


@pytest.mark.parametrize("ntasks", [2, 5, 8])
def test_halo_exchange_lists_exact(mesh, ntasks):
    """HaloTest: owned elements carry their global ID, halos are poisoned;
    after the exchange every local element equals its global ID."""
    locs = decomp.build_all(mesh, ntasks, 3)
    lists = parallel.build_halo_lists(locs)
    for kind in ("Cell", "Edge", "Vertex"):
        arrays = []
        for L in locs:
            n_owned, n_size = parallel._n(L, kind, "Owned"), parallel._n(L, kind, "Size")
            a = np.full((2, n_size, 3), -1.0)
            ids = L[kind + "ID"].astype(np.float64)
            a[:, :n_owned, :] = ids[None, :n_owned, None] * np.array([1.0, 2.0])[:, None, None] + np.arange(3)[None, None, :]
            arrays.append(a)
        parallel.exchange_inprocess(lists, kind, arrays)
        for L, a in zip(locs, arrays):
            n_all = parallel._n(L, kind, "All")
            ids = L[kind + "ID"].astype(np.float64)
            exp = ids[None, :n_all, None] * np.array([1.0, 2.0])[:, None, None] + np.arange(3)[None, None, :]
            assert np.array_equal(a[:, :n_all], exp), kind
            assert np.all(a[:, n_all] == -1.0), "boundary slot must not be touched"
    for t, h in enumerate(lists):
        assert h["neighbors"] == sorted(h["neighbors"]) and t not in h["neighbors"]


def _multirank_oracle(mesh, ntasks, K, T, cfg, nsteps, dt, stepper, halo_width):
    locs = decomp.build_all(mesh, ntasks, K, halo_width)
    lists = parallel.build_halo_lists(locs)
    orcs = []
    for L in locs:
        h, u, tr = analytic.tc5(L, T)
        o = Oracle(L, cfg, T)
        o.make_state(h, u, tr)
        orcs.append(o)
    nph = 2 if stepper == capi.RK4 else 1
    for s in range(nsteps):
        co = capi.rk_coeffs(stepper, dt, t0=s * dt)
        for p in range(nph):
            for o in orcs:
                o.step_phase(stepper, co, p)
            per_rank = [o.exchange_arrays(stepper, p) for o in orcs]
            for i, (kind, _) in enumerate(per_rank[0]):
                parallel.exchange_inprocess(lists, kind, [pr[i][1] for pr in per_rank])
        for o in orcs:
            o.rotate()
    return locs, orcs


def partition_error(ntasks, stepper, halo_width, land=False, **cfgkw):
    """max over fields of rel-max difference between the N-rank emulation of
    the reference algorithm and its 1-rank run, over ALL local elements."""
    K, T, nsteps = 3, 2, 4
    mesh = cases.land_mesh(4) if land else cases.sphere_mesh(4)
    L1, cfg, (h, u, tr), dt = (cases.land_case if land else cases.sphere_case)(4, K, T, **cfgkw)
    ref = Oracle(L1, cfg, T)
    ref.make_state(h, u, tr)
    for s in range(nsteps):
        ref.do_step(stepper, capi.rk_coeffs(stepper, dt, t0=s * dt))
    gh, gu, gt = ref.current()
    cell_of, edge_of = L1.CellID[:-1] - 1, L1.EdgeID[:-1] - 1
    H = np.empty_like(gh[:-1]); H[cell_of] = gh[:-1]
    U = np.empty_like(gu[:-1]); U[edge_of] = gu[:-1]
    TR = np.empty_like(gt[:, :-1]); TR[:, cell_of] = gt[:, :-1]
    locs, orcs = _multirank_oracle(mesh, ntasks, K, T, cfg, nsteps, dt, stepper, halo_width)
    worst = 0.0
    for L, o in zip(locs, orcs):
        lh, lu, lt = o.current()
        ci, ei = L.CellID[:L.NCellsAll] - 1, L.EdgeID[:L.NEdgesAll] - 1
        for a, b in ((lh[:L.NCellsAll], H[ci]), (lu[:L.NEdgesAll], U[ei]), (lt[:, :L.NCellsAll], TR[:, ci])):
            worst = max(worst, cases.rel_max(a, b))
        o.close()
    ref.close()
    return worst


NO_DEL4 = dict(VelHyperDiffTendencyEnable=0, TracerHyperDiffTendencyEnable=0)


@pytest.mark.parametrize("ntasks", [2, 4])
@pytest.mark.parametrize("stepper,halo_width,cfgkw", [
    (capi.RK4, 3, NO_DEL4), (capi.RK2, 3, NO_DEL4), (capi.FB, 3, {}),
    (capi.RK4, 4, {}), (capi.RK2, 4, {})])
def test_results_do_not_depend_on_partition(ntasks, stepper, halo_width, cfgkw):
    """bit-for-bit partition independence (doc/design/OmegaV0ShallowWater.md:57-59)."""
    assert partition_error(ntasks, stepper, halo_width, **cfgkw) == 0.0


@pytest.mark.parametrize("ntasks", [2, 5])
def test_results_do_not_depend_on_partition_with_a_coastline(ntasks):
    """the same on the mesh with land (cases.land_mesh): partitions that cut the coast"""
    assert partition_error(ntasks, capi.RK4, 4, land=True) == 0.0


def test_reference_needs_halo_width_4_with_del4():
    """Finding, not a bug of this repo: with the biharmonic terms on, one RHS
    consumes two halo layers, RK4 evaluates two RHS between exchanges
    (src/timeStepping/RungeKutta4Stepper.cpp:107 "TODO ... depends on halo width"),
    so the reference algorithm with the default HaloWidth 3 is itself partition
    dependent at ~1e-6; HaloWidth >= 4 restores bit-exact independence."""
    err3 = partition_error(2, capi.RK4, 3)
    assert 0.0 < err3 < 1e-4


def _gloo_worker(rank, world, port, q):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    m = cases.sphere_mesh(3)
    L, lists = parallel.local_problem(m, world, rank, 2)
    ex = parallel.HostHaloExchanger(lists)
    ok = True
    for kind in ("Cell", "Edge", "Vertex"):
        n_owned, n_all, n_size = (parallel._n(L, kind, w) for w in ("Owned", "All", "Size"))
        ids = L[kind + "ID"].astype(np.float64)
        a = np.full((n_size, 2), -1.0)
        a[:n_owned] = ids[:n_owned, None] + np.array([0.0, 0.5])
        ex.exchange(a, kind)
        ok &= bool(np.array_equal(a[:n_all], ids[:n_all, None] + np.array([0.0, 0.5])))
    q.put((rank, ok))
    dist.destroy_process_group()


def test_gloo_two_process_exchange():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=_gloo_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
    assert all(ok for _, ok in res), res
