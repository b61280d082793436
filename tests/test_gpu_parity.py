"""GPU parity tests proper: the CUDA path, called through the C ABI, against
the CPU oracle on the same inputs.  Thresholds are BASELINE.json's: indices,
permutations and derived sign/mask arrays bit-exact; per-field tendencies
after one right-hand side <= 1e-12 relative max (max|a-b|/max|b|); prognostic
state after many steps <= 1e-10 relative L2."""
import numpy as np
import pytest

from omega_b200 import capi
from tests import cases
from tests.cases import rel_l2, rel_max

pytestmark = pytest.mark.gpu

TOL_TEND = 1e-12
TOL_STATE = 1e-10


def gpu_objects(L, cfg, T, reorder=True):
    from omega_b200 import model
    mesh = model.HorzMesh(L, reorder=reorder)
    return mesh, model.Stepper(mesh, cfg, T), model.OceanState(mesh, T)


CASES = {
    "sphere3_K4_T2": lambda: cases.sphere_case(3, 4, 2),
    "sphere2_K3_T1": lambda: cases.sphere_case(2, 3, 1),      # odd level count -> scalar path
    "sphere3_K64_T3": lambda: cases.sphere_case(3, 64, 3),    # warp == one element row
    "plane16_K1_T0": lambda: cases.plane_case(16, 1, 0),      # config-1 shape: 1 level, no tracers
    "plane12_K10_T2": lambda: cases.plane_case(12, 10, 2),
    # K % 16 == 0 -> the shared-memory tiled cell pass (fused_tile.cu); 41 tiles incl. the 12 pentagons
    "sphere4_K16_T3": lambda: cases.sphere_case(4, 16, 3),
    "plane16_K32_T2": lambda: cases.plane_case(16, 32, 2),
    "sphere3_K16_T1": lambda: cases.sphere_case(3, 16, 1),    # one tracer: the short copy schedule of k_fCT2
    # tracer hyperdiffusion off -> the tiled pass without its Del2 phase
    "sphere3_K16_T2_nodel4": lambda: cases.sphere_case(3, 16, 2, TracerHyperDiffTendencyEnable=0),
    # shapes of BASELINE configs 5 and 4 on a small mesh: 32 tracers walk the four rotating tracer buffers of
    # k_fCT2 eight times per chunk; 80 levels are five chunks (copy stream index it0 = chunk * T)
    "sphere3_K64_T32": lambda: cases.sphere_case(3, 64, 32),
    "sphere3_K80_T2": lambda: cases.sphere_case(3, 80, 2),
    # level counts that are not multiples of 16: the last chunk of the tile kernels is partial (masked lanes).
    # 60 = 3 chunks + 12 levels, 24 = 1 chunk + 8 levels; K = 4 and K = 10 above are a single partial chunk
    "sphere3_K60_T2": lambda: cases.sphere_case(3, 60, 2),
    "sphere3_K24_T3": lambda: cases.sphere_case(3, 24, 3),
    # a coastline (cases.land_mesh): coastal edges with one cell and EdgeMask = 0, vertices with missing cells and
    # edges, cells with land neighbours -- boundary-slot reads that change OWNED results, which a closed mesh never has
    "land3_K4_T2": lambda: cases.land_case(3, 4, 2),
    "land4_K16_T3": lambda: cases.land_case(4, 16, 3),
    "land3_K3_T1": lambda: cases.land_case(3, 3, 1),
    # heptagons (cases.hept_mesh: maxEdges = 7, rows of 11 and 12 edgesOnEdge): the tile kernels are written for at
    # most six edges per cell and stand down (mesh.cu: MaxEdgesUsed > 6), the general fused kernels step
    "hept3_K4_T2": lambda: cases.hept_case(3, 4, 2),
    "hept3_K16_T3": lambda: cases.hept_case(3, 16, 3),
}
TILED = {"sphere4_K16_T3", "plane16_K32_T2", "sphere3_K16_T1", "sphere3_K16_T2_nodel4", "sphere3_K64_T32",
         "sphere3_K80_T2", "sphere3_K64_T3", "sphere3_K60_T2", "sphere3_K24_T3", "sphere3_K4_T2", "plane12_K10_T2",
         "land3_K4_T2", "land4_K16_T3"}


@pytest.fixture(scope="module", params=list(CASES))
def case(request):
    return request.param, CASES[request.param]()


@pytest.mark.parametrize("xfer", ["full", "slab", "zerocopy"])
def test_field_transfer_modes_bit_exact(monkeypatch, xfer):
    """omega_field_to_device / to_host with the permutation applied, all three transfer paths (full-size staging,
    64 MB slabs, kernel access to pinned host memory), pageable and pinned host arrays, several slabs per field."""
    import torch
    monkeypatch.setenv("OMEGA_XFER", xfer)
    L, cfg, (h, u, tr), dt = cases.sphere_case(5, 64, 12)      # tracers: 12 x 10 243 x 64 x 8 B = 63 MB -> slabs
    big = np.ascontiguousarray(np.repeat(tr, 2, axis=0))       # 126 MB: the last slab is a partial one
    mesh, st, state = gpu_objects(L, cfg, 0)
    p, _ = mesh.perm(capi.ON_CELL)
    for host in (big, u):
        kind = capi.ON_CELL if host.ndim == 3 else capi.ON_EDGE
        d = mesh.to_device(kind, host)
        pe, _ = mesh.perm(kind)
        assert np.array_equal(d.cpu().numpy(), host[..., pe, :]), "device order must be host[perm]"
        assert np.array_equal(mesh.to_host(kind, d), host)
        src = torch.from_numpy(host).pin_memory()
        dst = torch.empty(host.shape, dtype=torch.float64).pin_memory()
        d2 = mesh.to_device(kind, src.numpy())
        mesh.to_host(kind, d2, dst.numpy(), sync=False)
        torch.cuda.current_stream().synchronize()
        assert np.array_equal(dst.numpy(), host)
    st.close(); mesh.close()


@pytest.mark.parametrize("reorder", [True, False])
def test_permutation_and_transfer_bit_exact(case, reorder):
    _, (L, cfg, (h, u, tr), dt) = case
    mesh, st, state = gpu_objects(L, cfg, tr.shape[0], reorder)
    for kind in range(3):
        p, ip = mesh.perm(kind)
        n = mesh.NSize[kind]
        assert np.array_equal(np.sort(p), np.arange(n)), "perm is not a bijection"
        assert np.array_equal(ip[p], np.arange(n))
        assert p[-1] == n - 1, "boundary slot must stay last"
        own = mesh.NOwned[kind]
        assert np.array_equal(np.sort(p[:own]), np.arange(own)), "owned block must stay first"
    for kind, a in ((capi.ON_CELL, h), (capi.ON_EDGE, u)):
        d = mesh.to_device(kind, a)
        p, _ = mesh.perm(kind)
        assert np.array_equal(d.cpu().numpy(), a[p]), "device order must be host[perm]"
        assert np.array_equal(mesh.to_host(kind, d), a)
    if tr.shape[0]:
        d = mesh.to_device(capi.ON_CELL, tr)
        assert np.array_equal(mesh.to_host(capi.ON_CELL, d), tr)
    # omega_field_to_host_async: pinned destination, ordered by the stream, on a side stream
    import torch
    pin = torch.empty(u.shape, dtype=torch.float64).pin_memory()
    d = mesh.to_device(capi.ON_EDGE, u)
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        mesh.to_host(capi.ON_EDGE, d, pin.numpy(), sync=False)
    side.synchronize()
    assert np.array_equal(pin.numpy(), u)
    # pinned SOURCE: the library reads it over PCIe from the kernel itself (no staging), permutation applied
    src = torch.from_numpy(np.ascontiguousarray(u)).pin_memory()
    d2 = mesh.to_device(capi.ON_EDGE, src.numpy())
    assert np.array_equal(mesh.to_host(capi.ON_EDGE, d2), u)
    if tr.shape[0]:
        srct = torch.from_numpy(np.ascontiguousarray(tr)).pin_memory()
        pint = torch.empty(tr.shape, dtype=torch.float64).pin_memory()
        dt_ = mesh.to_device(capi.ON_CELL, srct.numpy())
        mesh.to_host(capi.ON_CELL, dt_, pint.numpy(), sync=False)
        torch.cuda.current_stream().synchronize()
        assert np.array_equal(pint.numpy(), tr)
    st.close(); mesh.close()


def test_derived_mesh_fields_bit_exact(case):
    from tests.oracle_lib import Oracle
    _, (L, cfg, (h, u, tr), dt) = case
    mesh, st, state = gpu_objects(L, cfg, tr.shape[0])
    o = Oracle(L, cfg, tr.shape[0])
    for a, b in zip(mesh.derived(), o.derived()):
        assert np.array_equal(a, b)
    o.close(); st.close(); mesh.close()


@pytest.mark.parametrize("reorder", [True, False])
def test_aux_and_tendencies_one_rhs(case, reorder):
    from tests.oracle_lib import Oracle
    name, (L, cfg, (h, u, tr), dt) = case
    T = tr.shape[0]
    mesh, st, state = gpu_objects(L, cfg, T, reorder)
    o = Oracle(L, cfg, T)
    o.tend_all(h, u, tr, time=0.0)
    dh, du = mesh.to_device(capi.ON_CELL, h), mesh.to_device(capi.ON_EDGE, u)
    dtr = mesh.to_device(capi.ON_CELL, tr) if T else None
    st.computeAllTendencies(dh, du, dtr, 0.0)
    aux = st.aux_to_host()
    # the TC5 flow is non-divergent: VelocityDivCell (and Del2DivCell) are cancellation residue
    # (>= 1000x below the terms that cancel), so their error is measured against the size of
    # those terms (|u|/dc, |u|/dc^3) instead of against the residue itself
    dc, umax = float(np.mean(L.DcEdge[:-1])), float(np.abs(u).max())
    floor = {"VelocityDivCell": umax / dc, "Del2DivCell": umax / dc**3}
    for n in capi.AUX_FIELDS:
        if n in ("ZonalStressCell", "MeridStressCell", "NormalStressEdge"):
            continue
        err = rel_max(aux[n], o.aux[n])
        if n in floor and np.abs(o.aux[n]).max() < 1e-3 * floor[n]:
            err = np.abs(aux[n] - o.aux[n]).max() / floor[n]
        assert err <= TOL_TEND, (name, n, err)
    tend = st.tend_to_host()
    for n in tend:
        assert rel_max(tend[n], o.tend[n]) <= TOL_TEND, (name, n, rel_max(tend[n], o.tend[n]))
        if tend[n].size:
            assert np.abs(o.tend[n]).max() > 0, "degenerate case: tendency is identically zero"
    o.close(); st.close(); mesh.close()


@pytest.mark.parametrize("land", [False, True])
def test_all_optional_terms_one_rhs(land):
    """wind forcing + bottom drag + upwind fluxes (off in Default.yml); with a coastline: upwind picks between an
    ocean cell and the boundary slot, the wind stress is interpolated over vertices with missing cells."""
    from tests.oracle_lib import Oracle
    L, cfg, (h, u, tr), dt, (zonal, merid) = cases.all_terms_case(land=land)
    T = tr.shape[0]
    mesh, st, state = gpu_objects(L, cfg, T)
    o = Oracle(L, cfg, T)
    o.set_wind_stress(zonal, merid)
    st.set_wind_stress(zonal, merid)
    o.tend_all(h, u, tr)
    dh, du, dtr = mesh.to_device(capi.ON_CELL, h), mesh.to_device(capi.ON_EDGE, u), mesh.to_device(capi.ON_CELL, tr)
    st.computeAllTendencies(dh, du, dtr)
    aux, tend = st.aux_to_host(), st.tend_to_host()
    for n in capi.AUX_FIELDS:
        assert rel_max(aux[n], o.aux[n]) <= TOL_TEND, (n, rel_max(aux[n], o.aux[n]))
    for n in tend:
        assert rel_max(tend[n], o.tend[n]) <= TOL_TEND, (n, rel_max(tend[n], o.tend[n]))
    o.close(); st.close(); mesh.close()


def test_manufactured_solution_rhs():
    """config 1: single layer planar hex with the custom source terms."""
    from omega_b200 import analytic
    from tests.oracle_lib import Oracle
    L, cfg, _, dt = cases.plane_case(16, 1, 0, ManufacturedSolutionTendency=1, ManufacturedWavelengthX=16 * 5.0e6 / 64,
                                     ManufacturedWavelengthY=16 * 5.0e6 / 64 * np.sqrt(3) / 2,
                                     TracerHorzAdvTendencyEnable=0, TracerDiffTendencyEnable=0,
                                     TracerHyperDiffTendencyEnable=0)
    h, u = analytic.manufactured(L, cfg)
    tr = np.zeros((0, h.shape[0], 1))
    mesh, st, state = gpu_objects(L, cfg, 0)
    o = Oracle(L, cfg, 0)
    o.tend_all(h, u, tr, time=123.0)
    st.computeAllTendencies(mesh.to_device(capi.ON_CELL, h), mesh.to_device(capi.ON_EDGE, u), None, 123.0)
    tend = st.tend_to_host()
    for n in ("LayerThicknessTend", "NormalVelocityTend"):
        assert rel_max(tend[n], o.tend[n]) <= TOL_TEND, (n, rel_max(tend[n], o.tend[n]))
    o.close(); st.close(); mesh.close()


PATH_FUSED, PATH_TILED_CELL, PATH_CELL_V2, PATH_TILED_EDGE = 1, 2, 4, 8


def run_steps(L, cfg, init, dt, stepper, nsteps, fused, compat, reorder=True, want_path=0, t0=0.0):
    """nsteps of `stepper` on the GPU and in the oracle; relative L2 errors of (h, u[, tracers]).
    want_path: OMEGA_PATH_* bits the GPU step must have run (omega_stepper_last_path)."""
    from tests.oracle_lib import Oracle
    h, u, tr = init
    T = tr.shape[0]
    mesh, st, state = gpu_objects(L, cfg, T, reorder)
    o = Oracle(L, cfg, T)
    o.make_state(h, u, tr)
    state.copyToDevice(h, u, tr)
    for i in range(nsteps):
        co = capi.rk_coeffs(stepper, dt, t0=t0 + i * dt)
        o.do_step(stepper, co, compat=int(compat))
        st.doStep(state, stepper, co, fused=fused, compat=compat)
    assert st.last_path() & want_path == want_path, (st.last_path(), want_path)
    got, exp = state.copyToHost(0), o.current()
    errs = [rel_l2(got[0], exp[0]), rel_l2(got[1], exp[1])]
    if T:
        errs.append(rel_l2(got[2], exp[2]))
    assert np.all(np.isfinite(exp[0])) and np.all(np.isfinite(exp[1]))
    o.close(); st.close(); mesh.close()
    return errs


@pytest.mark.parametrize("stepper", [capi.RK4, capi.RK2, capi.FB])
@pytest.mark.parametrize("compat", [True, False])
def test_steppers_kernel_per_term_path(case, stepper, compat):
    name, (L, cfg, init, dt) = case
    errs = run_steps(L, cfg, init, dt, stepper, 5, fused=False, compat=compat)
    assert max(errs) <= TOL_STATE, (name, errs)


@pytest.mark.parametrize("compat", [True, False])
def test_rk4_fused_path(case, compat):
    name, (L, cfg, init, dt) = case
    # heptagons: Fused=1 steps on the general kernels (step.cu: the fused ones unroll six edges per cell)
    want = 0 if name.startswith("hept") else PATH_FUSED
    if name in TILED:   # the kernels bench.py times: k_fCT2 (k_fCT without tracer Del4) and k_fET
        want |= PATH_TILED_CELL | PATH_TILED_EDGE | (0 if "nodel4" in name else PATH_CELL_V2)
    errs = run_steps(L, cfg, init, dt, capi.RK4, 5, fused=True, compat=compat, want_path=want)
    assert max(errs) <= TOL_STATE, (name, errs)


def test_fused_kernels_right_hand_side_direct(case):
    """The <= 1e-12 per-field tendency gate ON the fused (tiled) kernels themselves.  They never store a stage's
    right-hand side R, but with stage coefficients B = (2^20, 0, 0, 0), A = 0 the first RK4 phase leaves
    next = cur + 2^20 R(cur) (tracers: cur h + 2^20 R in thickness-weighted form, TimeStepper.cpp:446-524), the
    second stage adding 0 x R(cur); scaling by a power of two is exact, so R is recovered to rounding."""
    from tests.oracle_lib import Oracle
    name, (L, cfg, (h, u, tr), dt) = case
    T = tr.shape[0]
    mesh, st, state = gpu_objects(L, cfg, T)
    o = Oracle(L, cfg, T)
    o.tend_all(h, u, tr, time=0.0)
    state.copyToDevice(h, u, tr)
    big = float(2 ** 20)
    co = capi.StepCoeffs()
    co.NStages, co.Dt = 4, 1.0
    for s in range(4):
        co.A[s], co.B[s], co.StageTime[s] = 0.0, (big if s == 0 else 0.0), 0.0
    st.stepPhase(state, capi.RK4, co, 0, fused=True, compat=True)
    if name in TILED:
        assert st.last_path() & (PATH_TILED_CELL | PATH_TILED_EDGE) == PATH_TILED_CELL | PATH_TILED_EDGE, st.last_path()
    hn = mesh.to_host(capi.ON_CELL, state.LayerThickness[1])
    un = mesh.to_host(capi.ON_EDGE, state.NormalVelocity[1])
    nc, ne = int(L.NCellsAll), int(L.NEdgesAll)
    got = {"LayerThicknessTend": (hn - h) / big, "NormalVelocityTend": (un - u) / big}
    if T:
        trn = mesh.to_host(capi.ON_CELL, state.Tracers[1])
        got["TracerTend"] = (trn - tr * h[None]) / big
    for n, g in got.items():
        w = o.tend[n]
        rows = ne if n == "NormalVelocityTend" else nc
        err = rel_max(g[..., :rows, :], w[..., :rows, :])
        assert err <= TOL_TEND, (name, n, err)
    o.close(); st.close(); mesh.close()


@pytest.mark.parametrize("stepper", [capi.RK2, capi.FB])
@pytest.mark.parametrize("compat", [True, False])
def test_rk2_and_forward_backward_fused_path(case, stepper, compat):
    """RungeKutta2Stepper.cpp:27-73 and ForwardBackwardStepper.cpp:49-68 on the fused kernels
    (tiled where the case allows).  Forward-backward with raw time-level indexing takes the kernel-per-term
    path on odd steps (its tracer fluxes then see another thickness than its thickness flux), so five steps
    mix both paths there."""
    name, (L, cfg, init, dt) = case
    errs = run_steps(L, cfg, init, dt, stepper, 5, fused=True, compat=compat)
    assert max(errs) <= TOL_STATE, (name, errs)
    if name in TILED and not (stepper == capi.FB and compat):
        # the last step ran fused on the tiled kernels
        errs = run_steps(L, cfg, init, dt, stepper, 2, fused=True, compat=compat,
                         want_path=PATH_FUSED | PATH_TILED_CELL | PATH_TILED_EDGE)
        assert max(errs) <= TOL_STATE, (name, errs)


def test_rk4_100_steps_state_parity():
    """BASELINE.json: prognostic state after 100 RK4 steps, relative L2 <= 1e-10 (untiled kernels: K = 8)."""
    L, cfg, init, dt = cases.sphere_case(3, 8, 3)
    for fused in (False, True):
        errs = run_steps(L, cfg, init, dt, capi.RK4, 100, fused=fused, compat=True)
        assert max(errs) <= TOL_STATE, (fused, errs)


@pytest.mark.parametrize("compat", [True, False])
def test_rk4_100_steps_state_parity_tiled_kernels(compat):
    """The same gate on the kernels the bench times: K = 16 runs k_fCT2 + k_fET (asserted)."""
    L, cfg, init, dt = cases.sphere_case(4, 16, 3)
    errs = run_steps(L, cfg, init, dt, capi.RK4, 100, fused=True, compat=compat,
                     want_path=PATH_FUSED | PATH_TILED_CELL | PATH_CELL_V2 | PATH_TILED_EDGE)
    assert max(errs) <= TOL_STATE, errs


@pytest.mark.parametrize("compat", [True, False])
@pytest.mark.parametrize("terms", ["upwind", "wind_drag", "all", "all_land"])
def test_fused_tiled_kernels_with_optional_terms(terms, compat):
    """upwind flux options, wind forcing and bottom drag ON THE TILED fused kernels (k_fCT2 /
    k_fET coefficient selection, k_edge_layer_terms), five RK4 steps against the oracle; the path is asserted."""
    from tests.oracle_lib import Oracle
    kw = {}
    if terms in ("upwind", "all", "all_land"):
        kw.update(FluxThicknessType=1, FluxTracerType=1)
    if terms in ("wind_drag", "all", "all_land"):
        kw.update(WindForcingTendencyEnable=1, BottomDragTendencyEnable=1, BottomDragCoeff=1.0e-3)
    L, cfg, (h, u, tr), dt = (cases.land_case if terms == "all_land" else cases.sphere_case)(3, 16, 3, **kw)
    zonal = np.ascontiguousarray(0.1 * np.cos(L.LatCell)); zonal[-1] = 0.0
    merid = np.ascontiguousarray(0.05 * np.sin(2 * L.LonCell) * np.cos(L.LatCell)); merid[-1] = 0.0
    T = tr.shape[0]
    mesh, st, state = gpu_objects(L, cfg, T)
    o = Oracle(L, cfg, T)
    o.set_wind_stress(zonal, merid); st.set_wind_stress(zonal, merid)
    o.make_state(h, u, tr); state.copyToDevice(h, u, tr)
    co = capi.rk_coeffs(capi.RK4, dt)
    for _ in range(5):
        o.do_step(capi.RK4, co, compat=int(compat))
        st.doStep(state, capi.RK4, co, fused=True, compat=compat)
    want = PATH_FUSED | PATH_TILED_CELL | PATH_CELL_V2 | PATH_TILED_EDGE
    assert st.last_path() & want == want, st.last_path()
    got, exp = state.copyToHost(0), o.current()
    errs = [rel_l2(a, b) for a, b in zip(got, exp)]
    o.close(); st.close(); mesh.close()
    assert max(errs) <= TOL_STATE, (terms, errs)


@pytest.mark.parametrize("T", [1, 2])
def test_fused_tiled_kernels_with_manufactured_sources(T):
    """The manufactured source terms on the tiled kernels (k_ms_sources feeds k_fCT2 / k_fET), K = 16."""
    from omega_b200 import analytic
    L, cfg, _, dt = cases.plane_case(16, 16, T, ManufacturedSolutionTendency=1, ManufacturedWavelengthX=16 * 5.0e6 / 64,
                                     ManufacturedWavelengthY=16 * 5.0e6 / 64 * np.sqrt(3) / 2)
    h, u = analytic.manufactured(L, cfg)
    tr = np.ascontiguousarray(analytic.planar_state(L, T)[2])
    want = PATH_FUSED | PATH_TILED_CELL | PATH_CELL_V2 | PATH_TILED_EDGE
    errs = run_steps(L, cfg, (h, u, tr), dt, capi.RK4, 5, fused=True, compat=True, want_path=want)
    assert max(errs) <= TOL_STATE, errs


def test_fused_manufactured_one_tracer_with_del4():
    """Manufactured thickness source + ONE tracer + tracer hyperdiffusion at K = 32 (this used to
    reach k_fCT, whose Del2 role needs two tracers, and read past the tracer array).  k_fCT2 now takes the
    source per cell from k_ms_sources; with OMEGA_TILE_V2=0 the case must fall back to the untiled cell pass."""
    L, cfg, _, dt = cases.plane_case(16, 32, 1, ManufacturedSolutionTendency=1, ManufacturedWavelengthX=16 * 5.0e6 / 64,
                                     ManufacturedWavelengthY=16 * 5.0e6 / 64 * np.sqrt(3) / 2)
    from omega_b200 import analytic
    h, u = analytic.manufactured(L, cfg)
    tr = np.ascontiguousarray(analytic.planar_state(L, 1)[2])
    tr[0, :-1] = 2.0 + np.cos(2 * np.pi * L.XCell[:-1, None] / float(L.Lx))
    for fused in (True, False):
        errs = run_steps(L, cfg, (h, u, tr), dt, capi.RK4, 5, fused=fused, compat=True)
        assert max(errs) <= TOL_STATE, (fused, errs)


def config1_problem(nx):
    """BASELINE configs[0]: planar hex nx x nx cells, ONE level, no tracers, the manufactured solution of
    src/ocn/CustomTendencyTerms.cpp:18-210 with Default.yml's wavelength = domain size (5000 km)."""
    from omega_b200 import analytic, decomp, meshgen
    dc = 5.0e6 / nx
    m = meshgen.planar_hex(nx, nx, dc, f0=1.0e-4, bottom_depth=1000.0)
    L = decomp.build_all(m, 1, 1)[0]
    cfg = capi.TendConfig.default(ManufacturedSolutionTendency=1, ManufacturedWavelengthX=float(L.Lx),
                                  ManufacturedWavelengthY=float(L.Ly), TracerHorzAdvTendencyEnable=0,
                                  TracerDiffTendencyEnable=0, TracerHyperDiffTendencyEnable=0)
    h, u = analytic.manufactured(L, cfg)
    return L, cfg, h, u


@pytest.mark.parametrize("fused", [True, False])
def test_config1_manufactured_rk4_stepping_and_convergence(fused):
    """Config 1 as a stepping test on the GPU: RK4 from t = 0 to a quarter period on nx = 16, 32, 64 with
    dt ~ dc; (a) state against the oracle <= 1e-10, (b) the error against the exact solution converges at
    second order in space (doc/design/OmegaV0ShallowWater.md:656-666; test/timeStepping/TimeStepperTest.cpp:
    375-388 is the same rate check for the time discretisation)."""
    from omega_b200 import analytic
    t_end, errs_exact = 9600.0, []
    for nx, dt in ((16, 480.0), (32, 240.0), (64, 120.0)):
        L, cfg, h, u = config1_problem(nx)
        tr = np.zeros((0, h.shape[0], 1))
        n = int(round(t_end / dt))
        mesh, st, state = gpu_objects(L, cfg, 0)
        state.copyToDevice(h, u, None)
        from tests.oracle_lib import Oracle
        o = Oracle(L, cfg, 0)
        o.make_state(h, u, tr)
        for i in range(n):
            co = capi.rk_coeffs(capi.RK4, dt, t0=i * dt)
            st.doStep(state, capi.RK4, co, fused=fused, compat=True)
            o.do_step(capi.RK4, co, compat=1)
        gh, gu, _ = state.copyToHost(0)
        eh, eu, _ = o.current()
        assert rel_l2(gh, eh) <= TOL_STATE and rel_l2(gu, eu) <= TOL_STATE, (nx, rel_l2(gh, eh), rel_l2(gu, eu))
        xh, xu = analytic.manufactured(L, cfg, t=n * dt)
        H0 = float(L.BottomDepth[0])
        errs_exact.append((np.sqrt(np.mean((gh[:-1] - xh[:-1]) ** 2)) / cfg.ManufacturedAmplitude,
                           np.sqrt(np.mean((gu[:-1] - xu[:-1]) ** 2)) / cfg.ManufacturedAmplitude))
        assert abs((L.AreaCell[:-1] * (gh[:-1, 0] - H0)).sum()) <= 1e-9 * (L.AreaCell[:-1] * H0).sum()
        o.close(); st.close(); mesh.close()
    rates = [np.log2(np.array(errs_exact[i]) / np.array(errs_exact[i + 1])) for i in range(2)]
    print("config 1 errors (h, u) / eta0:", errs_exact, "rates:", rates)
    assert all(abs(r - 2.0) <= 0.35 for r in rates[-1]), rates


def test_wind_stress_anisotropic_interpolation():
    """WindStressInterpType = 1 (HorzOperators.h:137-187, anisotropic branch): NormalStressEdge and the
    velocity tendency against the oracle."""
    from tests.oracle_lib import Oracle
    L, cfg, (h, u, tr), dt, (zonal, merid) = cases.all_terms_case()
    cfg.WindStressInterpType = 1
    T = tr.shape[0]
    mesh, st, state = gpu_objects(L, cfg, T)
    o = Oracle(L, cfg, T)
    o.set_wind_stress(zonal, merid)
    st.set_wind_stress(zonal, merid)
    o.tend_all(h, u, tr)
    st.computeAllTendencies(mesh.to_device(capi.ON_CELL, h), mesh.to_device(capi.ON_EDGE, u), mesh.to_device(capi.ON_CELL, tr))
    aux, tend = st.aux_to_host(("NormalStressEdge",)), st.tend_to_host()
    assert np.abs(o.aux["NormalStressEdge"]).max() > 0
    assert rel_max(aux["NormalStressEdge"], o.aux["NormalStressEdge"]) <= TOL_TEND
    assert rel_max(tend["NormalVelocityTend"], o.tend["NormalVelocityTend"]) <= TOL_TEND
    # and it differs from the isotropic interpolation, so the switch is live
    cfg0 = capi.TendConfig.from_buffer_copy(cfg)
    cfg0.WindStressInterpType = 0
    o.set_config(cfg0)
    o.tend_all(h, u, tr)
    assert rel_max(aux["NormalStressEdge"], o.aux["NormalStressEdge"]) > 1e-6
    o.close(); st.close(); mesh.close()


def test_tiled_and_untiled_fused_paths_agree(monkeypatch):
    """fused_tile.cu against the fused.cu kernels it replaces (OMEGA_TILE=0, OMEGA_TILE_E=0), same inputs."""
    L, cfg, init, dt = cases.sphere_case(4, 32, 4)
    res = []
    for tile in ("1", "0"):
        monkeypatch.setenv("OMEGA_TILE", tile)
        monkeypatch.setenv("OMEGA_TILE_E", tile)
        from omega_b200 import model
        h, u, tr = init
        mesh, st, state = gpu_objects(L, cfg, tr.shape[0])
        state.copyToDevice(h, u, tr)
        co = capi.rk_coeffs(capi.RK4, dt)
        for _ in range(3):
            st.doStep(state, capi.RK4, co, fused=True, compat=True)
        res.append(state.copyToHost(0))
        st.close(); mesh.close()
    for a, b in zip(*res):
        assert rel_l2(a, b) <= 1e-13, rel_l2(a, b)


@pytest.mark.parametrize("fused,K,land", [(False, 4, False), (True, 4, False), (True, 16, False), (True, 16, True)])
@pytest.mark.parametrize("ntasks", [2, 3])
def test_multirank_emulated_on_one_gpu(ntasks, fused, K, land):
    """N ranks' partitions stepped on one GPU in one process: the library's round pack / unpack kernels,
    exchange lists in device order, phase-split stepping with the halo attached (boundary tiles first,
    halo tiles skipped before an exchange), with device-to-device copies standing in for NCCL.  Checked
    against the 1-rank oracle on every local element (HaloWidth 4: see
    test_multirank_cpu.test_reference_needs_halo_width_4_with_del4)."""
    import ctypes as C
    from omega_b200 import analytic, decomp, model, parallel
    from tests.oracle_lib import Oracle
    T, nsteps = 2, 3
    level = 4 if K == 16 else 3          # level 4: 2562 cells, enough tiles for all three tile classes
    mesh_g = cases.land_mesh(level) if land else cases.sphere_mesh(level)
    L1, cfg, (h, u, tr), dt = (cases.land_case if land else cases.sphere_case)(level, K, T)
    ref = Oracle(L1, cfg, T)
    ref.make_state(h, u, tr)
    for s in range(nsteps):
        ref.do_step(capi.RK4, capi.rk_coeffs(capi.RK4, dt, t0=s * dt))
    gh, gu, gt = ref.current()
    cell_of, edge_of = L1.CellID[:-1] - 1, L1.EdgeID[:-1] - 1
    H = np.empty_like(gh[:-1]); H[cell_of] = gh[:-1]
    U = np.empty_like(gu[:-1]); U[edge_of] = gu[:-1]
    TR = np.empty_like(gt[:, :-1]); TR[:, cell_of] = gt[:, :-1]
    locs = decomp.build_all(mesh_g, ntasks, K, 4)
    lists = parallel.build_halo_lists(locs)
    ranks = []
    for L, hl in zip(locs, lists):
        m = model.HorzMesh(L, reorder=True)
        st, state = model.Stepper(m, cfg, T), model.OceanState(m, T)
        state.copyToDevice(*analytic.tc5(L, T))
        ex = parallel.HaloExchanger(m, hl, T)
        ex.attach(st)
        ranks.append((L, m, st, state, ex))
    lib = ranks[0][1].lib
    stream = C.c_void_p(0)
    import torch
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    split_seen = False
    for s in range(nsteps):
        co = capi.rk_coeffs(capi.RK4, dt, t0=s * dt)
        for p in range(2):
            rounds = []
            for L, m, st, state, ex in ranks:
                st.stepPhase(state, capi.RK4, co, p, fused=fused, compat=True)
                split_seen |= bool(st.last_path() & 16)
                ptrs = ex.exchange_list(st, state, capi.RK4, p)
                sbuf, sc, rc = ex.pack_round(ptrs)
                rounds.append((ptrs, sbuf, sc, rc, ex.recv_buffer(ptrs)))
            # transport: my message to neighbour nbr is nbr's message from me (offsets in neighbour order)
            for t, (rk, (ptrs, sbuf, sc, rc, rbuf)) in enumerate(zip(ranks, rounds)):
                soff = np.concatenate(([0], np.cumsum(sc)))
                for i, nbr in enumerate(rk[4].nbrs):
                    _, _, sc2, rc2, rbuf2 = rounds[nbr]
                    j = ranks[nbr][4].nbrs.index(t)
                    roff2 = np.concatenate(([0], np.cumsum(rc2)))
                    assert rc2[j] == sc[i]
                    assert lib.omega_dev_copy(C.c_void_p(rbuf2 + 8 * int(roff2[j])), C.c_void_p(sbuf + 8 * int(soff[i])),
                                              C.c_ulonglong(8 * sc[i]), stream) == 0
            for (L, m, st, state, ex), (ptrs, *_rest) in zip(ranks, rounds):
                ex.unpack_round(ptrs)
        for L, m, st, state, ex in ranks:
            m.lib.omega_state_rotate(C.byref(state.c))
    if fused and K == 16:
        assert split_seen, "the attached halo must switch the tiled path to boundary-first launches"
        nb, nown, nall = ranks[0][4].tile_classes()
        assert 0 < nb < nown < nall, (nb, nown, nall)
    worst = 0.0
    for L, m, st, state, ex in ranks:
        lh, lu, lt = state.copyToHost(0)
        ci, ei = L.CellID[:L.NCellsAll] - 1, L.EdgeID[:L.NEdgesAll] - 1
        for a, b in ((lh[:L.NCellsAll], H[ci]), (lu[:L.NEdgesAll], U[ei]), (lt[:, :L.NCellsAll], TR[:, ci])):
            worst = max(worst, rel_l2(a, b))
        ex.close(); st.close(); m.close()
    ref.close()
    assert worst <= TOL_STATE, worst


@pytest.mark.parametrize("T,v2", [(1, "1"), (2, "1"), (3, "1"), (5, "1"), (10, "1"), (2, "0"), (5, "0")])
def test_tiled_kernels_bit_reproducible_and_stay_in_bounds(monkeypatch, T, v2):
    """compute-sanitizer is closed on this GPU pool, so the named-barrier / setmaxnreg / cp.async protocols
    of fused_tile.cu are checked the way a race or a stray store would show: the same three fused steps run
    three times from the same state must give bit-identical results (k_fCT2 with its short and long copy
    schedules, and the older k_fCT with OMEGA_TILE_V2=0) and the boundary-slot rows (row N*All, zero by
    construction, Decomp.cpp:1081-1082) must stay exactly zero
    untouched where the step must not write."""
    monkeypatch.setenv("OMEGA_TILE_V2", v2)
    L, cfg, (h, u, tr), dt = cases.sphere_case(4, 32, T)
    co = capi.rk_coeffs(capi.RK4, dt)
    runs = []
    mesh, st, state = gpu_objects(L, cfg, T)
    for rep in range(3):
        state.c.CurTimeIndex = 0
        state.c.TracerCurTimeIndex = 0
        state.copyToDevice(h, u, tr)
        for _ in range(3):
            st.doStep(state, capi.RK4, co, fused=True, compat=True)
        assert st.last_path() & (PATH_TILED_CELL | PATH_TILED_EDGE) == PATH_TILED_CELL | PATH_TILED_EDGE
        gh, gu, gt = state.copyToHost(0)
        runs.append((gh.copy(), gu.copy(), gt.copy()))
        for a in (gh, gu):
            assert np.all(a[-1] == 0.0), "boundary slot row written"
        assert np.all(gt[:, -1] == 0.0), "boundary slot row of a tracer written"
    for a, b in zip(runs[0], runs[1]):
        assert np.array_equal(a, b)
    for a, b in zip(runs[0], runs[2]):
        assert np.array_equal(a, b)
    st.close(); mesh.close()


@pytest.mark.parametrize("nranks", [2, 4, 8])
def test_nccl_ranks_match_oracle(nranks):
    """The NCCL path itself (library exchange rounds inside omega_stepper_do_step, overlapped with the interior
    tiles): N ranks, one GPU each, against the 1-rank oracle on every local element (tests/nccl_check.py)."""
    import os
    import subprocess
    import sys
    import torch
    if torch.cuda.device_count() < nranks:
        pytest.skip(f"needs {nranks} GPUs (gpurun --gpus {nranks})")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={nranks}", "--master-addr", "127.0.0.1",
           "--master-port", str(29600 + nranks), os.path.join(root, "tests", "nccl_check.py"), "5" if nranks > 2 else "4", "16", "3", "4"]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert out.returncode == 0 and "NCCL_CHECK_OK" in out.stdout, (out.stdout[-2000:], out.stderr[-2000:])


def test_config3_full_size_level_slices_and_conservation():
    """BASELINE.json config 3 at FULL size (655 362 cells x 64 levels x 10 tracers), two RK4 steps of the
    bench path (fused, tiled kernels).  Levels do not interact, so the oracle can check the full horizontal
    mesh one level at a time: levels 0 and 37 of the GPU result against oracle runs with a single level
    (relative L2 <= 1e-10, raw time-level indexing as the reference runs it).  Size-independent
    properties on all 64 levels, with consistent time levels (RawTimeLevelCompat = 0; the raw indexing
    feeds the tracer advection another velocity level than the thickness equation on odd steps, SURVEY
    3.5 P1, and then neither property holds in the reference either): sum(A h) per level is conserved and
    the constant tracer stays 1; the drift of sum(A h phi) must not exceed the oracle's."""
    from tests.oracle_lib import Oracle
    from omega_b200 import analytic, decomp
    K, T, dt, nsteps = 64, 10, 60.0, 2
    m = cases.sphere_mesh(8)
    dc = float(np.mean(m.dcEdge))
    cfg = capi.TendConfig.default(ViscDel2=1.0e3 * (dc / 30.0e3) ** 2, ViscDel4=1.2e11 * (dc / 30.0e3) ** 4,
                                  EddyDiff2=10.0 * (dc / 30.0e3) ** 2, EddyDiff4=1.0e10 * (dc / 30.0e3) ** 4)
    L = decomp.build_all(m, 1, K)[0]
    L1 = decomp.build_all(m, 1, 1)[0]
    h, u, tr = analytic.tc5(L, T)
    co = capi.rk_coeffs(capi.RK4, dt)
    nC = int(L.NCellsOwned)
    A = np.asarray(L.AreaCell)[:nC, None]

    def drift(h0, t0, h1, t1):
        """largest relative change of sum(A h) and of sum(A h phi_l) over levels / tracers"""
        dm = np.max(np.abs((A * h1).sum(0) - (A * h0).sum(0)) / (A * h0).sum(0))
        dq = max(np.max(np.abs((A * h1 * t1[l]).sum(0) - (A * h0 * t0[l]).sum(0)) / np.abs((A * h0 * t0[l]).sum(0)))
                 for l in range(T))
        return dm, dq

    mesh, st, state = gpu_objects(L, cfg, T)
    for compat in (True, False):
        state.copyToDevice(h, u, tr)
        for _ in range(nsteps):
            st.doStep(state, capi.RK4, co, fused=True, compat=compat)
        gh, gu, gtr = state.copyToHost(0)
        assert np.isfinite(gh).all() and np.isfinite(gu).all() and np.isfinite(gtr).all()
        if not compat:
            dm, _ = drift(h[:nC], tr[:, :nC], gh[:nC], gtr[:, :nC])
            one = np.max(np.abs(gtr[T - 1, :nC] - 1.0))
            print("config3, consistent time levels: mass drift", dm, "max |constant tracer - 1|", one)
            assert dm <= 1e-12, dm
            assert one <= 1e-11, one
        # the oracle on the full horizontal mesh, one level at a time
        for k in (0, 37) if compat else (37,):
            o = Oracle(L1, cfg, T)
            o.make_state(np.ascontiguousarray(h[:, k:k + 1]), np.ascontiguousarray(u[:, k:k + 1]),
                         np.ascontiguousarray(tr[:, :, k:k + 1]))
            for _ in range(nsteps):
                o.do_step(capi.RK4, co, compat=int(compat))
            eh, eu, etr = o.current()
            errs = [rel_l2(gh[:, k:k + 1], eh), rel_l2(gu[:, k:k + 1], eu), rel_l2(gtr[:, :, k:k + 1], etr)]
            odm, odq = drift(h[:nC, k:k + 1], tr[:, :nC, k:k + 1], eh[:nC], etr[:, :nC])
            gdm, gdq = drift(h[:nC, k:k + 1], tr[:, :nC, k:k + 1], gh[:nC, k:k + 1], gtr[:, :nC, k:k + 1])
            print("compat", compat, "level", k, "rel L2 (h, u, tr)", errs, "drift oracle", odm, odq, "gpu", gdm, gdq)
            o.close()
            assert max(errs) <= TOL_STATE, (compat, k, errs)
            assert gdq <= 2.0 * odq + 1e-13, (compat, k, gdq, odq)
    st.close(); mesh.close()
