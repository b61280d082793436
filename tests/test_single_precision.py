"""The single-precision build (libomega_b200_sp.so: the same sources with -DOMEGA_SINGLE_PRECISION, Real = float as in
the reference's src/base/DataTypes.h:30-34) against the fp64 oracle.  fp32 round-off bounds what can be asked:
fields and per-step state to a few 1e-6 relative, tendencies (differences of nearly equal numbers) to 1e-3 of their
largest entry.  Transfers and halo packing move bits and stay exact."""
import ctypes
import os
import re

import numpy as np
import pytest

from omega_b200 import capi
from tests import cases
from tests.cases import rel_l2, rel_max

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

TOL_SP_FIELD = 2e-5      # auxiliary fields after one right-hand side, max|a-b| / max|b|
TOL_SP_TEND = 5e-4       # tendencies after one right-hand side (measured: 7e-5 on the velocity tendency)
TOL_SP_STATE = 5e-6      # state after 5 RK4 / RK2 / forward-backward steps, relative L2 (measured: 1.05e-6)


def _sp_lib():
    if not os.path.exists(capi.LIB_PATH_SP):
        import __graft_entry__
        __graft_entry__.build()
    return capi.load_library(capi.LIB_PATH_SP)


def test_sp_library_exports_every_declared_symbol():
    src = re.sub(r"/\*.*?\*/", "", open(os.path.join(ROOT, "include", "omega_b200.h")).read(), flags=re.S)
    names = sorted(set(re.findall(r"\b(omega_[a-z0-9_]+)\s*\(", src)))
    lib = _sp_lib()
    assert not [n for n in names if not hasattr(lib, n)]
    assert lib.omega_b200_real_bytes() == 4
    assert ctypes.CDLL(capi.LIB_PATH).omega_b200_real_bytes() == 8


def _objects(L, cfg, T, reorder=True):
    from omega_b200 import model
    mesh = model.HorzMesh(L, reorder=reorder, lib=_sp_lib())
    assert mesh.np_dtype == np.float32
    return mesh, model.Stepper(mesh, cfg, T), model.OceanState(mesh, T)


@pytest.mark.gpu
def test_sp_transfers_bit_exact():
    import torch
    L, cfg, (h, u, tr), dt = cases.sphere_case(4, 16, 3)
    mesh, st, state = _objects(L, cfg, 3)
    for kind, a in ((capi.ON_CELL, h), (capi.ON_EDGE, u), (capi.ON_CELL, tr)):
        a32 = a.astype(np.float32)
        d = mesh.to_device(kind, a32)
        assert d.dtype == torch.float32
        p, _ = mesh.perm(kind)
        assert np.array_equal(d.cpu().numpy(), a32[..., p, :])
        assert np.array_equal(mesh.to_host(kind, d), a32)
    got = torch.empty(tr.shape, dtype=torch.float32, device="cuda:0")
    from omega_b200 import model
    a32 = np.ascontiguousarray(tr, np.float32)
    rc = mesh.lib.omega_field_to_device_noalloc(mesh.h, capi.ON_CELL, 3, 16, a32.ctypes.data_as(capi.c_dbl_p), model._dptr(got), model._stream())
    torch.cuda.synchronize()
    assert rc == 0 and torch.equal(got, mesh.to_device(capi.ON_CELL, a32))
    st.close(); mesh.close()


@pytest.mark.gpu
@pytest.mark.parametrize("which", ["default", "all_terms"])
def test_sp_one_rhs_against_fp64_oracle(which):
    from tests.oracle_lib import Oracle
    if which == "default":
        L, cfg, (h, u, tr), dt = cases.sphere_case(4, 16, 3)
        wind = None
    else:
        L, cfg, (h, u, tr), dt, wind = cases.all_terms_case()
    T = tr.shape[0]
    mesh, st, state = _objects(L, cfg, T)
    o = Oracle(L, cfg, T)
    if wind:
        o.set_wind_stress(*wind)
        st.set_wind_stress(*wind)
    o.tend_all(h, u, tr, time=0.0)
    dh, du, dtr = mesh.to_device(capi.ON_CELL, h), mesh.to_device(capi.ON_EDGE, u), mesh.to_device(capi.ON_CELL, tr)
    st.computeAllTendencies(dh, du, dtr, 0.0)
    aux, tend = st.aux_to_host(), st.tend_to_host()
    dc, umax = float(np.mean(L.DcEdge[:-1])), float(np.abs(u).max())
    # fields that are residues of cancelling terms are measured against the size of those terms
    floor = {"VelocityDivCell": umax / dc, "Del2DivCell": umax / dc**3, "RelVortVertex": umax / dc,
             "Del2Edge": umax / dc**2, "Del2RelVortVertex": umax / dc**3,
             "Del2TracersCell": float(np.abs(h).max() * np.abs(tr).max()) / dc**2 if tr.size else 0.0}
    worst = {}
    for n in capi.AUX_FIELDS:
        if o.aux[n].size == 0 or np.abs(o.aux[n]).max() == 0:
            continue
        assert aux[n].dtype == np.float32
        err = np.abs(aux[n] - o.aux[n]).max() / max(np.abs(o.aux[n]).max(), floor.get(n, 0.0))
        worst[n] = err
        assert err <= TOL_SP_FIELD, (n, err)
    for n in tend:
        if tend[n].size:
            worst[n] = rel_max(tend[n], o.tend[n])
            assert worst[n] <= TOL_SP_TEND, (n, worst[n])
    print("single precision, one RHS:", {k: float("%.2g" % v) for k, v in worst.items()})
    o.close(); st.close(); mesh.close()


@pytest.mark.gpu
@pytest.mark.parametrize("stepper", [capi.RK4, capi.RK2, capi.FB])
@pytest.mark.parametrize("fused", [True, False])
def test_sp_five_steps_against_fp64_oracle(stepper, fused):
    from tests.oracle_lib import Oracle
    L, cfg, (h, u, tr), dt = cases.sphere_case(4, 16, 3)
    T = tr.shape[0]
    mesh, st, state = _objects(L, cfg, T)
    o = Oracle(L, cfg, T)
    o.make_state(h, u, tr)
    state.copyToDevice(h, u, tr)
    for i in range(5):
        co = capi.rk_coeffs(stepper, dt, t0=i * dt)
        o.do_step(stepper, co, compat=1)
        st.doStep(state, stepper, co, fused=fused, compat=True)
    if fused and stepper != capi.FB:
        assert st.last_path() == 1, "the single-precision build steps through the untiled fused kernels"
    gh, gu, gtr = state.copyToHost(0)
    oh, ou, otr = o.current()
    errs = (rel_l2(gh, oh), rel_l2(gu, ou), rel_l2(gtr, otr))
    print("single precision, 5 steps:", stepper, fused, errs)
    assert max(errs) <= TOL_SP_STATE, errs
    o.close(); st.close(); mesh.close()


@pytest.mark.gpu
def test_sp_two_ranks_emulated_on_one_gpu():
    """The halo round kernels and the phase-split stepper of the single-precision build: two partitions stepped on one
    GPU, device-to-device copies standing in for NCCL (messages counted in 4-byte reals), against the 1-rank oracle."""
    import ctypes as C
    import torch
    from omega_b200 import analytic, decomp, model, parallel
    from tests.oracle_lib import Oracle
    T, K, nsteps, ntasks = 2, 4, 3, 2
    mesh_g = cases.sphere_mesh(3)
    L1, cfg, (h, u, tr), dt = cases.sphere_case(3, K, T)
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
    lib = _sp_lib()
    ranks = []
    for L, hl in zip(locs, lists):
        m = model.HorzMesh(L, reorder=True, lib=lib)
        st, state = model.Stepper(m, cfg, T), model.OceanState(m, T)
        state.copyToDevice(*analytic.tc5(L, T))
        ex = parallel.HaloExchanger(m, hl, T)
        ex.attach(st)
        ranks.append((L, m, st, state, ex))
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    RB = 4
    for s in range(nsteps):
        co = capi.rk_coeffs(capi.RK4, dt, t0=s * dt)
        for p in range(2):
            rounds = []
            for L, m, st, state, ex in ranks:
                st.stepPhase(state, capi.RK4, co, p, fused=True, compat=True)
                ptrs = ex.exchange_list(st, state, capi.RK4, p)
                sbuf, sc, rc = ex.pack_round(ptrs)
                rounds.append((ptrs, sbuf, sc, rc, ex.recv_buffer(ptrs)))
            for t, (rk, (ptrs, sbuf, sc, rc, rbuf)) in enumerate(zip(ranks, rounds)):
                soff = np.concatenate(([0], np.cumsum(sc)))
                for i, nbr in enumerate(rk[4].nbrs):
                    _, _, sc2, rc2, rbuf2 = rounds[nbr]
                    j = ranks[nbr][4].nbrs.index(t)
                    roff2 = np.concatenate(([0], np.cumsum(rc2)))
                    assert rc2[j] == sc[i]
                    assert lib.omega_dev_copy(C.c_void_p(rbuf2 + RB * int(roff2[j])), C.c_void_p(sbuf + RB * int(soff[i])),
                                              C.c_ulonglong(RB * sc[i]), stream) == 0
            for (L, m, st, state, ex), (ptrs, *_rest) in zip(ranks, rounds):
                ex.unpack_round(ptrs)
        for L, m, st, state, ex in ranks:
            m.lib.omega_state_rotate(C.byref(state.c))
    worst = 0.0
    for L, m, st, state, ex in ranks:
        lh, lu, lt = state.copyToHost(0)
        ci, ei = L.CellID[:L.NCellsAll] - 1, L.EdgeID[:L.NEdgesAll] - 1
        for a, b in ((lh[:L.NCellsAll], H[ci]), (lu[:L.NEdgesAll], U[ei]), (lt[:, :L.NCellsAll], TR[:, ci])):
            worst = max(worst, rel_l2(a, b))
        ex.close(); st.close(); m.close()
    ref.close()
    assert worst <= TOL_SP_STATE, worst
