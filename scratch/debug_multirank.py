import ctypes as C, os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from omega_b200 import analytic, capi, decomp, model, parallel
from tests import cases
from tests.oracle_lib import Oracle
from tests.cases import rel_l2

def run(ntasks, level, K, T, nsteps, attach):
    mesh_g = cases.sphere_mesh(level)
    L1, cfg, (h, u, tr), dt = cases.sphere_case(level, K, T)
    ref = Oracle(L1, cfg, T); ref.make_state(h, u, tr)
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
        if attach: ex.attach(st)
        ranks.append((L, m, st, state, ex))
    lib = ranks[0][1].lib
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
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
                    lib.omega_dev_copy(C.c_void_p(rbuf2 + 8 * int(roff2[j])), C.c_void_p(sbuf + 8 * int(soff[i])), C.c_ulonglong(8 * sc[i]), stream)
            for (L, m, st, state, ex), (ptrs, *_r) in zip(ranks, rounds):
                ex.unpack_round(ptrs)
        for L, m, st, state, ex in ranks:
            m.lib.omega_state_rotate(C.byref(state.c))
    out = []
    for L, m, st, state, ex in ranks:
        lh, lu, lt = state.copyToHost(0)
        ci, ei = L.CellID[:L.NCellsAll] - 1, L.EdgeID[:L.NEdgesAll] - 1
        no, neo = int(L.NCellsOwned), int(L.NEdgesOwned)
        out.append(("all", rel_l2(lh[:L.NCellsAll], H[ci]), rel_l2(lu[:L.NEdgesAll], U[ei]), rel_l2(lt[:, :L.NCellsAll], TR[:, ci]),
                    "owned", rel_l2(lh[:no], H[ci[:no]]), rel_l2(lu[:neo], U[ei[:neo]]), rel_l2(lt[:, :no], TR[:, ci[:no]]), st.last_path()))
        ex.close(); st.close(); m.close()
    ref.close()
    return out

if __name__ == "__main__":
    ntasks, level, K, nsteps, attach = (int(x) for x in sys.argv[1:6])
    for r in run(ntasks, level, K, 2, nsteps, attach):
        print(os.environ.get("OMEGA_TILE", "-"), os.environ.get("OMEGA_TILE_E", "-"), os.environ.get("OMEGA_HALO_OVERLAP", "-"), "attach", attach, r)
