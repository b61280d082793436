"""Run under torch.distributed.run with N ranks (one GPU each): every rank steps its partition through
HaloExchanger.do_step -> omega_stepper_do_step with the library's NCCL exchange rounds (overlapped with
the interior tiles) and compares ALL its local elements -- owned and halo -- with the 1-rank CPU oracle.
Prints 'NCCL_CHECK_OK <worst rel L2>' on rank 0; exit code non-zero on any mismatch.
Used by tests/test_gpu_parity.py::test_nccl_ranks_match_oracle (needs >= 2 GPUs)."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    import torch
    import torch.distributed as dist
    from omega_b200 import analytic, capi, decomp, model, parallel
    from tests import cases
    from tests.oracle_lib import Oracle
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    level, K, T, nsteps = (int(x) for x in (sys.argv[1:5] + ["4", "16", "3", "4"][len(sys.argv) - 1:]))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    mesh_g = cases.sphere_mesh(level)
    L1, cfg, (h, u, tr), dt = cases.sphere_case(level, K, T)
    L, hl = parallel.local_problem(mesh_g, world, rank, K, halo_width=4)
    mesh = model.HorzMesh(L, reorder=True, device=dev)
    st, state = model.Stepper(mesh, cfg, T), model.OceanState(mesh, T)
    state.copyToDevice(*analytic.tc5(L, T))
    ex = parallel.HaloExchanger(mesh, hl, T)
    ex.init_comm(world, rank)
    split = False
    for s in range(nsteps):
        ex.do_step(st, state, capi.RK4, capi.rk_coeffs(capi.RK4, dt, t0=s * dt), fused=True, compat=True)
        split |= bool(st.last_path() & 16)
    torch.cuda.synchronize()
    lh, lu, lt = state.copyToHost(0)
    ref = Oracle(L1, cfg, T)
    ref.make_state(h, u, tr)
    for s in range(nsteps):
        ref.do_step(capi.RK4, capi.rk_coeffs(capi.RK4, dt, t0=s * dt))
    gh, gu, gt = ref.current()
    cell_of, edge_of = L1.CellID[:-1] - 1, L1.EdgeID[:-1] - 1
    H = np.empty_like(gh[:-1]); H[cell_of] = gh[:-1]
    U = np.empty_like(gu[:-1]); U[edge_of] = gu[:-1]
    TR = np.empty_like(gt[:, :-1]); TR[:, cell_of] = gt[:, :-1]
    ci, ei = L.CellID[:L.NCellsAll] - 1, L.EdgeID[:L.NEdgesAll] - 1
    worst = max(cases.rel_l2(a, b) for a, b in ((lh[:L.NCellsAll], H[ci]), (lu[:L.NEdgesAll], U[ei]), (lt[:, :L.NCellsAll], TR[:, ci])))
    # test/base/HaloTest.cpp:41-99 over NCCL: arrays of every element type hold global IDs in their owned entries
    # only, omega_halo_exchange_any fills the halo entries, the result must equal the fully initialised array
    import ctypes as C
    bad = 0
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    for kind, ids, nall, nown in ((capi.ON_CELL, L.CellID, L.NCellsAll, L.NCellsOwned), (capi.ON_EDGE, L.EdgeID, L.NEdgesAll, L.NEdgesOwned),
                                  (capi.ON_VERTEX, L.VertexID, L.NVerticesAll, L.NVerticesOwned)):
        for npdt, nouter, ninner in ((np.int32, 1, 1), (np.int64, 2, 3), (np.float32, 1, 5), (np.float64, 3, 2)):
            n = mesh.NSize[kind]
            o, i = np.arange(nouter).reshape(nouter, 1, 1), np.arange(ninner).reshape(1, 1, ninner)
            full = ((np.asarray(ids, np.int64)[:n].reshape(1, n, 1) * 7 + o) * 5 + i).astype(npdt)
            full[:, int(nall):] = 0
            part = full.copy()
            part[:, int(nown):int(nall)] = npdt(-1)
            p, _ = mesh.perm(kind)
            d = torch.from_numpy(np.ascontiguousarray(part[:, p])).to(dev)
            rc = mesh.lib.omega_halo_exchange_any(ex.handle, C.c_void_p(d.data_ptr()), kind, nouter, ninner, np.dtype(npdt).itemsize, stream)
            torch.cuda.synchronize()
            if rc != 0 or not np.array_equal(d.cpu().numpy(), full[:, p]):
                bad += 1
    t = torch.tensor([worst, 0.0 if split else 1.0, float(bad)], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ok = float(t[0]) <= 1e-10 and (K % 16 != 0 or float(t[1]) == 0.0) and float(t[2]) == 0.0
    if rank == 0:
        print(("NCCL_CHECK_OK" if ok else "NCCL_CHECK_FAILED"), float(t[0]), "overlap split on every rank:", float(t[1]) == 0.0,
              "any-type exchanges wrong:", int(t[2]), flush=True)
    ex.close(); st.close(); mesh.close()
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
