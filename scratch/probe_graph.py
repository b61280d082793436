"""probe: config 1 / config 2 step time with plain launches (profiling off) and with the step captured in CUDA graphs"""
import sys, time
import numpy as np
import torch
sys.path.insert(0, ".")
import bench
from omega_b200 import capi, decomp, model

def run(name, n=400):
    m, cfg, K, T, dt, ks = bench.build_problem(name, 1)
    co = capi.rk_coeffs(capi.RK4, dt)
    L = decomp.build_all(m, 1, K)[0]
    h, u, tr = bench.initial_state(name, L, cfg, T)
    mesh = model.HorzMesh(L, reorder=True)
    st = model.Stepper(mesh, cfg, T)
    state = model.OceanState(mesh, T)
    state.copyToDevice(h, u, tr if T else None)
    for _ in range(4):
        st.doStep(state, capi.RK4, co, fused=True, compat=True)
    torch.cuda.synchronize()
    n0 = st.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter(); e0.record()
    for _ in range(n):
        st.doStep(state, capi.RK4, co, fused=True, compat=True)
    e1.record(); torch.cuda.synchronize(); t1 = time.perf_counter()
    print(name, "plain launches: %.4f ms/step (device), %.4f host wall, %d launches/step, path %d" % (e0.elapsed_time(e1) / n, (t1 - t0) * 1e3 / n, (st.launch_count() - n0) // n, st.last_path()), flush=True)
    ref = [x.clone() for x in (state.getLayerThickness(0), state.getNormalVelocity(0))]
    # two graphs: one per parity of the time-level ring
    graphs = []
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        for par in range(2):
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=s):
                st.doStep(state, capi.RK4, co, fused=True, compat=True)
            graphs.append(g)
    torch.cuda.current_stream().wait_stream(s)
    torch.cuda.synchronize()
    # capture does not execute: the ring indices advanced twice on the host, data did not move
    e0.record()
    for i in range(n):
        graphs[i & 1].replay()
    e1.record(); torch.cuda.synchronize()
    print(name, "graph replay:   %.4f ms/step (device)" % (e0.elapsed_time(e1) / n), flush=True)
    assert bool(torch.isfinite(state.getLayerThickness(0)).all())
    st.close(); mesh.close()

for nm in sys.argv[1:] or ["config1", "config2"]:
    run(nm)
