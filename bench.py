#!/usr/bin/env python
"""bench.py -- Omega RK4 step throughput on B200 (BASELINE.json metric).

A "step" is one RK4 step (TimeStepper::doStep) of the whole hot path: aux
state + every tendency term + stage updates (+ halo exchange for N > 1) on the
workload of BASELINE.json configs[2]: synthetic icosahedral mesh, 655,362 cells
x 64 levels, 10 tracers, Del2 + Del4 on velocity and tracers, dt = 60 s.

    python bench.py --gpus N --steps K --warmup W [--impl reference]

Prints ONE JSON line (rank 0).  See DESIGN.md "Measurement" for every field.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {
    # name: (icosahedral level | -nx of the planar hex mesh, levels, tracers, dt seconds, CPU-sample levels)
    "config3": (8, 64, 10, 60.0, 8),
    "config2": (6, 64, 2, 240.0, 16),
    "config4": (9, 80, 2, 30.0, 8),       # BASELINE configs[3]: 2,621,442 cells x 80 levels, 2 tracers
    "config5": (8, 64, 32, 60.0, 4),      # BASELINE configs[4]: BGC-like, 32 tracers
    "config1": (-64, 1, 0, 120.0, 1),     # BASELINE configs[0]: planar hex 64 x 64, 1 level, manufactured solution
    "small": (5, 16, 3, 480.0, 16),
}


def workload_cfg(capi, dc):
    """Default.yml terms + tracer hyperdiffusion on (configs[2]: 'Del2+Del4')."""
    return capi.TendConfig.default(ViscDel2=1.0e3 * (dc / 30.0e3) ** 2, ViscDel4=1.2e11 * (dc / 30.0e3) ** 4,
                                   EddyDiff2=10.0 * (dc / 30.0e3) ** 2, EddyDiff4=1.0e10 * (dc / 30.0e3) ** 4)


def build_problem(name, ntasks):
    from omega_b200 import analytic, capi, decomp, meshgen
    level, K, T, dt, ks = WORKLOADS[name]
    if level < 0:   # config 1: doubly periodic planar hex, manufactured solution (Default.yml wavelength = domain)
        nx = -level
        m = meshgen.planar_hex(nx, nx, 5.0e6 / nx, f0=1.0e-4, bottom_depth=1000.0)
        cfg = capi.TendConfig.default(ManufacturedSolutionTendency=1, ManufacturedWavelengthX=float(m.Lx),
                                      ManufacturedWavelengthY=float(m.Ly), TracerHorzAdvTendencyEnable=0,
                                      TracerDiffTendencyEnable=0, TracerHyperDiffTendencyEnable=0)
        return m, cfg, K, T, dt, ks
    m = meshgen.icosahedral(level)
    m.bottomDepth = analytic.tc5_bottom_depth(m)
    cfg = workload_cfg(capi, float(np.mean(m.dcEdge)))
    return m, cfg, K, T, dt, ks


def initial_state(name, L, cfg, T):
    from omega_b200 import analytic
    if WORKLOADS[name][0] < 0:
        h, u = analytic.manufactured(L, cfg)
        return h, u, np.zeros((0, h.shape[0], h.shape[1]))
    return analytic.tc5(L, T)


def describe(name, m, K, T, dt):
    if WORKLOADS[name][0] < 0:
        return (f"{name}: planar hex {m.nCells} cells x {K} level, no tracers, RK4 dt={dt:g}s, manufactured-solution "
                "source terms, Del2+Del4 on velocity")
    return (f"{name}: icosahedral {m.nCells} cells x {K} levels, {T} tracers, RK4 dt={dt:g}s, "
            "Del2+Del4 on velocity and tracers, TC5-style state")


def host_threads():
    """Threads the CPU arm may use: the cores this process is allowed on (torch.distributed.run exports
    OMP_NUM_THREADS=1 to its workers, which is not the machine's core count)."""
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


class ClockSampler:
    """nvidia-smi clocks during the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.p = None
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(index), f"--query-gpu={self.Q}",
                                       "--format=csv,noheader,nounits", "-lms", "100"],
                                      stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except OSError:
            pass

    def stop(self):
        if self.p is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.p.terminate()
        try:
            out = self.p.communicate(timeout=5)[0]
        except subprocess.TimeoutExpired:
            self.p.kill()
            out = self.p.communicate()[0]
        sm, mx, reasons = [], [], set()
        names = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")
        for line in out.splitlines():
            f = [x.strip() for x in line.split(",")]
            if len(f) < 6:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1]))
            except ValueError:
                continue
            for n, v in zip(names, f[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def cpu_reference_rate(name, steps, warmup, full_depth):
    """The reference's CPU implementation of the same step on the host cores: the reference's own sources over
    the host Kokkos stand-in with OpenMP (oracle/_ref/libomega_ref_fast.so) when built, else the plain-C port
    (oracle/).  full_depth: all K levels (the --impl reference arm); otherwise a bounded sample, the full
    horizontal mesh with `ks` of the K levels (levels are independent in Omega-0, cost is linear in them)."""
    nthr = host_threads()
    os.environ["OMP_NUM_THREADS"] = str(nthr)      # before libgomp initialises (first load of the oracle library)
    os.environ.setdefault("OMP_PROC_BIND", "spread")
    os.environ.setdefault("OMP_PLACES", "threads")
    import ctypes
    from omega_b200 import capi, decomp
    from tests import oracle_lib
    m, cfg, K, T, dt, ks = build_problem(name, 1)
    if full_depth:
        # footprint of the reference's arrays ((177 + 24 T) doubles per cell-layer incl. HTracersEdge)
        need = m.nCells * K * 8.0 * (60 + 9 * T) * 1.3
        avail = 0.0
        try:
            for ln in open("/proc/meminfo"):
                if ln.startswith("MemAvailable"):
                    avail = float(ln.split()[1]) * 1024
        except OSError:
            pass
        if avail and need > 0.8 * avail:
            full_depth = False
    kl = K if full_depth else min(ks, K)
    L = decomp.build_all(m, 1, kl)[0]
    h, u, tr = initial_state(name, L, cfg, T)
    if os.path.exists(oracle_lib.REF_FAST) and kl % 8 == 0:
        path, kind = oracle_lib.REF_FAST, "reference"
    else:
        if not os.path.exists(oracle_lib.FAST):
            oracle_lib.build_oracle()
        path, kind = oracle_lib.FAST, "port"
    o = oracle_lib.Oracle(L, cfg, T, lib_path=path)
    try:
        # libgomp may already be initialised (torch loads it) with torchrun's OMP_NUM_THREADS=1: set the
        # thread count of this thread's parallel regions explicitly, then read back what will be used
        gomp = ctypes.CDLL("libgomp.so.1")
        gomp.omp_set_num_threads(nthr)
        used = int(gomp.omp_get_max_threads())
    except OSError:
        used = 1
    o.make_state(h, u, tr)
    co = capi.rk_coeffs(capi.RK4, dt)
    for _ in range(warmup):
        o.do_step(capi.RK4, co, compat=1)
    t0 = time.perf_counter()
    for _ in range(steps):
        o.do_step(capi.RK4, co, compat=1)
    el = time.perf_counter() - t0
    o.close()
    rate = m.nCells * kl * steps / el
    what = "all" if kl == K else f"{kl} of the"
    return {"value": rate, "unit": "cell-layer updates/s", "cores": used, "kind": kind,
            "sample": f"{m.nCells} cells x {what} {K} levels, {T} tracers, {warmup} warm-up + {steps} timed RK4 step(s), "
                      f"{el / steps:.2f} s/step, OpenMP threads {used} (omp_get_max_threads; {nthr} cores available)",
            "extrapolated": kl != K}, el / steps, dt, m, kl


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    level, K, T, _, _ = WORKLOADS[args.workload]
    # a full-depth config-3 step costs ~25 s on the box's host cores: at most 3 timed steps after one warm-up
    big = level >= 8
    steps, warm = (max(1, min(args.steps, 3)), min(args.warmup, 1)) if big else (args.steps, args.warmup)
    cb, s_per_step, dt, m, kl = cpu_reference_rate(args.workload, steps, warm, full_depth=True)
    full = s_per_step * K / kl
    line = {"impl": "reference", "metric": "cell-layer updates/s", "value": cb["value"], "unit": "cell-layer updates/s",
            "n_gpus": args.gpus, "steps": steps, "warmup": warm, "ms_per_step": full * 1e3, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "sypd": dt / (365.0 * full),
            "config": {"workload": describe(args.workload, m, K, T, dt), "path": "reference CPU sources, OpenMP host shim",
                       "levels_run": kl},
            "cpu_baseline": cb,
            "e2e": {"value": cb["value"], "unit": "cell-layer updates/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="config3", choices=list(WORKLOADS))
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--unfused", action="store_true", help="time the kernel-per-term path instead")
    ap.add_argument("--precision", default="double", choices=["double", "single"],
                    help="single: libomega_b200_sp.so (Real = float as in the reference's OMEGA_SINGLE_PRECISION build; "
                         "untiled fused kernels); not the BASELINE metric, which is fp64")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)

    import torch
    from omega_b200 import analytic, capi, decomp, model

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (the B200 path has no CPU fallback)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)
    W = max(args.warmup, 3)
    K_steps = args.steps

    m, cfg, K, T, dt, ks = build_problem(args.workload, world)
    co = capi.rk_coeffs(capi.RK4, dt)
    if world == 1:
        L = decomp.build_all(m, 1, K)[0]
        halo = None
    else:
        from omega_b200 import parallel
        # HaloWidth 4: with the biharmonic terms on, the default 3 makes the reference algorithm itself
        # partition dependent (tests/test_multirank_cpu.py::test_reference_needs_halo_width_4_with_del4)
        L, halo_desc = parallel.local_problem(m, world, rank, K, halo_width=4)
    h, u, tr = initial_state(args.workload, L, cfg, T)
    sp = args.precision == "single"
    mesh = model.HorzMesh(L, reorder=True, device=dev, lib=capi.load_library(capi.LIB_PATH_SP) if sp else None)
    rb = mesh.real_bytes
    st = model.Stepper(mesh, cfg, T)
    state = model.OceanState(mesh, T)
    state.copyToDevice(h, u, tr if T else None)
    if world > 1:
        halo = parallel.HaloExchanger(mesh, halo_desc, T)
        halo.init_comm(world, rank)      # NCCL communicator of the library; torch.distributed only carries the id

    fused = not args.unfused

    def step():
        if world == 1:
            st.doStep(state, capi.RK4, co, fused=fused, compat=True)
        else:
            halo.do_step(st, state, capi.RK4, co, fused=fused, compat=True)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    for _ in range(W):
        step()
    barrier()
    n0 = st.launch_count()
    if halo is not None:
        halo.set_profiling(True)
    sampler = ClockSampler(local) if rank == 0 else None
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(K_steps):
        step()
    e1.record()
    barrier()
    ms = e0.elapsed_time(e1)
    clocks = sampler.stop() if sampler else None
    launches = st.launch_count() - n0
    halo_raw = None
    if world > 1:
        halo_raw = halo.profile()      # CUDA events around pack .. unpack on the exchange stream, timed region only
        halo.set_profiling(False)
    # per-kernel breakdown: a second, shorter pass with CUDA events around every launch of the fused pipeline.
    # Kept out of the timed region: the events sit between the launches and switch off the overlap of
    # independent consecutive kernels (programmatic stream serialization, csrc/fused.cu)
    import ctypes as C
    kms, kcnt = (C.c_double * 5)(), (C.c_int * 5)()
    KP = min(K_steps, 5)
    st.lib.omega_stepper_set_profiling(st.h, 1 if fused else 0)
    p0, p1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    p0.record()
    for _ in range(KP):
        step()
    p1.record()
    barrier()
    ms_prof = p0.elapsed_time(p1)
    st.lib.omega_stepper_get_profile(st.h, kms, kcnt, 5)
    st.lib.omega_stepper_set_profiling(st.h, 0)
    halo_prof = None
    if world > 1:
        hms, hrounds, hbytes = halo_raw
        t = torch.tensor([ms, hms / max(hrounds, 1), float(hbytes) / max(hrounds, 1)], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t[0].item())
        halo_prof = (float(t[1].item()), float(t[2].item()), halo.tile_classes())
    # sanity: the state must still be finite after W + K steps
    hh = state.getLayerThickness(0)
    assert bool(torch.isfinite(hh).all()), "state blew up"
    # checksum over OWNED elements, reduced over ranks: sum(A h), sum(A h phi_l) per tracer, sum(u^2).  The result
    # of W + K steps must not depend on the partition (HaloWidth 4, tests/test_multirank_cpu.py), so the N-GPU
    # values equal the one-GPU values to rounding of the reduction order: a wrong exchange list shows here.
    pc, _ = mesh.perm(capi.ON_CELL)
    nco, neo = mesh.NOwned[0], mesh.NOwned[1]
    area = torch.from_numpy(np.ascontiguousarray(np.asarray(L.AreaCell)[pc][:nco])).to(dev)
    ah = area[:, None] * hh[:nco]
    cs = [ah.sum()]
    cs += [(ah * state.getTracers(0)[l, :nco]).sum() for l in range(T)]
    cs.append((state.getNormalVelocity(0)[:neo] ** 2).sum())
    cs = torch.stack(cs)
    if world > 1:
        dist.all_reduce(cs, op=dist.ReduceOp.SUM)
    checksum = [float(x) for x in cs.cpu()]

    ncl = m.nCells * K                       # cell-layers of the whole job
    t_step = ms / 1e3 / K_steps
    value = ncl / t_step
    peaks = {}
    pk_path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(pk_path):
        peaks = json.load(open(pk_path))
    peak, peak_src = (peaks["hbm_gbs"], "measured") if "hbm_gbs" in peaks else (6650.0, "fallback")
    alg_step = 16.0 * rb * (4 + T)           # bytes per cell-layer per RK4 step (128 (4 + T) in fp64)

    # ---- e2e: HOST buffers in, HOST buffers out, every step -----------------
    # The public API a driver calls: omega_field_to_device (pinned host -> device order),
    # omega_stepper_do_step, omega_field_to_host_async.  Every step uploads its full input state
    # and downloads its full result; the three legs of consecutive steps run on three streams over
    # three device state slots (upload of step i+1 | step i | download of step i-1), so PCIe works
    # in both directions while the GPU computes.  The timed region covers all three legs of every
    # counted step and ends with a device synchronisation.
    e2e = None
    if not args.no_e2e and world == 1:
        hp = [torch.from_numpy(np.ascontiguousarray(a, mesh.np_dtype)).pin_memory() for a in ((h, u, tr) if T else (h, u))]
        NS = 3
        slots = [state] + [model.OceanState(mesh, T) for _ in range(NS - 1)]
        outp = [[torch.empty_like(x).pin_memory() for x in hp] for _ in range(2)]
        s_in, s_cmp, s_out = (torch.cuda.Stream(device=dev) for _ in range(3))
        ev_in = [torch.cuda.Event() for _ in range(NS)]
        ev_cmp = [torch.cuda.Event() for _ in range(NS)]
        ev_out = [torch.cuda.Event() for _ in range(NS)]
        ne = min(K_steps, 40)   # the pipeline's fill and drain (two extra legs) are inside the timed region: (ne + 2) / ne

        def run_pipeline(n):
            for i in range(n + 2):
                if i < n:                       # upload the input state of step i
                    sl = slots[i % NS]
                    with torch.cuda.stream(s_in):
                        if i >= NS:
                            s_in.wait_event(ev_out[i % NS])   # the slot's previous result has left
                        mesh.to_device(capi.ON_CELL, hp[0].numpy(), sl.getLayerThickness(0), sync=False)
                        mesh.to_device(capi.ON_EDGE, hp[1].numpy(), sl.getNormalVelocity(0), sync=False)
                        if T:
                            mesh.to_device(capi.ON_CELL, hp[2].numpy(), sl.getTracers(0), sync=False)
                        ev_in[i % NS].record(s_in)
                j = i - 1
                if 0 <= j < n:                  # step j
                    with torch.cuda.stream(s_cmp):
                        s_cmp.wait_event(ev_in[j % NS])
                        st.doStep(slots[j % NS], capi.RK4, co, fused=fused, compat=True)
                        ev_cmp[j % NS].record(s_cmp)
                j = i - 2
                if 0 <= j < n:                  # download the result of step j
                    sl, o = slots[j % NS], outp[j % 2]
                    with torch.cuda.stream(s_out):
                        s_out.wait_event(ev_cmp[j % NS])
                        mesh.to_host(capi.ON_CELL, sl.getLayerThickness(0), o[0].numpy(), sync=False)
                        mesh.to_host(capi.ON_EDGE, sl.getNormalVelocity(0), o[1].numpy(), sync=False)
                        if T:
                            mesh.to_host(capi.ON_CELL, sl.getTracers(0), o[2].numpy(), sync=False)
                        ev_out[j % NS].record(s_out)
            torch.cuda.synchronize()

        run_pipeline(2)
        t0 = time.perf_counter()
        run_pipeline(ne)
        te = (time.perf_counter() - t0) / ne
        assert bool(np.isfinite(outp[(ne - 1) % 2][0].numpy()).all())
        nbytes = sum(x.numel() * rb for x in hp)
        e2e = {"value": ncl / te, "unit": "cell-layer updates/s", "h2d_bytes_per_step": nbytes,
               "d2h_bytes_per_step": nbytes, "ms_per_step": te * 1e3, "steps": ne,
               "api": "omega_field_to_device x3 | omega_stepper_do_step | omega_field_to_host_async x3 per step, pinned host "
                      "state, three streams over three device state slots (upload i+1, step i, download i-1 overlap)"}
        del slots[1:]

    if world > 1:
        bytes_round = halo.bytes_per_round
        dist.barrier()
        dist.destroy_process_group()
    if rank != 0:
        return
    names = ["fA_cell_vertex_aux", "fB_del2_aux", "fT2_tracer_del2", "fC_cell_thick_tracers", "fE_edge_velocity"]
    kern = {n: {"ms_per_launch": kms[i] / kcnt[i], "launches": kcnt[i], "share": kms[i] / ms_prof}
            for i, n in enumerate(names) if kcnt[i]}
    roofline = None
    if fused and kcnt[3]:
        # dominant kernel fC: per launch it owns the (1+T) cell-centred fields of one stage
        # = (3+5+5+3)/4 * 8 * (1+T) bytes per cell-layer on average
        cl_local = mesh.NAll[0] * K
        alg_fc = 4.0 * rb * (1 + T) * cl_local
        alg_stage = [float(rb) * n * (1 + T) * cl_local for n in (3, 5, 5, 3)]   # per RK4 stage
        ach = alg_fc / (kms[3] / kcnt[3] * 1e-3) / 1e9
        # DRAM bytes per launch of that kernel from the committed ncu --set full capture (profiles/)
        traffic = None
        tr_path = os.path.join(ROOT, "profiles", "traffic.json")
        if os.path.exists(tr_path) and args.workload == "config3" and world == 1 and not sp:
            traffic = json.load(open(tr_path)).get("k_fCT_dram_bytes_per_launch")
        path = st.last_path()
        kname = ("k_fCT2 (tiled cell pass: thickness + all tracer tendencies + RK updates)" if path & 4 else
                 "k_fCT (tiled cell pass without tracer hyperdiffusion)" if path & 2 else
                 "k_fC_center (untiled cell pass: level count not a multiple of 16)")
        roofline = {"bound": "hbm", "kernel": kname,
                    "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
                    "traffic": traffic, "peak_source": peak_src,
                    "algorithmic_bytes_per_launch": alg_fc, "algorithmic_bytes_by_stage": alg_stage,
                    "traffic_stage": 1 if traffic else None,
                    "traffic_over_algorithmic": traffic / alg_stage[1] if traffic else None}
    ach_step = alg_step * ncl / world / t_step / 1e9
    line = {"metric": "cell-layer updates/s", "value": value, "unit": "cell-layer updates/s", "n_gpus": world,
            "steps": K_steps, "warmup": W, "ms_per_step": t_step * 1e3, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f32" if sp else "f64", "data": "synthetic",
            "sypd": dt / (365.0 * t_step),
            "config": {"workload": describe(args.workload, m, K, T, dt),
                       "path": ("fused" if fused else "kernel-per-term") + (", single-precision build" if sp else ""), "parallelism": f"domain decomposition x{world}",
                       "l2": ("state per step (>= 5 GB) exceeds the 126 MB L2; no flush needed" if ncl * (4 + T) * 8 > 2.5e8
                              else "state fits L2 (launch-bound workload); no flush")},
            "checksum": {"sum_A_h": checksum[0], "sum_A_h_phi": checksum[1:-1], "sum_u2": checksum[-1],
                         "what": "owned elements, summed over ranks, after warmup + steps"},
            "clocks": clocks, "gpu_launches": launches, "e2e": e2e, "roofline": roofline,
            "roofline_step": {"bound": "hbm", "achieved": ach_step, "peak": peak, "unit": "GB/s", "frac": ach_step / peak,
                              "algorithmic_bytes_per_cell_layer_step": alg_step, "per_gpu": True},
            "kernels": kern,
            "kernels_pass": {"steps": KP, "ms_per_step": ms_prof / KP,
                             "what": "separate pass after the timed region, CUDA events around every launch"}}
    if world > 1:
        # two exchange rounds per RK4 step; bytes one rank sends per round, against NVLink 900 GB/s/direction
        hms, hb, (nb, nown, ntl) = halo_prof
        # bytes one rank moves per round (sent + received) / time of the round (pack .. unpack, device events,
        # max over ranks) against NVLink: 900 GB/s nominal per direction, i.e. send and receive each at most that
        gbs = hb / 2 / (hms * 1e-3) / 1e9 if hms > 0 else None
        line["halo"] = {"rounds_per_step": 2, "send_bytes_per_round_rank0": bytes_round, "halo_width": 4,
                        "ms_per_round": hms, "bytes_per_round_max_rank": hb,
                        "achieved_gbs_per_direction": gbs, "frac_of_nvlink_900": gbs / 900.0 if gbs else None,
                        "tiles_rank0": {"boundary": nb, "owned": nown, "all": ntl},
                        "transport": "library: one pack kernel + grouped ncclSend/ncclRecv + one unpack kernel per round "
                                     "(csrc/halo.cu), on its own stream behind the boundary tiles, overlapped with the "
                                     "interior tiles"}
    if world == 1 and not args.no_cpu_baseline and not sp:
        cb, *_ = cpu_reference_rate(args.workload, 2, 1, full_depth=False)
        line["cpu_baseline"] = cb
    else:
        line["cpu_baseline"] = None
    print(json.dumps(line))


if __name__ == "__main__":
    main()
