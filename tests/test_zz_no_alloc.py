"""omega_field_to_device_noalloc (include/omega_b200.h): the permuting field upload with its staging block in a
static device array of the library, for callers that may not allocate.  Called through the C ABI and compared
bit for bit with omega_field_to_device (the tolerance of the transfer tests: exact) on calls in a row that share
the static block.  Kept in a file that pytest collects last."""
import ctypes as C

import numpy as np
import pytest

from omega_b200 import capi
from tests import cases

pytestmark = pytest.mark.gpu

MAX_BYTES = 32 << 20     # OMEGA_NOALLOC_MAX_BYTES


def _objects():
    import torch
    from omega_b200 import model
    L, cfg, (h, u, tr), dt = cases.sphere_case(4, 64, 12)      # tracers 12 x 2 563 x 64 x 8 B = 15.7 MB
    mesh = model.HorzMesh(L, reorder=True)
    return torch, model, mesh, (h, u, tr)


def _upload(torch, model, mesh, fn, kind, host, out):
    host = np.ascontiguousarray(host, np.float64)
    nouter = host.shape[0] if host.ndim == 3 else 1
    rc = fn(mesh.h, kind, nouter, host.shape[-1], host.ctypes.data_as(capi.c_dbl_p), model._dptr(out), model._stream())
    torch.cuda.current_stream().synchronize()
    assert rc == 0, mesh.lib.omega_last_error().decode()
    return out


def test_noalloc_upload_equals_the_allocating_upload_on_calls_in_a_row():
    torch, model, mesh, (h, u, tr) = _objects()
    lib = mesh.lib
    rng = np.random.default_rng(7)
    # big field, then a smaller one (stale rows of the first stay behind it in the block), then the big shape
    # again with other values, then the same call twice
    seq = [(capi.ON_CELL, tr), (capi.ON_EDGE, u), (capi.ON_CELL, h), (capi.ON_CELL, rng.standard_normal(tr.shape)),
           (capi.ON_EDGE, u), (capi.ON_EDGE, u)]
    for kind, host in seq:
        assert host.nbytes <= MAX_BYTES
        want = torch.full(host.shape, np.nan, dtype=torch.float64, device="cuda:0")
        got = torch.full(host.shape, np.nan, dtype=torch.float64, device="cuda:0")
        _upload(torch, model, mesh, lib.omega_field_to_device, kind, host, want)
        _upload(torch, model, mesh, lib.omega_field_to_device_noalloc, kind, host, got)
        assert torch.equal(got, want), "static staging block: result differs from omega_field_to_device"
        p, _ = mesh.perm(kind)
        assert np.array_equal(got.cpu().numpy(), np.asarray(host)[..., p, :]), "device order must be host[perm]"
    mesh.close()


def test_noalloc_upload_rejects_a_field_beyond_the_static_block():
    torch, model, mesh, (h, u, tr) = _objects()
    lib = mesh.lib
    rows, K = mesh.NSize[capi.ON_CELL], 64
    nouter = MAX_BYTES // (rows * K * 8) + 1                    # one slice too many; rejected before any access
    dummy = torch.zeros(16, dtype=torch.float64, device="cuda:0")
    hostbuf = np.zeros(16)
    rc = lib.omega_field_to_device_noalloc(mesh.h, capi.ON_CELL, nouter, K, hostbuf.ctypes.data_as(capi.c_dbl_p),
                                           model._dptr(dummy), model._stream())
    assert rc == 2
    msg = lib.omega_last_error().decode()
    assert "exceeds the static staging block" in msg and str(MAX_BYTES) in msg, msg
    # and the largest field that fits is accepted
    fit = MAX_BYTES // (rows * K * 8)
    host = np.random.default_rng(3).standard_normal((fit, rows, K))
    want = torch.empty(host.shape, dtype=torch.float64, device="cuda:0")
    got = torch.empty(host.shape, dtype=torch.float64, device="cuda:0")
    _upload(torch, model, mesh, lib.omega_field_to_device, capi.ON_CELL, host, want)
    _upload(torch, model, mesh, lib.omega_field_to_device_noalloc, capi.ON_CELL, host, got)
    assert torch.equal(got, want)
    mesh.close()


def test_noalloc_upload_on_an_unpermuted_mesh_is_a_plain_copy():
    import torch
    from omega_b200 import model
    L, cfg, (h, u, tr), dt = cases.sphere_case(3, 16, 2)
    mesh = model.HorzMesh(L, reorder=False)
    got = torch.empty(u.shape, dtype=torch.float64, device="cuda:0")
    _upload(torch, model, mesh, mesh.lib.omega_field_to_device_noalloc, capi.ON_EDGE, u, got)
    assert np.array_equal(got.cpu().numpy(), u)
    mesh.close()


def test_noalloc_upload_is_capturable_in_a_cuda_graph():
    """What the entry point is for: under stream capture a cudaMalloc is an error; this call makes none.  Pinned source."""
    torch, model, mesh, (h, u, tr) = _objects()
    lib = mesh.lib
    src = torch.from_numpy(np.ascontiguousarray(u)).pin_memory()
    want = torch.empty(u.shape, dtype=torch.float64, device="cuda:0")
    got = torch.zeros(u.shape, dtype=torch.float64, device="cuda:0")
    _upload(torch, model, mesh, lib.omega_field_to_device, capi.ON_EDGE, u, want)
    _upload(torch, model, mesh, lib.omega_field_to_device_noalloc, capi.ON_EDGE, src.numpy(), got)   # symbol lookup outside capture
    got.zero_()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        rc = lib.omega_field_to_device_noalloc(mesh.h, capi.ON_EDGE, 1, u.shape[-1], C.cast(C.c_void_p(src.data_ptr()), capi.c_dbl_p),
                                               model._dptr(got), model._stream())
    assert rc == 0, lib.omega_last_error().decode()
    for rep in range(2):
        got.zero_()
        src.mul_(1.0)            # same values; the graph reads the pinned block at replay
        g.replay()
        torch.cuda.synchronize()
        assert torch.equal(got, want), f"graph replay {rep}"
    mesh.close()
