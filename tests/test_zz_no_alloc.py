"""GPU: mb2_dot_noalloc, the allocation-free entry point (no counterpart in the reference; added
for callers that cannot call cudaMalloc). It runs the same deterministic reduction as the norm that
mb2_newton_residual_device returns (NonlinearSolid::Mult's residual norm, operators/nonlinear_solid.cpp:172-205), with the
block partials in a file-scope device array instead of an allocated buffer. Checked through the C ABI against that existing
entry point: bit-equal (the tests' bar for this reduction against itself; 1e-14 relative against numpy), on two calls in a
row on different vectors, then on the first vector again (the shared scratch keeps the previous call's contents).
This file is collected last on purpose."""
import ctypes as C

import numpy as np
import pytest

from tests.cases import setup_ctx
from tests.test_gpu_newton import SWEEP_CASES, load_sweeps_golden

pytestmark = pytest.mark.gpu


def _r_ptr(ctx):
    u, r, A = C.c_void_p(), C.c_void_p(), C.c_void_p()
    ctx._ck(ctx.lib.mb2_device_pointers(ctx.h, C.byref(u), C.byref(r), C.byref(A)))
    return r.value


def test_dot_noalloc_equals_the_newton_norm_twice_in_a_row(ctx_factory):
    import torch
    D, W = load_sweeps_golden(SWEEP_CASES[0])
    ctx = setup_ctx(ctx_factory(), D, "neohookean")
    ctx.set_mass(W["mass"]); ctx.set_viscosity(W["visc"]); ctx.set_rhs(W["rhs"]); ctx.set_essential_dofs(W["ess"])
    NV = D["NV"]
    # two different residual vectors with the norms the EXISTING entry point reports for them
    nrm1 = ctx.newton_residual_device(W["x"], W["v"], W["a"], W["fac0"], W["fac1"])
    y1 = ctx.get_r()
    a2 = np.ascontiguousarray(0.5 * W["a"])
    nrm2 = ctx.newton_residual_device(W["x"], W["v"], a2, W["fac0"], W["fac1"])
    y2 = ctx.get_r()
    assert nrm1 != nrm2 and nrm1 > 0 and nrm2 > 0
    d1 = torch.from_numpy(y1).cuda()
    d2 = torch.from_numpy(y2).cuda()
    out = torch.full((3,), -1.0, dtype=torch.float64, device="cuda")
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    # two calls in a row, then the first vector again: no host synchronisation in between (one stream orders them)
    ctx.dot_noalloc(d1.data_ptr(), d1.data_ptr(), NV, out.data_ptr())
    ctx.dot_noalloc(d2.data_ptr(), d2.data_ptr(), NV, out.data_ptr() + 8)
    ctx.dot_noalloc(d1.data_ptr(), d1.data_ptr(), NV, out.data_ptr() + 16)
    ctx.sync()
    print("\nfree device memory before / after the three calls: %d / %d bytes" % (free0, torch.cuda.mem_get_info()[0]))
    o = out.cpu().numpy()
    assert np.sqrt(o[0]) == nrm1 and np.sqrt(o[1]) == nrm2 and o[2] == o[0]
    assert abs(o[0] - y1 @ y1) <= 1e-14 * o[0] and abs(o[1] - y2 @ y2) <= 1e-14 * o[1]
    # a mixed product on the library's own resident residual (device pointer from mb2_device_pointers), a short vector and
    # an empty one: stale partials of the longer calls above must not leak in
    ctx.dot_noalloc(_r_ptr(ctx), d1.data_ptr(), NV, out.data_ptr())
    ctx.dot_noalloc(d1.data_ptr(), d2.data_ptr(), 7, out.data_ptr() + 8)
    ctx.dot_noalloc(d1.data_ptr(), d2.data_ptr(), 0, out.data_ptr() + 16)
    ctx.sync()
    o = out.cpu().numpy()
    assert abs(o[0] - y2 @ y1) <= 1e-14 * abs(y2 @ y1) and abs(o[1] - y1[:7] @ y2[:7]) <= 1e-14 * max(abs(y1[:7] @ y2[:7]), 1e-300)
    assert o[2] == 0.0


def test_dot_noalloc_spans_many_blocks_and_rejects_sizes_beyond_its_scratch(ctx_factory):
    import torch
    from mimi_b200 import capi
    ctx = ctx_factory()
    n = 3 * 4096 * 257 + 11                                  # 772 block partials: the second reduction stage loops
    g = torch.Generator(device="cpu").manual_seed(5)
    a = torch.randn(n, dtype=torch.float64, generator=g)
    b = torch.randn(n, dtype=torch.float64, generator=g)
    da, db = a.cuda(), b.cuda()
    out = torch.zeros(2, dtype=torch.float64, device="cuda")
    torch.cuda.synchronize()
    ctx.dot_noalloc(da.data_ptr(), db.data_ptr(), n, out.data_ptr())
    ctx.dot_noalloc(da.data_ptr(), db.data_ptr(), n, out.data_ptr() + 8)
    ctx.sync()
    o = out.cpu().numpy()
    ref = float(a.numpy() @ b.numpy())
    assert o[0] == o[1] and abs(o[0] - ref) <= 1e-12 * np.sqrt(float(a @ a) * float(b @ b))
    MAXN = 64 * 1024 * 1024
    with pytest.raises(capi.Mb2Error, match="exceeds MB2_NOALLOC_MAX_N"):
        ctx.dot_noalloc(da.data_ptr(), db.data_ptr(), MAXN + 1, out.data_ptr())
    with pytest.raises(capi.Mb2Error, match="null pointer"):
        ctx.dot_noalloc(0, db.data_ptr(), 8, out.data_ptr())
