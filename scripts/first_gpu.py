"""Ad-hoc GPU probe: FP64 pipe peaks and a first timing of the FD tangent kernel on a reference-dumped mesh."""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mimi_b200 import capi
from oracle import oracle_py as O
from tests.cases import ref_case, setup_ctx

ctx = capi.Context(0)
print("fp64 peak TFLOP/s: dmul+dadd", ctx.measure_fp64_peak(0), " dfma", ctx.measure_fp64_peak(1), flush=True)
n = int(sys.argv[1]) if len(sys.argv) > 1 else 16
t0 = time.time()
D = ref_case("refine", 2, n, "neohookean")
print("reference dump (incl. its own 1-thread assembly) took %.1fs" % (time.time() - t0), flush=True)
setup_ctx(ctx, D, "neohookean")
ctx.set_u(D["u"])
for i in range(6):
    ctx.set_r(None); ctx.set_A(D["mass"])
    ctx.assemble_residual_and_grad_device(1.0)
    t = ctx.last_timing_ms()
    print("res+grad ms: element %.3f gather %.3f total %.3f -> %.3e qp/s" % (t[0], t[1], t[2], D["NE"] * D["nq"] / (t[2] * 1e-3)), flush=True)
A = ctx.get_A(); r = ctx.get_r()
print("bit-exact vs reference: r", np.array_equal(r, D["r2"]), "A", np.array_equal(A, D["A"]), np.abs(A - D["A"]).max())
for i in range(3):
    ctx.set_r(None)
    ctx.assemble_residual_device()
    t = ctx.last_timing_ms()
    print("residual ms: element %.3f gather %.3f" % (t[0], t[1]))
