#!/usr/bin/env python
"""bench.py -- quadrature points / second of one residual + tangent CSR assembly (BASELINE.json metric).

Workload at N=1: configs[2] of BASELINE.json = C3 of SURVEY.md section 8: single-patch unit cube, B-spline p=2,
100^3 = 1e6 elements, 64 qp each, compressible NeoHookean (E=2100, nu=0.3), u_i = 1e-3 (U_i - 0.5) from
splitmix64(12345 + i), grad_factor = 1. A "step" is one AddDomainResidualAndGrad-equivalent:
element FD-tangent kernel + deterministic gather into r[NV] and the CSR values A[nnz].

  python bench.py [--gpus N] [--steps K] [--warmup W] [--n SPANS] [--impl reference]

Contract keys: value = device-resident step (u, tables, CSR in HBM); e2e = the host-buffer C-ABI call
mb2_add_domain_residual_and_grad (u H2D, r and A D2H from pinned host memory inside the timed region);
roofline = FP64 CUDA-core pipe of the element kernel; cpu_baseline = the unmodified reference
(oracle/_ref/mimi_ref, OpenMP, all host cores) on a bounded sample.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
MG_LATE = -1   # --mg-late
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = json.load(open(os.path.join(ROOT, "BASELINE.json")))["metric"] if os.path.exists(os.path.join(ROOT, "BASELINE.json")) else \
    "quadrature pts/sec for residual+tangent CSR assembly (3D p=2 NeoHookean) @1/2/4/8 GPU"
UNIT = "qp/s"


# flops per quadrature point, counted mul/add/div = 1 each
STRESS_FLOPS = {"neohookean": 204, "stvk": 166}


def flops_per_qp(p, dim=3, material="neohookean"):
    nd = (p + 1) ** dim
    tdof = dim * nd
    F = 2 * dim * dim * nd + dim           # MultAtB + AddDiagonal
    stress = STRESS_FLOPS[material]        # NeoHookean: det 14, mu/J 1, FF^T 54, coeff 3, axpby 27, inverse 42, sigma F^-T 54, J* 9; StVK 166
    acc = dim * dim * (1 + 2 * nd) + 1     # AddMult_a_ABt + (w * detJ)
    algorithmic = (tdof + 1) * (F + stress + acc)
    # executed by the CUDA kernel: a perturbed dof only changes one row of F (the others are bit-identical
    # to the base F and are re-used), and w*detJ is folded into the tables at setup.
    F_row = 2 * dim * nd + 1
    executed = (F + stress + acc - 1) + tdof * (F_row + stress + acc - 1)
    fd = 2 * tdof * tdof                   # (R+ - R) * (1/h), per element -> caller divides by nq
    return algorithmic, executed, fd


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md)."""

    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _read(self):
        for ln in self.proc.stdout:
            self.lines.append(ln.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        self.t.join(timeout=2)
        sm, mx, pw, reasons = [], [], [], set()
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1])); pw.append(float(f[2]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        # under load = samples at or above the median power draw
        med_p = float(np.median(pw))
        load = [s for s, p_ in zip(sm, pw) if p_ >= med_p] or sm
        return {"sm_mhz": float(np.median(load)), "sm_max_mhz": float(max(mx)), "power_w_max": float(max(pw)),
                "samples": len(sm), "reasons": sorted(reasons)}


# ----------------------------------------------------------------------------------------------------
# reference arm / cpu_baseline: the UNMODIFIED reference compiled into oracle/_ref/mimi_ref (test
# infrastructure; this is the one place bench.py executes anything under oracle/).
# ----------------------------------------------------------------------------------------------------
def run_reference_sample(n, reps, nthreads, warmup=0, p=2, material="neohookean"):
    from mimi_b200 import meshgen
    from oracle import oracle_py as O
    if not O.have_ref():
        raise RuntimeError("oracle/_ref/mimi_ref is not present (build it with `make -f oracle/Makefile.ref` "
                           "where /root/reference exists; it ships to the GPU box with the snapshot)")
    with tempfile.TemporaryDirectory() as tmp:
        mesh = os.path.join(tmp, "m.mesh")
        meshgen.write_box_patch(mesh, p, (n, n, n))
        T = O.run_ref_time(mesh, 0, 0, material, nthreads, reps, warmup)
    qp = T["NE"] * T["nq"]
    return {"qp": qp, "NE": T["NE"], "t_mean": T["t_res_grad_mean"], "t_best": T["t_res_grad_best"],
            "t_res_best": T["t_res_best"], "value_mean": qp / T["t_res_grad_mean"], "value_best": qp / T["t_res_grad_best"],
            "t_precompute": T["t_precompute"], "omp_team": int(T.get("omp_team", 0)), "omp_max_threads": int(T.get("omp_max_threads", 0))}


def host_cores():
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    cores = host_cores()
    n = args.ref_n
    reps = args.steps
    t0 = time.time()
    S = run_reference_sample(n, reps, cores, warmup=args.warmup)
    value = S["qp"] / S["t_mean"]
    team = S["omp_team"] or cores
    sample = ("unmodified reference AddDomainResidualAndGrad (MFEM 4.7.1 + mimi integrators, OpenMP, -O3), "
              "p=2 unit cube %d^3 = %d elements (%d qp) per step, OpenMP team of %d threads (%d requested = host cores); "
              "mean of %d steps after %d warm-up" % (n, S["NE"], S["qp"], team, cores, reps, args.warmup))
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": S["t_mean"] * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": "C3 sample: p=2 NeoHookean unit cube %d^3 (bounded sample of the 100^3 workload)" % n,
                       "elements": S["NE"], "qp": S["qp"]},
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": team, "kind": "reference", "sample": sample},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0, "wall_s": time.time() - t0}
    emit(line)
    return 0


# ----------------------------------------------------------------------------------------------------
# GPU arm
# ----------------------------------------------------------------------------------------------------
def build_problem(ctx, n, p, rank=0, world=1, uid=None, material="neohookean", slab=None, nz_total=None):
    """Synthetic C3-style block in MFEM's numbering; tables regenerated on the device. world > 1: the global mesh
    is n x n x (n*world) spans (weak scaling), cut into z-slabs of n layers; rank g computes slab g, owns the nodes
    whose lowest element it computes, and receives the raw element matrices of the next slab's first p layers."""
    from mimi_b200 import capi, meshgen, partition
    t0 = time.time()
    nq1 = (2 * p + 3) // 2 + 1
    if world == 1:
        M = meshgen.patch_connectivity(p, (n, n, n))
        M["n_compute"] = M["NE"]
        owned_rows = None
    else:
        # weak scaling: n layers per rank of an n x n x (n world) patch; strong (slab given): slab layers of n x n x n
        # (nz_total given: that many layers in all, cut into balanced, possibly uneven slabs)
        M = partition.slab_rank_problem(p, (n, n, nz_total or (slab or n) * world), world, rank)
        owned_rows = np.repeat(M["owned"], 3)
    ctx.set_connectivity(3, M["nd"], nq1 ** 3, M["NV"], M["v_dofs"])
    if world > 1:
        ctx.mg_setup(M)
        ctx.mg_init_nccl(rank, world, uid)
    ctx.set_nodes(M["x_ref"], None)
    ctx.generate_patch_tables(p, M["ncp"], M["knots"], M["el_ijk"][:M["n_compute"]])
    nnz = ctx.build_pattern()
    young, poisson = 2100.0, 0.3
    lam = young * poisson / ((1 + poisson) * (1 - 2 * poisson))
    mu = young / (2.0 * (1.0 + poisson))
    ctx.set_material({"neohookean": capi.MB2_MAT_NEOHOOKEAN, "stvk": capi.MB2_MAT_STVK}[material], [lam, mu])
    u = meshgen.synth_u(M["NV"])
    # face x = 0 clamped: exactly 0 (h = 1e-10 branch of the FD rule)
    x = M["x_ref"].reshape(-1, 3)
    u.reshape(-1, 3)[x[:, 0] == 0.0, :] = 0.0
    M.update(u=u, nnz=nnz, nq=nq1 ** 3, setup_s=time.time() - t0, owned_rows=owned_rows)
    return M


def build_problem_c4(ctx, n, rank=0, world=1, uid=None):
    """BASELINE config 4: multi-patch p = 3 StVenantKirchhoff beam, patches of n^3 knot spans glued along x, element-partitioned:
    every rank computes n x n x n/2 elements (N = 1: half a patch; N = 2: one patch; N = 4: two; N = 8: the four patches =
    4e6 elements at n = 100). Lexicographic numbering over the glued control grid (mimi_b200/partition.py beam_rank_problem);
    the parity fixtures with MFEM's own multi-patch numbering are tests/golden/beam4_p3.npz / beam3_p2.npz."""
    from mimi_b200 import capi, meshgen, partition
    t0 = time.time()
    p = 3
    M_patches = max(1, world // 2)
    spans = (n, n, n if world >= 2 else n // 2)
    M = partition.beam_rank_problem(p, spans, M_patches, world, rank)
    nq1 = (2 * p + 3) // 2 + 1
    ctx.set_connectivity(3, M["nd"], nq1 ** 3, M["NV"], M["v_dofs"])
    if world > 1:
        ctx.mg_setup(M)
        ctx.mg_init_nccl(rank, world, uid)
        # the x-interfaces between patches put outgoing elements (every i < p column) all through a rank's element range:
        # hold the NCCL exchange back until the element kernels are done (see mb2_set_option "mg_exchange_late")
        ctx.set_option("mg_exchange_late", MG_LATE if MG_LATE >= 0 else (1 if world > 2 else 0))
    ctx.set_nodes(M["x_ref"], None)
    for (e0, cnt, q) in M["patch_segments"]:
        ctx.generate_patch_tables(p, M["ncp"], M["knots"], M["el_ijk"][e0:e0 + cnt], e0=e0)
    nnz = ctx.build_pattern()
    young, poisson = 2100.0, 0.3
    lam = young * poisson / ((1 + poisson) * (1 - 2 * poisson))
    mu = young / (2.0 * (1.0 + poisson))
    ctx.set_material(capi.MB2_MAT_STVK, [lam, mu])
    u = meshgen.synth_u(M["NV"])
    x = M["x_ref"].reshape(-1, 3)
    u.reshape(-1, 3)[x[:, 0] == 0.0, :] = 0.0
    M.update(u=u, nnz=nnz, nq=nq1 ** 3, setup_s=time.time() - t0, n_patches=M_patches, spans=spans)
    return M


def _code_only(text):
    """Source text without comments, blank lines and trailing blanks: a comment edit must not orphan an ncu capture."""
    import re
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    text = re.sub(r"//[^\n]*", "", text)
    return "\n".join(l.rstrip() for l in text.split("\n") if l.strip())


def source_md5(kernel_key=""):
    """md5 of the kernel's source CODE (comments stripped): a committed ncu capture only speaks for a binary built from the
    same kernel code. p=2: element_fd_kernel (Elem2Cfg .. the next section) + mb2_math.cuh; p=3: element_fd_split_kernel;
    fast mode: mb2_fast.cu + mb2_math.cuh."""
    import hashlib
    csrc = os.path.join(ROOT, "mimi_b200", "csrc")
    h = hashlib.md5()
    if "fast" in kernel_key:
        h.update(_code_only(open(os.path.join(csrc, "mb2_fast.cu")).read()).encode())
    else:
        src = open(os.path.join(csrc, "mb2_assembly.cu")).read()
        a, b = ("struct Elem3Cfg", "// History commit:") if "split" in kernel_key else ("struct Elem2Cfg", "// FD-tangent element kernel, column-split")
        i = src.find(a)
        j = src.find(b, i)
        h.update(_code_only(src[i:j] if i >= 0 and j > i else src).encode())
    h.update(_code_only(open(os.path.join(csrc, "mb2_math.cuh")).read()).encode())
    return h.hexdigest()


def ncu_traffic(kernel_key, launch_elems):
    """DRAM bytes per launch of the dominant kernel from the committed `ncu --set full` capture (profiles/traffic.json,
    written by scripts/ncu_traffic.py from the .ncu-rep): dram__bytes_read.sum + dram__bytes_write.sum of a capture of
    THIS source revision, scaled to the elements one launch of this run processes. No capture of the current sources ->
    null (never a stale constant)."""
    try:
        T = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
    except (OSError, ValueError):
        return None, "no profiles/traffic.json"
    ent = T.get(kernel_key)
    if not ent:
        return None, "no ncu capture of %s in profiles/traffic.json" % kernel_key
    if ent.get("source_md5") != source_md5(kernel_key):
        return None, "profiles/traffic.json holds a capture of another source revision of %s (%s)" % (kernel_key, ent.get("capture"))
    per_el = (ent["dram_read_bytes"] + ent["dram_write_bytes"]) / float(ent["launch_elements"])
    return per_el * launch_elems, "%s: %d-element launch, read %.1f MB + write %.1f MB, scaled to %d elements per launch" % (
        ent.get("capture"), ent["launch_elements"], ent["dram_read_bytes"] / 1e6, ent["dram_write_bytes"] / 1e6, launch_elems)


def nccl_parity_check(rank, world, local, dist):
    """Before any timing at N > 1: a small partitioned residual + tangent assembly over the LIVE NCCL exchange, every
    rank's owned rows compared bit for bit with the unmodified reference's dump (tests/golden/cube_p2_n3_patch.npz:
    27 p=2 elements, NeoHookean). Several assemblies are enqueued back to back without a host sync in between, so a
    sender running ahead of its receiver would show. Fails the run on any mismatch."""
    from mimi_b200 import capi, partition as P
    z = np.load(os.path.join(ROOT, "tests", "golden", "cube_p2_n3_patch.npz"))
    D = {k: z[k] for k in z.files}
    dim, nd, nq, NE, NV, nnz = (int(x) for x in D["meta"])
    uid = [capi.mg_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(uid, src=0)
    R = P.rank_problem(D["v_dofs"], dim, NV, P.element_ranges(NE, world), rank)
    ctx = capi.Context(local)
    loc = R["local_elems"]
    ctx.set_connectivity(dim, nd, nq, NV, R["v_dofs"])
    ctx.upload_quad_tables(D["dN_dX"][loc], D["w"][loc], D["detJ"][loc])
    ctx.mg_setup(R)
    ctx.mg_init_nccl(rank, world, uid[0])
    ctx.build_pattern()
    ctx.set_material(capi.MB2_MAT_NEOHOOKEAN, list(D["lambda_mu"]))
    rows = np.repeat(R["owned"], dim)
    gI = D["I"].astype(np.int64)
    own = np.nonzero(rows)[0]
    idx = np.concatenate([np.arange(gI[v], gI[v + 1]) for v in own]) if own.size else np.zeros(0, np.int64)
    ok = True
    gf = float(D["gf"])
    for rep in range(2):
        ctx.set_u(D["u"]); ctx.set_r(None); ctx.set_A(np.ascontiguousarray(D["mass"][idx]))
        ctx.assemble_residual_and_grad_device(gf)
        ok = ok and np.array_equal(ctx.get_r()[rows], D["r2_neohookean"][rows]) and np.array_equal(ctx.get_A(), D["A_neohookean"][idx])
    # three assemblies in flight without a host sync: A = mass + 3 gf K must equal the sequentially added reference values
    ctx.set_r(None); ctx.set_A(np.ascontiguousarray(D["mass"][idx]))
    for _ in range(3):
        ctx.assemble_residual_and_grad_device(gf)
    K = D["A_neohookean"][idx] - D["mass"][idx]
    A3 = ctx.get_A()
    ok = ok and bool(np.allclose(A3, D["mass"][idx] + 3.0 * K, rtol=1e-12, atol=1e-12 * float(np.abs(K).max() if K.size else 1.0)))
    ctx.close()
    import torch
    t = torch.tensor([int(ok)], device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MIN)
    return {"ranks": world, "ok": bool(int(t) == 1), "fixture": "tests/golden/cube_p2_n3_patch.npz (reference dump), owned rows of r and CSR values bit-equal on every rank over the live NCCL exchange"}


def timed_steps(ctx, stream, steps, warmup, barrier, world, dist, sampler=None):
    """W untimed + K timed device-resident assemblies bracketed by barrier + synchronize; CUDA events on the launching
    stream; max over ranks."""
    import torch
    for _ in range(warmup):
        ctx.assemble_residual_and_grad_device(1.0)
    ctx.sync()
    barrier()
    if sampler:
        sampler.start()
    t_el, t_sc, launches = [], [], 0
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    ev0.record(stream)
    for _ in range(steps):
        ctx.assemble_residual_and_grad_device(1.0)
        el, sc, tot = ctx.last_timing_ms()      # CUDA events on the library's stream (synchronises on the last)
        t_el.append(el); t_sc.append(sc)
        launches += ctx.last_launch_count()
    ev1.record(stream)
    barrier()
    wall = time.perf_counter() - t0
    clocks = sampler.stop() if sampler else None
    dev_ms = float(ev0.elapsed_time(ev1))
    if world > 1:
        tt = torch.tensor([dev_ms, wall], device="cuda", dtype=torch.float64)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        dev_ms, wall = float(tt[0]), float(tt[1])
    return {"ms_per_step": dev_ms / steps, "wall": wall, "t_el": t_el, "t_sc": t_sc, "launches": launches, "clocks": clocks}


def gpu_arm(args):
    import torch
    import torch.distributed as dist
    from mimi_b200 import capi

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py needs a CUDA device: libmimi_b200 has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def maxr(x):
        if world > 1:
            tt = torch.tensor([x], device="cuda", dtype=torch.float64)
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            return float(tt[0])
        return x

    parity = None
    if world > 1:
        parity = nccl_parity_check(rank, world, local, dist)
        if not parity["ok"]:
            raise RuntimeError("partitioned assembly over NCCL differs from the reference's golden vectors: not timing a wrong result")

    c4 = args.workload == "c4"
    if c4:
        args.p, args.material = 3, "stvk"
    n, p = args.n, args.p
    stream = torch.cuda.Stream()            # the library launches on this stream, so torch events see its work
    ctx = capi.Context(local, stream=stream.cuda_stream)
    ctx.set_option("elem_variant", args.variant)
    ctx.set_option("park_acc", args.park)
    ctx.set_option("res_variant", args.res_variant)
    fast = args.tangent == "analytic"
    if fast:
        ctx.set_option("tangent_mode", 1)          # MB2_TANGENT_ANALYTIC (secondary line: the headline stays the parity mode)
        ctx.set_option("fast_gemm", args.fast_gemm)
    uid = [None]
    if world > 1:
        uid = [capi.mg_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(uid, src=0)
    M = build_problem_c4(ctx, n, rank, world, uid[0]) if c4 else build_problem(ctx, n, p, rank, world, uid[0], args.material)
    NE, nq, NV, nnz = M["n_compute"], M["nq"], M["NV"], M["nnz"]     # NE = elements this rank computes
    ctx.set_u(M["u"])
    ctx.set_r(None)
    ctx.set_A(None)
    ctx.sync()

    # FP64 pipe peak of this very GPU (MEASURED_PEAKS.json carries no FP64 figure): DMUL+DADD pairs, the
    # instruction mix a bit-exact (non-FMA) kernel is allowed to issue.
    peak_nofma = ctx.measure_fp64_peak(0)
    peak_fma = ctx.measure_fp64_peak(1)
    ctx_peaks = {"dmma": ctx.measure_fp64_peak(2)}

    # ---- device-resident steps (weak scaling: n^3 elements per GPU) ----
    T = timed_steps(ctx, stream, args.steps, args.warmup, barrier, world, dist, ClockSampler(local))
    ms_per_step, wall, t_el, t_sc, launches, clocks = T["ms_per_step"], T["wall"], T["t_el"], T["t_sc"], T["launches"], T["clocks"]
    total_qp = float(NE) * nq * world
    value = total_qp / (ms_per_step * 1e-3)

    # ---- residual-only pass (SURVEY 8(d) "also report residual-only qp/s"): streams the per-qp tables once, HBM-bound
    res_only = None
    if not args.no_res_only:
        for _ in range(2):
            ctx.assemble_residual_device()
        ctx.sync()
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(max(args.steps, 3)):
            ctx.assemble_residual_device()
        e1.record(stream)
        barrier()
        ro_ms = maxr(float(e0.elapsed_time(e1)) / max(args.steps, 3))
        tdof_ = 3 * M["nd"]
        ro_bytes = float(NE) * (8.0 * nq * (tdof_ + 1) + 4 * tdof_ + 8 * tdof_ + 8 * tdof_)   # tables + v_dofs + u gather + R_e
        res_only = {"value": float(NE) * nq * world / (ro_ms * 1e-3), "unit": UNIT, "ms_per_step": ro_ms,
                    "algorithmic_gb_per_gpu": ro_bytes / 1e9, "achieved_gbs_per_gpu": ro_bytes / 1e9 / (ro_ms * 1e-3)}

    # ---- end to end through the reference-facing C-ABI calls, pinned HOST buffers, copies inside the timed region ----
    # e2e           : the Newton-iteration assembly with the CSR consumed on the device (the only consumer that exists at
    #                 every N: a host mfem::SparseMatrix cannot index N x 1.15e9 entries with its int nnz) --
    #                 mb2_newton_residual_and_grad_device(x, v, a, fac0, fac1): x, a H2D; y = M a + R(x + fac0 a) and
    #                 J = M + fac0 K stay resident for mb2_newton_solve_gmres_device; y D2H (mb2_get_r).
    # e2e_host_csr  : the host mfem::SparseMatrix consumer with the mass resident (mb2_set_mass once): u, r H2D; r and the
    #                 whole CSR D2H (mb2_domain_residual_and_grad_on_mass).
    # e2e_roundtrip : round 1's number: the CSR values travel both ways (mb2_add_domain_residual_and_grad adds into A).
    e2e = e2e_host = e2e_rt = None
    if not args.no_e2e:
        e2e_steps = max(1, args.e2e_steps if args.e2e_steps > 0 else args.steps)
        one = capi.C.c_double(1.0)
        u_h = torch.from_numpy(M["u"]).pin_memory()
        a_h = torch.zeros(NV, dtype=torch.float64).pin_memory()
        r_h = torch.zeros(NV, dtype=torch.float64).pin_memory()
        A_h = torch.zeros(nnz, dtype=torch.float64).pin_memory()
        up, ap_, rp, Ap = (C_void(t) for t in (u_h, a_h, r_h, A_h))
        # mass, resident: assembled on the device (VectorMassIntegrator, rho = 1), never leaves it
        ctx.set_A(None)
        if not c4:       # c4: the device forms take one patch per call -- a zero mass keeps the same kernels in the timed call
            ctx.assemble_mass_device(p, M["ncp"], M["knots"], M["el_ijk"][:M["n_compute"]])
        ctx.newton_mass_from_A(False)
        ctx.sync()

        def run(fn, k):
            fn()                                  # warm-up
            barrier()
            t0_ = time.perf_counter()
            for _ in range(k):
                fn()
            barrier()
            return maxr((time.perf_counter() - t0_) / k)

        def step_newton():
            ctx._ck(ctx.lib.mb2_newton_residual_and_grad_device(ctx.h, up, None, ap_, one, capi.C.c_double(0.0)))
            ctx._ck(ctx.lib.mb2_get_r(ctx.h, rp))

        s = run(step_newton, e2e_steps)
        e2e = {"value": total_qp / s, "unit": UNIT, "h2d_bytes_per_step": 8 * 2 * NV, "d2h_bytes_per_step": 8 * NV,
               "ms_per_step": s * 1e3, "steps": e2e_steps, "consumer": "device (CSR stays resident for mb2_newton_solve_gmres_device)",
               "call": "mb2_newton_residual_and_grad_device(x, NULL, a, 1.0, 0.0) + mb2_get_r(y): pinned host x, a in; y = M a + R(x + a) out; "
                       "J = M + K(x + a) resident"}

        def step_host():
            ctx._ck(ctx.lib.mb2_domain_residual_and_grad_on_mass(ctx.h, up, one, rp, Ap))

        k2 = max(1, min(e2e_steps, 5))
        s = run(step_host, k2)
        e2e_host = {"value": total_qp / s, "unit": UNIT, "h2d_bytes_per_step": 8 * 2 * NV, "d2h_bytes_per_step": 8 * (NV + nnz),
                    "ms_per_step": s * 1e3, "steps": k2, "consumer": "host mfem::SparseMatrix (mass resident on the device)",
                    "call": "mb2_domain_residual_and_grad_on_mass(u, 1.0, r, A_out) with pinned host buffers"}

        def step_rt():
            ctx._ck(ctx.lib.mb2_add_domain_residual_and_grad(ctx.h, up, one, rp, Ap))

        s = run(step_rt, k2)
        e2e_rt = {"value": total_qp / s, "unit": UNIT, "h2d_bytes_per_step": 8 * (2 * NV + nnz), "d2h_bytes_per_step": 8 * (NV + nnz),
                  "ms_per_step": s * 1e3, "steps": k2, "call": "mb2_add_domain_residual_and_grad(u, 1.0, r, A) with pinned host buffers"}
        del A_h, Ap

    chunk_elems, n_chunks = ctx.plan_info()

    # ---- secondary: the same assembly in the fast tangent mode (MB2_TANGENT_ANALYTIC, NOT bit-comparable),
    # so that the default line carries both modes. p = 2 NeoHookean / StVK only; its own executed-flop count and peak.
    fast_mode = None
    if not fast and not c4 and p == 2 and not args.no_fast:
        ctx.set_option("tangent_mode", 1)
        ctx.set_option("fast_gemm", 0)
        ctx.set_r(None); ctx.set_A(None)
        Tf = timed_steps(ctx, stream, max(2, min(args.steps, 5)), 2, barrier, world, dist)
        ctx.set_option("tangent_mode", 0)
        per_el = 2.0 * 162 * 192 * 27 + 2.0 * 162 * 192 * 3 + nq * (2.0 * 27 * 9 + 600.0) + 2.0 * 81 * 192
        elf = float(np.mean(Tf["t_el"]))
        ach = float(NE) * per_el / (elf * 1e-3) / 1e12
        fast_mode = {"value": total_qp / (Tf["ms_per_step"] * 1e-3), "unit": UNIT, "ms_per_step": Tf["ms_per_step"],
                     "speedup_vs_parity_mode": ms_per_step / Tf["ms_per_step"],
                     "kernel": "element_fast_p2v3_kernel<%s>" % args.material, "kernel_ms": elf,
                     "flop_per_qp_executed": per_el / nq, "achieved": ach, "peak": ctx_peaks["dmma"], "unit_roofline": "TFLOP/s",
                     "frac": ach / ctx_peaks["dmma"],
                     "note": "MB2_TANGENT_ANALYTIC: closed-form tangent, B^T C B on the FP64 tensor cores (mma.sync.m8n8k4.f64); opt-in, "
                             "not bit-comparable: Jacobian within the reference's finite-difference noise, residual unchanged, Newton "
                             "needs the same or fewer iterations (tests/test_gpu_fast.py); peak = mb2_measure_fp64_peak(2) in this run"}
    ctx.close()

    # ---- strong scaling: the SAME 1e6-element block (n^3) cut into N z-slabs ----
    strong = None
    if world > 1 and args.scaling in ("strong", "both") and n // world >= p and not c4:
        ctx2 = capi.Context(local, stream=stream.cuda_stream)
        ctx2.set_option("elem_variant", args.variant)
        uid2 = [capi.mg_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(uid2, src=0)
        M2 = build_problem(ctx2, n, p, rank, world, uid2[0], args.material, nz_total=n)
        ctx2.set_u(M2["u"]); ctx2.set_r(None); ctx2.set_A(None); ctx2.sync()
        T2 = timed_steps(ctx2, stream, args.steps, args.warmup, barrier, world, dist)
        qp2 = float(n) ** 3 * M2["nq"]
        strong = {"value": qp2 / (T2["ms_per_step"] * 1e-3), "unit": UNIT, "ms_per_step": T2["ms_per_step"],
                  "elements_total": n ** 3, "elements_per_gpu": M2["n_compute"],
                  "workload": "the one %d^3-element block cut into %d z-slabs of %s layers" % (
                      n, world, "%d" % (n // world) if n % world == 0 else "%d or %d" % (n // world + 1, n // world))}
        ctx2.close()
    elif world == 1:
        strong = {"value": value, "unit": UNIT, "ms_per_step": ms_per_step, "elements_total": NE, "elements_per_gpu": NE,
                  "workload": "N = 1: identical to the weak value"}

    if rank == 0:
        alg, exe, fd = flops_per_qp(p, 3, args.material)
        el_ms = float(np.mean(t_el))
        qp_rate_kernel = float(NE) * nq / (el_ms * 1e-3)
        exe_qp = exe + fd / nq
        achieved = qp_rate_kernel * exe_qp / 1e12
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except OSError:
            pass
        hbm_peak = peaks.get("hbm_gbs", 6650.0)
        tdof = 3 * M["nd"]
        bytes_el = NE * (8.0 * nq * (tdof + 1) + 4 * tdof + 8 * tdof + 8 * tdof * tdof + 8 * tdof)
        kern = ("element_fd_kernel<3,27,64,%s>" if p == 2 else "element_fd_split_kernel<3,64,125,%s,4>") % args.material
        peak_src = ("mb2_measure_fp64_peak(fma=0) on this GPU in this run: DMUL+DADD pairs = the only "
                    "mix a bit-exact non-FMA kernel may issue; peak_fma = the same with DFMA (MEASURED_PEAKS.json has no FP64 entry)")
        if fast:
            # closed-form tangent: flops the kernel executes per element = the 162 x 192 x 27 contraction (6 component pairs,
            # K^{mj} mirrored) + T (162 x 192 x 3) + per point F (27 x 9 FMA), stress + dP/dF (~ 600) + residual (81 x 192 FMA)
            kern = ("element_fast_p2v3_kernel<%s>" % args.material if args.fast_gemm == 0 else
                    "element_fast_p2_kernel<%s,%s>" % (args.material, "dmma" if args.fast_gemm == 2 else "dfma"))
            per_el = 2.0 * 162 * 192 * 27 + 2.0 * 162 * 192 * 3 + nq * (2.0 * 27 * 9 + 600.0) + 2.0 * 81 * 192
            exe = alg = per_el / nq
            fd = 0.0
            exe_qp = exe
            achieved = qp_rate_kernel * exe_qp / 1e12
            peak_dmma = ctx_peaks["dmma"]
            peak_nofma = peak_dmma if args.fast_gemm != 1 else peak_fma
            peak_src = ("fast mode (FMA allowed): mb2_measure_fp64_peak(%s) on this GPU in this run; DFMA %.2f, tensor-core "
                        "m8n8k4 %.2f TFLOP/s" % ("fma=2: FP64 tensor cores" if args.fast_gemm != 1 else "fma=1: DFMA", peak_fma, peak_dmma))
        traffic, traffic_note = ncu_traffic(kern, min(chunk_elems, NE))
        roofline = {"bound": "fp64", "kernel": kern, "achieved": achieved, "peak": peak_nofma,
                    "unit": "TFLOP/s", "frac": achieved / peak_nofma,
                    "peak_fma": peak_fma, "frac_vs_fma_peak": achieved / peak_fma,
                    "frac_algorithmic": qp_rate_kernel * (alg + fd / nq) / 1e12 / peak_nofma,
                    "traffic": traffic, "traffic_note": traffic_note,
                    "algorithmic_bytes_per_launch": bytes_el / NE * min(chunk_elems, NE), "launches_per_step": n_chunks,
                    "peak_source": peak_src,
                    "flop_per_qp_executed": exe_qp, "flop_per_qp_algorithmic": alg + fd / nq,
                    "achieved_algorithmic": qp_rate_kernel * (alg + fd / nq) / 1e12,
                    "kernel_ms": el_ms, "scatter_ms": float(np.mean(t_sc)),
                    "hbm_secondary": {"algorithmic_gb": bytes_el / 1e9, "achieved_gbs": bytes_el / 1e9 / (el_ms * 1e-3),
                                      "peak_gbs": hbm_peak, "peak_source": "MEASURED_PEAKS.json hbm_gbs (of measured)" if peaks else "fallback"}}
        cpu = None
        if not args.no_cpu:
            try:
                cores = host_cores()
                S = run_reference_sample(args.ref_n, 2, cores, warmup=1, p=p, material=args.material)
                cpu = {"value": S["value_mean"], "unit": UNIT, "cores": S["omp_team"], "kind": "reference",
                       "sample": "unmodified reference (oracle/_ref/mimi_ref, OpenMP team of %d threads, %d requested) on p=%d %s cube %d^3 = %d el, "
                                 "mean of 2 steps after 1 warm-up; residual-only best %.3e qp/s" % (
                                     S["omp_team"], cores, p, args.material, args.ref_n, S["NE"], S["qp"] / S["t_res_best"])}
                if S["omp_team"] < cores:
                    cpu["flag"] = "OpenMP team smaller than the host core count"
            except Exception as ex:  # the checker binary is absent: report that, never substitute
                cpu = {"value": None, "unit": UNIT, "cores": host_cores(), "kind": "reference", "sample": "unavailable: %s" % ex}
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
                "data": "synthetic",
                "config": {"workload": ("C4: multi-patch beam, %d patch(es) of %dx%dx%d knot spans glued along x, p=3, %d elements x %d qp per GPU "
                                        "(%d elements in all), stvk, one residual+tangent assembly into CSR (nnz=%d per GPU), lexicographic "
                                        "numbering over the glued control grid" % (M["n_patches"], M["spans"][0], M["spans"][1], M["spans"][2], NE, nq,
                                                                                    M["NE_global"], nnz)) if c4 else
                                       "%s: 3D unit-cube B-spline patch p=%d, %d^3 = %d elements x %d qp per GPU, %s, "
                                       "one residual+tangent assembly into CSR (nnz=%d per GPU), MFEM dof numbering" % (
                                           "C3" if (p == 2 and args.material == "neohookean") else "C4-style block", p, n, NE, nq, args.material, nnz),
                           "elements_per_gpu": NE, "qp_per_gpu": NE * nq, "nnz_per_gpu": nnz, "tangent": "FD parity (bit-exact vs reference)" if not fast else "MB2_TANGENT_ANALYTIC: closed-form tangent, %s contraction (NOT the headline mode: agrees with the reference's FD Jacobian to its 1e-4 noise, same Newton iteration counts)" % ("FP64 tensor-core (DMMA)" if args.fast_gemm != 1 else "CUDA-core DFMA"),
                           "l2": "inputs larger than L2 (tables %.1f GB, K_e %.1f GB per step)" % (8e-9 * NE * nq * (tdof + 1), 8e-9 * NE * tdof * tdof),
                           "parallelism": "1 GPU" if world == 1 else ("element ranges of %d elements per rank: every patch cut into 2 z-slabs, raw K_e/R_e of the p layers under every cut and of the first element column behind every patch interface sent to the owner rank by NCCL inside the step" % NE) if c4 else "z-slabs of %d^3 elements per GPU of one %dx%dx%d patch; raw K_e/R_e of the p interface layers sent to the owner rank by NCCL send/recv inside the step (bit-identical to the 1-thread reference order)" % (n, n, n, n * world),
                           "setup_s": M["setup_s"]},
                "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e, "e2e_host_csr": e2e_host, "e2e_roundtrip": e2e_rt,
                "strong_scaling": strong, "gpu_launches": launches, "clocks": clocks, "wall_s_timed_region": wall}
        if parity:
            line["parity_check"] = parity
        if res_only:
            res_only["hbm_peak_gbs"] = hbm_peak
            res_only["frac_hbm"] = res_only["achieved_gbs_per_gpu"] / hbm_peak
            line["residual_only"] = res_only
        if fast_mode:
            line["fast_mode"] = fast_mode
        emit(line)
    if world > 1:
        dist.destroy_process_group()
    return 0


# ----------------------------------------------------------------------------------------------------
# --workload c5: J2 elastoplasticity (history variables per qp) + rigid contact on the top face, one GPU.
# A "step" = what one Newton iteration of an implicit time step assembles: domain residual + FD tangent, then the
# contact residual + FD tangent on the same resident r / CSR values; DomainPostTimeAdvance is timed separately
# (once per converged step in the reference). Secondary workload: the headline bench line stays C3.
# ----------------------------------------------------------------------------------------------------
def c5_arm(args):
    import torch
    import torch.distributed as dist
    from mimi_b200 import capi, meshgen, partition
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py needs a CUDA device: libmimi_b200 has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def maxr(x):
        if world > 1:
            tt = torch.tensor([x], device="cuda", dtype=torch.float64)
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            return float(tt[0])
        return x

    parity = None
    if world > 1:
        parity = nccl_parity_check(rank, world, local, dist)
        if not parity["ok"]:
            raise RuntimeError("partitioned assembly over NCCL differs from the reference's golden vectors")
    n, p = args.n, 2
    stream = torch.cuda.Stream()
    ctx = capi.Context(local, stream=stream.cuda_stream)
    ctx.set_option("park_acc", args.park)
    t0 = time.time()
    nq1 = p + 2
    spans = (n, n, n * world)          # weak scaling: n^3 elements (and their history variables) per GPU, z-slabs
    if world == 1:
        M = meshgen.patch_connectivity(p, spans)
        M["n_compute"] = M["NE"]
        gtop = meshgen.patch_node_grid(M["ncp"])[:, :, -1]
    else:
        M = partition.slab_rank_problem(p, spans, world, rank)
        gtop = M["node_grid_top"]
    ctx.set_connectivity(3, M["nd"], nq1 ** 3, M["NV"], M["v_dofs"])
    if world > 1:
        uid = [capi.mg_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(uid, src=0)
        ctx.mg_setup(M)
        ctx.mg_init_nccl(rank, world, uid[0])
    ctx.set_nodes(M["x_ref"], None)
    ctx.generate_patch_tables(p, M["ncp"], M["knots"], M["el_ijk"][:M["n_compute"]])
    nnz = ctx.build_pattern()
    young, poisson = 2100.0, 0.3
    lam = young * poisson / ((1 + poisson) * (1 - 2 * poisson))
    mu = young / (2.0 * (1.0 + poisson))
    K = young / (3.0 * (1.0 - (2.0 * poisson)))
    # J2 + JohnsonCookTemperatureAndRateDependentHardening (tests/test_nonlinear_solid.py:16-33), yield stress scaled so
    # that the synthetic 1e-3 displacement field yields at about half of the quadrature points
    hp = [2.0, 4.0, 0.2835, 0.015, 1.0e-5, 20.0, 0.0, 1.3558, 0.0, 0.0]
    ctx.set_material(capi.MB2_MAT_J2, [lam, mu, K, mu, 4] + hp + [0.05, 0.9, 1.0, 450.0, 20.0, 1500.0])
    # rigid contact on the global top face: every rank evaluates all marked faces (faces << elements), gathers its own rows
    F = meshgen.top_face_tables_from_grid(p, spans[:2], M["knots"], gtop)
    ctx.contact_set_faces(F["v_dofs"], F["N"], F["dN_dxi"], F["w"])
    ztop = float(M["x_ref"].reshape(-1, 3)[:, 2].max())
    body = np.zeros(16)
    body[:4] = [0.5, 0.5, ztop + 1.0 - 2e-4, 1.0]     # rigid sphere pressed 2e-4 into the top face
    ctx.contact_set_rigid_body(capi.MB2_RIGID_SPHERE, body, 1.0e4 * young)
    u = meshgen.synth_u(M["NV"])
    x = M["x_ref"].reshape(-1, 3)
    u.reshape(-1, 3)[x[:, 2] == 0.0, :] = 0.0       # clamped bottom
    setup_s = time.time() - t0
    NE, nq = M["n_compute"], nq1 ** 3
    tdof = 3 * M["nd"]
    ctx.set_u(u); ctx.set_r(None); ctx.set_A(None)
    # commit a first history so that the timed assemblies run on a non-virgin state
    ctx._ck(ctx.lib.mb2_post_time_advance_device(ctx.h))
    ctx.set_u(np.ascontiguousarray(1.5 * u))
    ctx.sync()

    def step():
        ctx.assemble_residual_and_grad_device(1.0)
        el, sc, tot = ctx.last_timing_ms()
        l0 = ctx.last_launch_count()
        ctx.contact_residual_and_grad_device(1.0)
        return el, sc, l0 + ctx.last_launch_count()

    for _ in range(args.warmup):
        step()
    ctx.sync()
    sampler = ClockSampler(local)
    barrier()
    sampler.start()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2 * args.steps + 2)]
    t_el, launches = [], 0
    ev[0].record(stream)
    for k in range(args.steps):
        ctx.assemble_residual_and_grad_device(1.0)
        el, sc, tot = ctx.last_timing_ms()
        t_el.append(el)
        launches += ctx.last_launch_count()
        ev[1 + 2 * k].record(stream)
        ctx.contact_residual_and_grad_device(1.0)
        launches += ctx.last_launch_count()
        ev[2 + 2 * k].record(stream)
    barrier()
    clocks = sampler.stop()
    dev_ms = maxr(float(ev[0].elapsed_time(ev[2 * args.steps])))
    contact_ms = float(np.mean([ev[1 + 2 * k].elapsed_time(ev[2 + 2 * k]) for k in range(args.steps)]))
    ctx.check_numeric()
    # ---- what the return mapping really did in one such assembly (measurement pass, atomics on, outside the timing)
    ctx.set_option("j2_counters", 1)
    ctx.j2_counters(reset=True)
    ctx.assemble_residual_and_grad_device(1.0)
    n_yield, n_hard = ctx.j2_counters(reset=True)
    ctx.set_option("j2_counters", 0)
    n_eval = float(NE) * nq * (tdof + 1)
    yielded_fraction = n_yield / n_eval
    hard_per_yield = n_hard / max(n_yield, 1)
    face_qp = F["NF"] * F["nqf"]
    ms_per_step = dev_ms / args.steps
    total_qp = float(NE) * nq * world + face_qp
    # end to end: host-buffer entries (domain + contact) with pinned buffers, on the same displacement and history as the timed steps
    # (before the time-step loop below moves the history on)
    e2e = None
    if not args.no_e2e:
        u_h = torch.from_numpy(np.ascontiguousarray(1.5 * u)).pin_memory()
        r_h = torch.zeros(M["NV"], dtype=torch.float64).pin_memory()
        A_h = torch.zeros(nnz, dtype=torch.float64).pin_memory()
        up, rp, Ap = (C_void(t) for t in (u_h, r_h, A_h))
        one = capi.C.c_double(1.0)

        def host_step():
            ctx._ck(ctx.lib.mb2_add_domain_residual_and_grad(ctx.h, up, one, rp, Ap))
            ctx._ck(ctx.lib.mb2_add_boundary_residual_and_grad(ctx.h, up, one, rp, Ap))

        host_step()
        barrier()
        t1 = time.perf_counter()
        k2 = max(1, min(args.steps, 3))
        for _ in range(k2):
            host_step()
        barrier()
        e2e_s = maxr((time.perf_counter() - t1) / k2)
        e2e = {"value": total_qp / e2e_s, "unit": UNIT, "h2d_bytes_per_step": 8 * (4 * M["NV"] + nnz),
               "d2h_bytes_per_step": 8 * (2 * M["NV"] + nnz), "ms_per_step": e2e_s * 1e3, "steps": k2,
               "call": "mb2_add_domain_residual_and_grad + mb2_add_boundary_residual_and_grad with pinned host buffers"}
    # ---- implicit-dynamics time steps: 2 Newton iterations (assembly + contact, two line-search residuals each) and the
    # history commit per step, the displacement growing from step to step (assembly share of a generalized-alpha step)
    n_ts = max(2, min(args.steps, 4))
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t_hist = 0.0
    e0.record(stream)
    for ts in range(n_ts):
        ctx.set_u(np.ascontiguousarray((1.5 + 0.1 * (ts + 1)) * u))
        for it in range(2):
            ctx.assemble_residual_and_grad_device(1.0)
            ctx.contact_residual_and_grad_device(1.0)
            for ls in range(2):
                ctx.assemble_residual_device()
                ctx._ck(ctx.lib.mb2_contact_residual_device(ctx.h))
        ctx._ck(ctx.lib.mb2_post_time_advance_device(ctx.h))
    e1.record(stream)
    barrier()
    ts_ms = maxr(float(e0.elapsed_time(e1)) / n_ts)
    ctx.check_numeric()
    ea, eb = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ea.record(stream)
    ctx._ck(ctx.lib.mb2_post_time_advance_device(ctx.h))
    eb.record(stream)
    torch.cuda.synchronize()
    pta_ms = float(ea.elapsed_time(eb))
    gap, area, pr = ctx.contact_nodal()
    value = total_qp / (ms_per_step * 1e-3)
    peak_nofma = ctx.measure_fp64_peak(0)
    peak_fma = ctx.measure_fp64_peak(1)
    el_ms = float(np.mean(t_el))
    # ---- executed flops of the J2 element kernel. The return mapping runs a data-dependent number of hardening-law evaluations,
    # so the count is a calibrated linear model (profiles/j2_flops.json, written by scripts/ncu_j2_flops.py from an ncu capture:
    # FP64 thread instructions DADD + DMUL + 2 DFMA, split by source line into a fixed part per qp and a part per hardening
    # evaluation) evaluated at THIS run's device counters (return mappings, hardening evaluations of one assembly)
    roof = {"bound": "fp64", "kernel": "element_fd_kernel<3,27,64,J2>", "achieved": None, "peak": None, "unit": "TFLOP/s", "frac": None,
            "traffic": None, "kernel_ms": el_ms, "yielded_fraction": yielded_fraction, "hardening_evals_per_return_mapping": hard_per_yield,
            "stress_evaluations_per_assembly": n_eval, "park_acc": args.park}
    try:
        Jf = json.load(open(os.path.join(ROOT, "profiles", "j2_flops.json")))
        evals_qp = n_eval * yielded_fraction * hard_per_yield / (float(NE) * nq)
        f_fix, f_var = Jf["flop_fixed_per_qp"], Jf["flop_per_hardening_eval"] * evals_qp
        flop_qp = f_fix + f_var
        # share of the flops issued as DFMA (libdevice pow / log / divisions are fused whatever -fmad says)
        mix = (f_fix * Jf["fused_share_fixed"] + f_var * Jf["fused_share_variable"]) / flop_qp
        peak_mix = 1.0 / ((1.0 - mix) / peak_nofma + mix / peak_fma)
        achieved = float(NE) * nq / (el_ms * 1e-3) * flop_qp / 1e12
        roof.update(achieved=achieved, peak=peak_mix, frac=achieved / peak_mix,
                    flop_per_qp_executed=flop_qp, hardening_evals_per_qp=evals_qp, fused_flop_share=mix,
                    peak_nofma=peak_nofma, peak_fma=peak_fma,
                    peak_source="harmonic mix of the DMUL+DADD and DFMA peaks measured in this run (mb2_measure_fp64_peak) at the "
                                "kernel's fused flop share %.3f" % mix,
                    flop_source="profiles/j2_flops.json: %.0f flop/qp fixed + %.1f flop per hardening evaluation x %.1f evaluations/qp "
                                "counted by this run; calibrated on %s" % (f_fix, Jf["flop_per_hardening_eval"], evals_qp, Jf.get("capture")))
    except (OSError, KeyError, ValueError):
        roof["note"] = "no profiles/j2_flops.json: data-dependent flop count not calibrated"
    if rank == 0:
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64",
                "data": "synthetic",
                "config": {"workload": "C5-style (secondary workload, not the headline): 3D B-spline block p=2, %d x %d x %d = %d "
                                       "elements x %d qp per GPU (%d GPUs: %d elements), J2 + JohnsonCookTemperatureAndRateDependent hardening with %d history "
                                       "doubles per qp (partitioned with the elements), rigid sphere on the %d faces (x %d qp) of the top face (evaluated by every rank); one step = domain "
                                       "residual+tangent + contact residual+tangent into the same CSR (nnz=%d per GPU)" % (
                                           n, n, n, NE, nq, world, NE * world, ctx.state_size(), F["NF"], F["nqf"], nnz),
                           "elements_per_gpu": NE, "qp_per_gpu": NE * nq, "face_qp": face_qp, "nnz_per_gpu": nnz,
                           "active_contact_nodes": int((pr != 0).sum()), "marked_nodes": int(pr.size),
                           "l2": "inputs larger than L2", "parallelism": "1 GPU" if world == 1 else "z-slabs, NCCL exchange of raw K_e / R_e; history local; contact replicated",
                           "setup_s": setup_s},
                "roofline": roof,
                "extra": {"domain_element_kernel_ms": el_ms, "contact_ms": contact_ms, "post_time_advance_ms": pta_ms,
                          "contact_share": contact_ms / ms_per_step,
                          "time_step": {"steps": n_ts, "newton_iterations_per_step": 2, "line_search_residuals_per_iteration": 2,
                                        "ms_per_time_step": ts_ms, "note": "assembly share of an implicit generalized-alpha step: 2 x (residual + tangent + contact), "
                                        "4 x (residual + contact), 1 history commit; the displacement grows every step"}},
                "cpu_baseline": None, "e2e": e2e, "gpu_launches": launches, "clocks": clocks}
        if parity:
            line["parity_check"] = parity
        if not args.no_cpu:
            try:
                cores = host_cores()
                S = run_reference_sample(min(args.ref_n, 24), 2, cores, warmup=1, p=2, material="j2_jcthermo")
                line["cpu_baseline"] = {"value": S["value_mean"], "unit": UNIT, "cores": S["omp_team"], "kind": "reference",
                                        "sample": "unmodified reference, domain integrator only (J2 + JC thermo-rate hardening), "
                                                  "p=2 cube %d^3 = %d el, virgin state, %d threads" % (min(args.ref_n, 24), S["NE"], S["omp_team"])}
            except Exception as ex:
                line["cpu_baseline"] = {"value": None, "unit": UNIT, "cores": host_cores(), "kind": "reference",
                                        "sample": "unavailable: %s" % ex}
        emit(line)
    ctx.close()
    if world > 1:
        dist.destroy_process_group()
    return 0


def C_void(t):
    import ctypes
    return ctypes.c_void_p(t.data_ptr())


_JSON_OUT = None


def emit(line):
    out = _JSON_OUT or sys.stdout
    out.write(json.dumps(line) + "\n")
    out.flush()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--n", "--size", dest="n", type=int, default=100, help="knot spans per direction (per GPU)")
    ap.add_argument("--ref-n", type=int, default=32, help="spans per direction of the CPU reference sample")
    ap.add_argument("--e2e-steps", type=int, default=0, help="timed e2e steps (0 = --steps)")
    ap.add_argument("--scaling", default="both", choices=["weak", "strong", "both"],
                    help="N > 1: weak = n^3 elements per GPU (the headline value); strong = the one n^3 block cut into N slabs (key strong_scaling)")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-res-only", action="store_true", help="skip the residual-only pass")
    ap.add_argument("--no-fast", action="store_true", help="skip the secondary fast-tangent-mode measurement (key fast_mode)")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--p", type=int, default=2, help="degree (2 = C3 headline; 3 = the C4 element type)")
    ap.add_argument("--material", default="neohookean", choices=["neohookean", "stvk"])
    ap.add_argument("--workload", default="c3", choices=["c3", "c4", "c5"],
                    help="c4: multi-patch p=3 StVK beam, 5e5 elements per GPU (config 4); c5: J2 + contact secondary workload (1 GPU)")
    ap.add_argument("--park", type=int, default=1, help="A/B: FD kernels park their accumulators in tensor memory during phase 1 (mb2_set_option park_acc; p=2: plastic materials, p=3: all)")
    ap.add_argument("--mg-late", type=int, default=-1, help="A/B (c4, N > 1): mb2_set_option mg_exchange_late (-1 = 1 when patches are split over ranks in x)")
    ap.add_argument("--res-variant", type=int, default=0, help="A/B: residual-only kernel (0 = element_res2_kernel, 1 = first form)")
    ap.add_argument("--variant", type=int, default=0, help="A/B: element kernel variant (mb2_set_option elem_variant)")
    ap.add_argument("--tangent", default="parity", choices=["parity", "analytic"],
                    help="analytic = MB2_TANGENT_ANALYTIC fast mode (secondary line; p=2, neohookean / stvk)")
    ap.add_argument("--fast-gemm", type=int, default=0, help="analytic mode: 0 = FP64 tensor cores, A fragments in registers (default), 1 = CUDA-core DFMA, 2 = tensor cores fed from T buffers")
    args = ap.parse_args()
    global MG_LATE
    MG_LATE = args.mg_late
    # The contract is ONE JSON line on stdout: libraries that write to fd 1 (NCCL prints its version banner there)
    # are sent to stderr; the JSON line goes to the original stdout.
    global _JSON_OUT
    sys.stdout.flush()
    _JSON_OUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    if args.warmup < 3 and args.impl == "b200":
        args.warmup = max(args.warmup, 3) if not os.environ.get("MB2_BENCH_ALLOW_SHORT") else args.warmup
    if args.impl == "reference":
        return reference_arm(args)
    if args.workload == "c5":
        return c5_arm(args)
    return gpu_arm(args)


if __name__ == "__main__":
    sys.exit(main())
