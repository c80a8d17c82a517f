"""ctypes binding of the C ABI in include/mimi_b200.h (the same symbols a cgo/pybind caller binds).

There is no CPU fallback: importing works anywhere, but every call needs libmimi_b200.so (built by
``__graft_entry__.build()`` / ``make -C mimi_b200/csrc``) and, for compute, an sm_100a GPU.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libmimi_b200.so")

MB2_MAT_NEOHOOKEAN, MB2_MAT_STVK, MB2_MAT_J2LINEAR, MB2_MAT_J2, MB2_MAT_J2SIMO, MB2_MAT_J2LOG = 0, 1, 2, 3, 4, 5
MB2_RIGID_PLANE, MB2_RIGID_SPHERE, MB2_RIGID_BEZIER, MB2_RIGID_BSPLINE = 0, 1, 2, 3
MB2_RIGID_RATIONAL_BEZIER, MB2_RIGID_NURBS = 4, 5

# every symbol include/mimi_b200.h declares (checked by tests/test_abi.py)
SYMBOLS = [
    "mb2_create", "mb2_destroy", "mb2_last_error", "mb2_version", "mb2_set_connectivity", "mb2_dot_noalloc", "mb2_newton_solve_direct_device",
    "mb2_upload_quad_tables", "mb2_set_nodes", "mb2_generate_patch_tables", "mb2_assemble_mass_device", "mb2_assemble_body_force_device", "mb2_assemble_viscosity_device", "mb2_assemble_traction_device", "mb2_newton_viscosity_from_A", "mb2_newton_get_viscosity", "mb2_newton_rhs_from_r", "mb2_newton_mass_from_A", "mb2_newton_get_mass", "mb2_download_shape_values", "mb2_download_quad_tables", "mb2_build_pattern",
    "mb2_get_pattern_i32", "mb2_get_pattern_i64", "mb2_get_A_ids_i32", "mb2_set_material",
    "mb2_add_domain_residual", "mb2_add_domain_residual_and_grad", "mb2_set_u", "mb2_set_r", "mb2_set_A",
    "mb2_get_r", "mb2_get_A", "mb2_assemble_residual_device", "mb2_assemble_residual_and_grad_device",
    "mb2_device_pointers", "mb2_last_timing_ms", "mb2_last_launch_count", "mb2_sync", "mb2_measure_fp64_peak", "mb2_set_option", "mb2_host_register", "mb2_host_unregister", "mb2_plan_info", "mb2_set_partition", "mb2_mg_set_send", "mb2_mg_set_recv", "mb2_mg_unique_id",
    "mb2_mg_init_nccl", "mb2_mg_set_manual", "mb2_mg_begin", "mb2_mg_export_send", "mb2_mg_import_recv", "mb2_mg_finish",
    "mb2_mg_ke_doubles", "mb2_state_size", "mb2_get_state", "mb2_set_state", "mb2_post_time_advance",
    "mb2_post_time_advance_device", "mb2_set_time_step", "mb2_check_numeric", "mb2_contact_set_faces", "mb2_contact_set_rigid_body", "mb2_contact_n_marked_nodes",
    "mb2_contact_get_marked_v_dofs", "mb2_contact_get_A_ids_i32", "mb2_add_boundary_residual",
    "mb2_add_boundary_residual_and_grad", "mb2_contact_residual_device", "mb2_contact_residual_and_grad_device",
    "mb2_contact_get_nodal", "mb2_contact_last_scalars", "mb2_contact_gap_norm", "mb2_set_mass", "mb2_set_viscosity",
    "mb2_set_rhs", "mb2_set_essential_dofs", "mb2_newton_residual_and_grad_device", "mb2_newton_residual_device", "mb2_newton_solve_gmres_device", "mb2_newton_get_solution",
    "mb2_domain_residual_and_grad_on_mass", "mb2_get_j2_counters", "mb2_plan_partition", "mb2_plan_free", "mb2_plan_size",
    "mb2_plan_local_elems", "mb2_plan_owned_nodes", "mb2_plan_peer", "mb2_plan_peer_send", "mb2_plan_apply", "mb2_mg_group_init",
    "mb2_mg_group_assemble",
]


class Mb2Error(RuntimeError):
    """Mirrors mimi::utils::PrintAndThrowError -> std::runtime_error -> Python RuntimeError
    (src/mimi/utils/print.hpp:49-56)."""


_lib = None


def mg_unique_id():
    buf = C.create_string_buffer(128)
    rc = load().mb2_mg_unique_id(buf)
    if rc:
        raise Mb2Error("mb2_mg_unique_id failed (libnccl not loadable?)")
    return buf.raw


def load():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise Mb2Error("libmimi_b200.so is not built (run __graft_entry__.build()); there is no CPU fallback")
        _lib = C.CDLL(LIB_PATH)
        _lib.mb2_last_error.restype = C.c_char_p
        _lib.mb2_last_error.argtypes = [C.c_void_p]
    return _lib


def _dp(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


class Context:
    """Thin owner of an mb2_ctx*."""

    def __init__(self, device=0, stream=None):
        self.lib = load()
        self.h = C.c_void_p()
        rc = self.lib.mb2_create(C.byref(self.h), C.c_int(device), C.c_void_p(stream or 0))
        if rc:
            raise Mb2Error(self.lib.mb2_last_error(None).decode())

    def _ck(self, rc):
        if rc:
            raise Mb2Error("mb2 error %d: %s" % (rc, self.lib.mb2_last_error(self.h).decode()))

    def close(self):
        if self.h:
            self.lib.mb2_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- setup
    def set_connectivity(self, dim, nd, nq, NV, v_dofs):
        v = np.ascontiguousarray(v_dofs, np.int32)
        self.dim, self.nd, self.nq, self.tdof, self.NE, self.NV = dim, nd, nq, dim * nd, v.shape[0], NV
        self._ck(self.lib.mb2_set_connectivity(self.h, dim, nd, nq, C.c_int64(self.NE), C.c_int64(NV), _dp(v)))

    def upload_quad_tables(self, dN_dX, w, detJ):
        a, b, c = (np.ascontiguousarray(x, np.float64) for x in (dN_dX, w, detJ))
        assert a.size == self.NE * self.nq * self.tdof and b.size == self.NE * self.nq == c.size
        self._ck(self.lib.mb2_upload_quad_tables(self.h, _dp(a), _dp(b), _dp(c)))

    def set_nodes(self, x_ref, weights=None):
        x = np.ascontiguousarray(x_ref, np.float64)
        w = None if weights is None else np.ascontiguousarray(weights, np.float64)
        assert x.size == self.NV
        self._ck(self.lib.mb2_set_nodes(self.h, _dp(x), _dp(w)))

    def generate_patch_tables(self, p, ncp, knots, el_ijk, e0=0, quad_order=-1, want_N=False):
        ncp = np.ascontiguousarray(ncp, np.int32)
        ijk = np.ascontiguousarray(el_ijk, np.int32)
        kv = [np.ascontiguousarray(k, np.float64) for k in knots]
        assert all(len(k) == n + p + 1 for k, n in zip(kv, ncp))
        arr = (C.c_void_p * len(kv))(*[k.ctypes.data for k in kv])
        self._ck(self.lib.mb2_generate_patch_tables(self.h, C.c_int64(e0), C.c_int64(ijk.shape[0]), C.c_int(p), _dp(ncp),
                                                    arr, _dp(ijk), C.c_int(quad_order), C.c_int(int(want_N))))

    def assemble_mass_device(self, p, ncp, knots, el_ijk, quad_order=-1, density=1.0):
        """A += M (VectorMassIntegrator) on the resident CSR values."""
        ncp = np.ascontiguousarray(ncp, np.int32)
        ijk = np.ascontiguousarray(el_ijk, np.int32)
        kv = [np.ascontiguousarray(k, np.float64) for k in knots]
        assert all(len(k) == n + p + 1 for k, n in zip(kv, ncp))
        arr = (C.c_void_p * len(kv))(*[k.ctypes.data for k in kv])
        self._ck(self.lib.mb2_assemble_mass_device(self.h, C.c_int(p), _dp(ncp), arr, _dp(ijk), C.c_int(quad_order),
                                                   C.c_double(density)))

    def assemble_body_force_device(self, p, ncp, knots, el_ijk, force, quad_order=-1):
        """r += b (VectorDomainLFIntegrator of a constant body force) on the resident vector."""
        ncp = np.ascontiguousarray(ncp, np.int32)
        ijk = np.ascontiguousarray(el_ijk, np.int32)
        kv = [np.ascontiguousarray(k, np.float64) for k in knots]
        f = np.ascontiguousarray(force, np.float64)
        assert f.size == self.dim and all(len(k) == n + p + 1 for k, n in zip(kv, ncp))
        arr = (C.c_void_p * len(kv))(*[k.ctypes.data for k in kv])
        self._ck(self.lib.mb2_assemble_body_force_device(self.h, C.c_int(p), _dp(ncp), arr, _dp(ijk), C.c_int(quad_order),
                                                         _dp(f)))

    def assemble_traction_device(self, face_vdofs, N, w, detW, traction):
        """r = ((r + f_1) + f_2) + ... (VectorBoundaryLFIntegrator of a constant traction on the uploaded faces)."""
        vd = np.ascontiguousarray(face_vdofs, np.int32)
        N = np.ascontiguousarray(N, np.float64)
        w = np.ascontiguousarray(w, np.float64)
        dw = np.ascontiguousarray(detW, np.float64)
        t = np.ascontiguousarray(traction, np.float64)
        NF, nqf, ndf = N.shape
        assert vd.shape == (NF, self.dim * ndf) and w.shape == (NF, nqf) and dw.shape == (NF, nqf) and t.size == self.dim
        self._ck(self.lib.mb2_assemble_traction_device(self.h, C.c_int64(NF), C.c_int(ndf), C.c_int(nqf), _dp(vd), _dp(N),
                                                       _dp(w), _dp(dw), _dp(t)))

    def newton_rhs_from_r(self, zero_essential=True):
        self._ck(self.lib.mb2_newton_rhs_from_r(self.h, C.c_int(int(zero_essential))))

    def assemble_viscosity_device(self, p, ncp, knots, el_ijk, viscosity, quad_order=-1):
        """A += S (VectorDiffusionIntegrator) on the resident CSR values."""
        ncp = np.ascontiguousarray(ncp, np.int32)
        ijk = np.ascontiguousarray(el_ijk, np.int32)
        kv = [np.ascontiguousarray(k, np.float64) for k in knots]
        assert all(len(k) == n + p + 1 for k, n in zip(kv, ncp))
        arr = (C.c_void_p * len(kv))(*[k.ctypes.data for k in kv])
        self._ck(self.lib.mb2_assemble_viscosity_device(self.h, C.c_int(p), _dp(ncp), arr, _dp(ijk), C.c_int(quad_order),
                                                        C.c_double(viscosity)))

    def newton_viscosity_from_A(self, form_system_matrix=True):
        self._ck(self.lib.mb2_newton_viscosity_from_A(self.h, C.c_int(int(form_system_matrix))))

    def newton_get_viscosity(self):
        S = np.empty(self.nnz)
        self._ck(self.lib.mb2_newton_get_viscosity(self.h, _dp(S)))
        return S

    def newton_mass_from_A(self, form_system_matrix=True):
        self._ck(self.lib.mb2_newton_mass_from_A(self.h, C.c_int(int(form_system_matrix))))

    def newton_get_mass(self):
        M = np.empty(self.nnz)
        self._ck(self.lib.mb2_newton_get_mass(self.h, _dp(M)))
        return M

    def download_shape_values(self):
        N = np.empty((self.NE, self.nq, self.nd))
        self._ck(self.lib.mb2_download_shape_values(self.h, _dp(N)))
        return N

    def download_quad_tables(self):
        dN = np.empty((self.NE, self.nq, self.dim, self.nd))
        wd = np.empty((self.NE, self.nq))
        self._ck(self.lib.mb2_download_quad_tables(self.h, _dp(dN), _dp(wd)))
        return dN, wd

    def build_pattern(self):
        nnz = C.c_int64()
        self._ck(self.lib.mb2_build_pattern(self.h, C.byref(nnz)))
        self.nnz = nnz.value
        return self.nnz

    def get_pattern(self, i64=False):
        I = np.empty(self.NV + 1, np.int64 if i64 else np.int32)
        J = np.empty(self.nnz, np.int32)
        f = self.lib.mb2_get_pattern_i64 if i64 else self.lib.mb2_get_pattern_i32
        self._ck(f(self.h, _dp(I), _dp(J)))
        return I, J

    def get_A_ids(self):
        A = np.empty((self.NE, self.tdof * self.tdof), np.int32)
        self._ck(self.lib.mb2_get_A_ids_i32(self.h, _dp(A)))
        return A

    def set_material(self, kind, params):
        p = np.ascontiguousarray(params, np.float64)
        self._ck(self.lib.mb2_set_material(self.h, kind, _dp(p), p.size))

    # ---- history variables (materials with state)
    def state_size(self):
        return self.lib.mb2_state_size(self.h)

    def get_state(self):
        st = np.empty((self.NE, self.nq, self.state_size()))
        self._ck(self.lib.mb2_get_state(self.h, _dp(st)))
        return st

    def set_state(self, state=None):
        st = None if state is None else np.ascontiguousarray(state, np.float64)
        self._ck(self.lib.mb2_set_state(self.h, _dp(st)))

    def set_time_step(self, dt):
        self._ck(self.lib.mb2_set_time_step(self.h, C.c_double(dt)))

    def check_numeric(self):
        self._ck(self.lib.mb2_check_numeric(self.h))

    def post_time_advance(self, u):
        self._ck(self.lib.mb2_post_time_advance(self.h, _dp(u)))

    # ---- rigid-body mortar contact
    def contact_set_faces(self, v_dofs, N, dN_dxi, w):
        v = np.ascontiguousarray(v_dofs, np.int32)
        N, dN, w = (np.ascontiguousarray(x, np.float64) for x in (N, dN_dxi, w))
        self.NF, self.nqf = w.shape
        self.ndf = v.shape[1] // self.dim
        assert N.size == self.NF * self.nqf * self.ndf and dN.size == N.size * (self.dim - 1)
        self._ck(self.lib.mb2_contact_set_faces(self.h, C.c_int64(self.NF), self.ndf, self.nqf, _dp(v), _dp(N), _dp(dN), _dp(w)))
        n = C.c_int64()
        self._ck(self.lib.mb2_contact_n_marked_nodes(self.h, C.byref(n)))
        self.n_marked = n.value

    def contact_set_rigid_body(self, kind, params, penalty):
        p = np.ascontiguousarray(params, np.float64)
        self._ck(self.lib.mb2_contact_set_rigid_body(self.h, kind, _dp(p), p.size, C.c_double(penalty)))

    def contact_marked_v_dofs(self):
        out = np.empty(self.n_marked * self.dim, np.int32)
        self._ck(self.lib.mb2_contact_get_marked_v_dofs(self.h, _dp(out)))
        return out

    def contact_A_ids(self):
        tdf = self.ndf * self.dim
        out = np.empty((self.NF, tdf * tdf), np.int32)
        self._ck(self.lib.mb2_contact_get_A_ids_i32(self.h, _dp(out)))
        return out

    def add_boundary_residual(self, u, r):
        self._ck(self.lib.mb2_add_boundary_residual(self.h, _dp(u), _dp(r)))

    def add_boundary_residual_and_grad(self, u, gf, r, A):
        self._ck(self.lib.mb2_add_boundary_residual_and_grad(self.h, _dp(u), C.c_double(gf), _dp(r), _dp(A)))

    def contact_residual_device(self):
        self._ck(self.lib.mb2_contact_residual_device(self.h))

    def contact_residual_and_grad_device(self, gf):
        self._ck(self.lib.mb2_contact_residual_and_grad_device(self.h, C.c_double(gf)))

    def contact_nodal(self):
        g, a, p = (np.empty(self.n_marked) for _ in range(3))
        self._ck(self.lib.mb2_contact_get_nodal(self.h, _dp(g), _dp(a), _dp(p)))
        return g, a, p

    def contact_last_scalars(self):
        out = np.empty(5)
        self._ck(self.lib.mb2_contact_last_scalars(self.h, _dp(out)))
        return out

    def contact_gap_norm(self, u):
        out = C.c_double()
        self._ck(self.lib.mb2_contact_gap_norm(self.h, _dp(u), C.byref(out)))
        return out.value

    # ---- Newton-iteration sweeps on the resident CSR
    def set_mass(self, M):
        self._ck(self.lib.mb2_set_mass(self.h, _dp(np.ascontiguousarray(M, np.float64))))

    def set_viscosity(self, S=None):
        self._ck(self.lib.mb2_set_viscosity(self.h, _dp(None if S is None else np.ascontiguousarray(S, np.float64))))

    def set_rhs(self, rhs=None):
        self._ck(self.lib.mb2_set_rhs(self.h, _dp(None if rhs is None else np.ascontiguousarray(rhs, np.float64))))

    def set_essential_dofs(self, tdofs):
        t = np.ascontiguousarray(tdofs, np.int32)
        self._ck(self.lib.mb2_set_essential_dofs(self.h, _dp(t), C.c_int64(t.size)))

    def newton_residual_and_grad_device(self, x, v, a, fac0, fac1):
        self._ck(self.lib.mb2_newton_residual_and_grad_device(self.h, _dp(x), _dp(v), _dp(a), C.c_double(fac0),
                                                              C.c_double(fac1)))

    # ---- host-buffer hot path
    def newton_residual_device(self, x, v, a, fac0, fac1, want_norm=True):
        nrm = C.c_double()
        self._ck(self.lib.mb2_newton_residual_device(self.h, _dp(x), _dp(v), _dp(a), C.c_double(fac0), C.c_double(fac1),
                                                     C.byref(nrm) if want_norm else None))
        return nrm.value

    def newton_solve_direct_device(self):
        """J c = y on the resident arrays by cuSOLVER's sparse QR (the reference's default is UMFPACK); -> singular row (-1)."""
        row = C.c_int(-1)
        self._ck(self.lib.mb2_newton_solve_direct_device(self.h, C.byref(row)))
        return row.value

    def dot_noalloc(self, d_a, d_b, n, d_out):
        """mb2_dot_noalloc: device addresses (ints) of a, b (n doubles) and of the one-double result; enqueued on the
        context's stream, allocates nothing."""
        self._ck(self.lib.mb2_dot_noalloc(self.h, C.c_void_p(d_a), C.c_void_p(d_b), C.c_int64(n), C.c_void_p(d_out)))

    def newton_solve_gmres_device(self, rel_tol=1e-8, abs_tol=1e-12, max_iter=300, kdim=50):
        it, conv, fn = C.c_int(), C.c_int(), C.c_double()
        self._ck(self.lib.mb2_newton_solve_gmres_device(self.h, C.c_double(rel_tol), C.c_double(abs_tol), max_iter, kdim,
                                                        C.byref(it), C.byref(fn), C.byref(conv)))
        return it.value, fn.value, bool(conv.value)

    def newton_get_solution(self):
        x = np.empty(self.NV)
        self._ck(self.lib.mb2_newton_get_solution(self.h, _dp(x)))
        return x

    def add_domain_residual(self, u, r):
        self._ck(self.lib.mb2_add_domain_residual(self.h, _dp(u), _dp(r)))

    def add_domain_residual_and_grad(self, u, gf, r, A):
        self._ck(self.lib.mb2_add_domain_residual_and_grad(self.h, _dp(u), C.c_double(gf), _dp(r), _dp(A)))

    def j2_counters(self, reset=True):
        out = (C.c_uint64 * 2)()
        self._ck(self.lib.mb2_get_j2_counters(self.h, out, int(reset)))
        return int(out[0]), int(out[1])

    def domain_residual_and_grad_on_mass(self, u, gf, r, A_out):
        """r += R(u); A_out = M + gf K(u) with the mass resident on the device (mb2_set_mass): no CSR upload."""
        self._ck(self.lib.mb2_domain_residual_and_grad_on_mass(self.h, _dp(u), C.c_double(gf), _dp(r), _dp(A_out)))

    # ---- device-resident hot path
    def set_u(self, u):
        self._ck(self.lib.mb2_set_u(self.h, _dp(u)))

    def set_r(self, r=None):
        self._ck(self.lib.mb2_set_r(self.h, _dp(r)))

    def set_A(self, A=None):
        self._ck(self.lib.mb2_set_A(self.h, _dp(A)))

    def get_r(self):
        r = np.empty(self.NV)
        self._ck(self.lib.mb2_get_r(self.h, _dp(r)))
        return r

    def get_A(self):
        A = np.empty(self.nnz)
        self._ck(self.lib.mb2_get_A(self.h, _dp(A)))
        return A

    def assemble_residual_device(self):
        self._ck(self.lib.mb2_assemble_residual_device(self.h))

    def assemble_residual_and_grad_device(self, gf):
        self._ck(self.lib.mb2_assemble_residual_and_grad_device(self.h, C.c_double(gf)))

    def last_timing_ms(self):
        t = (C.c_float * 3)()
        self._ck(self.lib.mb2_last_timing_ms(self.h, t))
        return list(t)

    def last_launch_count(self):
        return self.lib.mb2_last_launch_count(self.h)

    def set_option(self, name, value):
        self._ck(self.lib.mb2_set_option(self.h, name.encode(), C.c_int(value)))

    def plan_info(self):
        ec, n = C.c_int64(), C.c_int()
        self._ck(self.lib.mb2_plan_info(self.h, C.byref(ec), C.byref(n)))
        return ec.value, n.value

    def host_register(self, arr):
        self._ck(self.lib.mb2_host_register(self.h, _dp(arr), C.c_size_t(arr.nbytes)))

    def host_unregister(self, arr):
        self._ck(self.lib.mb2_host_unregister(self.h, _dp(arr)))

    # ---- multi-GPU partition
    def set_partition(self, n_compute, owned_nodes):
        m = np.ascontiguousarray(owned_nodes, np.uint8)
        assert m.size == self.NV // self.dim
        self._ck(self.lib.mb2_set_partition(self.h, C.c_int64(n_compute), _dp(m)))
        self.n_compute = n_compute

    def mg_setup(self, R, manual=False):
        """R: a dict from mimi_b200.partition.rank_problem."""
        self.set_partition(R["n_compute"], R["owned"])
        for peer, idx in R["send"].items():
            idx = np.ascontiguousarray(idx, np.int64)
            self._ck(self.lib.mb2_mg_set_send(self.h, int(peer), _dp(idx), C.c_int64(len(idx))))
        for peer, idx in R["recv"].items():
            assert (np.diff(idx) == 1).all()
            self._ck(self.lib.mb2_mg_set_recv(self.h, int(peer), C.c_int64(int(idx[0])), C.c_int64(len(idx))))
        if manual:
            self._ck(self.lib.mb2_mg_set_manual(self.h, 1))
        self.mg = R

    def mg_init_nccl(self, rank, world, uid):
        self._ck(self.lib.mb2_mg_init_nccl(self.h, rank, world, uid))

    def mg_begin(self, gf, grad=True):
        self._ck(self.lib.mb2_mg_begin(self.h, C.c_double(gf), int(grad)))

    def mg_finish(self, gf, grad=True):
        self._ck(self.lib.mb2_mg_finish(self.h, C.c_double(gf), int(grad)))

    def mg_export(self, peer, grad=True):
        n = len(self.mg["send"][peer])
        self.lib.mb2_mg_ke_doubles.restype = C.c_int64
        ked = self.lib.mb2_mg_ke_doubles(self.h)
        K = np.empty((n, ked)) if grad else None
        R = np.empty((n, self.tdof))
        self._ck(self.lib.mb2_mg_export_send(self.h, int(peer), _dp(K), _dp(R)))
        return K, R

    def mg_import(self, peer, K, R):
        self._ck(self.lib.mb2_mg_import_recv(self.h, int(peer), _dp(K), _dp(R)))

    def sync(self):
        self._ck(self.lib.mb2_sync(self.h))

    def measure_fp64_peak(self, fma):
        t = C.c_double()
        self._ck(self.lib.mb2_measure_fp64_peak(self.h, int(fma), C.byref(t)))
        return t.value


class Plan:
    """mb2_plan_partition: the library's planner for rank `rank` of `world` (the C++ twin of partition.rank_problem)."""

    def __init__(self, v_dofs, dim, NV, world, rank):
        self.lib = load()
        self.lib.mb2_plan_size.restype = C.c_int64
        self.lib.mb2_plan_size.argtypes = [C.c_void_p, C.c_int]
        v = np.ascontiguousarray(v_dofs, np.int32)
        self.v_dofs, self.dim, self.nd = v, dim, v.shape[1] // dim
        self.h = C.c_void_p()
        rc = self.lib.mb2_plan_partition(_dp(v), C.c_int(dim), C.c_int(self.nd), C.c_int64(v.shape[0]), C.c_int64(NV),
                                         C.c_int(world), C.c_int(rank), C.byref(self.h))
        if rc:
            raise Mb2Error("mb2_plan_partition failed (%d)" % rc)
        self.n_compute, self.n_local, self.e0, self.n_peers, self.Nn = (int(self.lib.mb2_plan_size(self.h, k)) for k in range(5))
        self.late_exchange = bool(self.lib.mb2_plan_size(self.h, 5))
        self.local_elems = np.zeros(self.n_local, np.int64)
        self.lib.mb2_plan_local_elems(self.h, _dp(self.local_elems))
        self.owned = np.zeros(self.Nn, np.uint8)
        self.lib.mb2_plan_owned_nodes(self.h, _dp(self.owned))
        self.send, self.recv = {}, {}
        for k in range(self.n_peers):
            peer, ns, rf, nr = C.c_int(), C.c_int64(), C.c_int64(), C.c_int64()
            self.lib.mb2_plan_peer(self.h, k, C.byref(peer), C.byref(ns), C.byref(rf), C.byref(nr))
            if ns.value:
                a = np.zeros(ns.value, np.int64)
                self.lib.mb2_plan_peer_send(self.h, k, _dp(a))
                self.send[peer.value] = a
            if nr.value:
                self.recv[peer.value] = np.arange(rf.value, rf.value + nr.value)

    def apply(self, ctx, nq):
        ctx._ck(self.lib.mb2_plan_apply(ctx.h, self.h, _dp(self.v_dofs), C.c_int(nq)))
        ctx.dim, ctx.nd, ctx.nq, ctx.tdof, ctx.NE, ctx.NV = self.dim, self.nd, nq, self.dim * self.nd, self.n_local, self.Nn * self.dim

    def close(self):
        if self.h:
            self.lib.mb2_plan_free(self.h)
            self.h = C.c_void_p()


def group_init(ctxs):
    arr = (C.c_void_p * len(ctxs))(*[c.h for c in ctxs])
    ctxs[0]._ck(load().mb2_mg_group_init(arr, len(ctxs)))
    return arr


def group_assemble(ctxs, arr, gf, grad=True):
    rc = load().mb2_mg_group_assemble(arr, len(ctxs), C.c_double(gf), int(grad))
    if rc:
        raise Mb2Error("mb2_mg_group_assemble: " + "; ".join(load().mb2_last_error(c.h).decode() for c in ctxs))
