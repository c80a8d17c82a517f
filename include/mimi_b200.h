/*
 * mimi_b200 -- C ABI of the B200-native residual + tangent CSR assembly path of j042/mimi.
 *
 * This is the drop-in boundary: plain C types, caller-owned host buffers, library-owned device
 * buffers, 0 = success / non-zero = error (message via mb2_last_error). A maintainer of the
 * reference binds these from a `mimi::integrators::NonlinearBase` subclass;
 * each entry cites the reference interface it replaces. There is NO CPU fallback: every compute
 * entry fails with MB2_ERR_CUDA when no sm_100a device / CUDA runtime is usable.
 *
 * Conventions shared with the reference:
 *  - all matrices column-major (mfem::DenseMatrix); FP64 throughout;
 *  - vdof ordering is mfem::Ordering::byVDIM (src/mimi/py/py_nonlinear_solid.cpp:63,74): global
 *    vdof = dim*node + component; element vdofs are component-major: v_dofs[c*nd + a];
 *  - entries ADD into the residual and into the CSR values they are given
 *    (src/mimi/integrators/nonlinear_base.hpp:90-151).
 */
#ifndef MIMI_B200_H
#define MIMI_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct mb2_ctx mb2_ctx;

enum {
  MB2_OK = 0,
  MB2_ERR_ARG = 1,      /* bad argument / call order */
  MB2_ERR_CUDA = 2,     /* CUDA runtime error or no device */
  MB2_ERR_UNSUPPORTED = 3,
  MB2_ERR_NUMERIC = 4   /* device-side failure flag (det F <= 0, scalar solve not converged) */
};

/* material ids: src/mimi/materials/materials.hpp (Name()) */
enum {
  MB2_MAT_NEOHOOKEAN = 0, /* CompressibleOgdenNeoHookean, materials.cpp:96-118; params {lambda, mu} */
  MB2_MAT_STVK = 1,       /* StVenantKirchhoff,           materials.cpp:73-94;  params {lambda, mu} */
  MB2_MAT_J2LINEAR = 2,   /* J2Linear, materials.hpp:142-249; params {lambda, mu, K, G, isotropic_hardening,
                             kinematic_hardening, sigma_y}; 19 history doubles per quadrature point */
  MB2_MAT_J2 = 3,         /* J2 with nonlinear isotropic hardening, materials.hpp:259-404; params {lambda, mu, K, G,
                             hardening kind (MB2_HARD_*), hp[10], dt, heat_fraction, density, specific_heat,
                             initial_temperature, melting_temperature}; 11 history doubles per quadrature point:
                             plastic_strain[9], accumulated_plastic_strain, temperature */
  MB2_MAT_J2SIMO = 4,     /* J2Simo (finite strain), materials.hpp:406-579; params as MB2_MAT_J2; 20 history doubles:
                             be_old[9], F_old[9], accumulated_plastic_strain, temperature */
  MB2_MAT_J2LOG = 5       /* J2Log (finite strain, logarithmic), materials.hpp:580-756; params as MB2_MAT_J2; 11 history
                             doubles: Fp_inv[9], accumulated_plastic_strain, temperature */
};
/* hardening laws, src/mimi/materials/material_hardening.hpp. hp[10] per law (unused entries 0):
 *   POWERLAW {sigma_y, n, eps0}; VOCE {sigma_y, sigma_sat, strain_constant};
 *   JC {A, B, n}; JC_RATE {A, B, n, C, eps0_dot}; JC_THERMO_RATE / JC_CONST_T {A, B, n, C, eps0_dot, T_ref, -, m} */
enum {
  MB2_HARD_POWERLAW = 0, MB2_HARD_VOCE = 1, MB2_HARD_JC = 2, MB2_HARD_JC_RATE = 3, MB2_HARD_JC_THERMO_RATE = 4,
  MB2_HARD_JC_CONST_T = 5
};

/* tangent modes */
enum {
  MB2_TANGENT_FD_PARITY = 0, /* forward finite difference in the reference's exact operation order
                                (src/mimi/integrators/nonlinear_solid.cpp:48-76), no FMA: bit-identical to the reference */
  MB2_TANGENT_ANALYTIC = 1   /* "fast mode" (mb2_set_option(ctx, "tangent_mode", MB2_TANGENT_ANALYTIC)): the closed-form
                                tangent the reference's difference quotient approximates, K_e = sum_q B^T (w detJ dP/dF) B,
                                FMA allowed; CompressibleOgdenNeoHookean and StVenantKirchhoff only (MB2_ERR_UNSUPPORTED
                                otherwise). Not bit-comparable: agrees with the reference's K to its finite-difference
                                noise (1e-6 .. 1e-4 relative) and keeps its Newton iteration counts (tests/test_gpu_fast.py).
                                The residual is the same quadrature with contracted FMAs (1e-15 relative).
                                p = 2 hexahedra run the contraction on the FP64 tensor cores ("fast_gemm" 0, default:
                                A fragments built in registers; 2: fed from shared-memory T buffers) or on the CUDA
                                cores ("fast_gemm" 1; same bits as 2). */
};

/* ---- lifetime ------------------------------------------------------------------------------- */
/* device: CUDA ordinal. stream: a cudaStream_t cast to void* (NULL = the library creates its own).
 * Replaces: construction of mimi::integrators::NonlinearSolid (nonlinear_solid.hpp:45-50). */
int mb2_create(mb2_ctx** out, int device, void* stream);
int mb2_destroy(mb2_ctx* ctx);
/* Error text of the last failing call on ctx (ctx may be NULL for create failures). */
const char* mb2_last_error(const mb2_ctx* ctx);
/* Library/ABI version, for the loader check. */
int mb2_version(void);

/* ---- setup: what PrecomputedData holds (src/mimi/utils/precomputed.hpp:58-130) ------------------ */
/* Element connectivity. v_dofs: NE x (dim*nd) int32, component-major per element, byVDIM global ids
 * (ElementData::v_dofs, precomputed.cpp:86). NV = number of vdofs = dim * number of nodes.
 * elem_offset: global id of this rank's first element (0 on one GPU); elements must be passed in
 * ascending global order -- the deterministic scatter adds contributions in that order. */
int mb2_set_connectivity(mb2_ctx* ctx, int dim, int nd, int nq, int64_t NE, int64_t NV, const int32_t* v_dofs);

/* Flat per-quadrature-point tables exactly as the reference precomputes them
 * (PrecomputeElementQuadData, precomputed.cpp:264-330): dN_dX NE x nq x (nd x dim col-major),
 * w and detJ NE x nq (QuadData::integration_weight, det_dX_dxi). Host pointers. */
int mb2_upload_quad_tables(mb2_ctx* ctx, const double* dN_dX, const double* w, const double* detJ);

/* Reference node coordinates (the mesh Nodes GridFunction, byVDIM: x_ref[dim*node + c]) and NURBS weights
 * per node (NULL = B-spline, all 1); what IsoparametricTransformation::PointMat and
 * NURBSFiniteElement::Weights() hold per element (third_party/mfem/fem/eltrans.cpp:406-420). */
int mb2_set_nodes(mb2_ctx* ctx, const double* x_ref, const double* weights);

/* Device-side regeneration of the same tables for the elements [elem_begin, elem_begin+elem_count) of one
 * tensor-product NURBS patch from its knot vectors, in MFEM's arithmetic order
 * (third_party/mfem/mesh/nurbs.cpp:297-376, fem/fe/fe_nurbs.cpp:125-158,269-311, precomputed.cpp:289-321),
 * bit-identical to PrecomputeElementQuadData. knots[d] has ncp[d]+p+1 entries; el_ijk is elem_count x dim,
 * the knot-span index of every element (NURBSExtension::el_to_IJK). quad_order < 0 selects the reference
 * default 2p+3. want_N != 0 also keeps the shape values N (QuadData::N). Call once per patch. */
int mb2_generate_patch_tables(mb2_ctx* ctx, int64_t elem_begin, int64_t elem_count, int p, const int32_t* ncp,
                              const double* const* knots, const int32_t* el_ijk, int quad_order, int want_N);
/* Mass matrix on the device: A += M on the resident CSR values, M = what
 * mass->AddDomainIntegrator(new mfem::VectorMassIntegrator(rho)); mass->Assemble(0) leaves in the BilinearForm
 * (src/mimi/py/py_nonlinear_solid.cpp:154-172; third_party/mfem/fem/bilininteg.cpp:1542-1620), bit for bit: rule
 * IntRules.Get(geom, 2p + OrderW) (quad_order < 0; OrderW = p*dim - 1), per point (N_i N_j) * ((ip.weight * Trans.Weight())
 * * density), points then elements added in ascending order. Needs mb2_set_nodes + mb2_build_pattern; the domain tables
 * are not touched. el_ijk: knot-span index of every element this context computes (one tensor-product patch).
 * In a partitioned run the ghost blocks travel like the K_e blocks. */
int mb2_assemble_mass_device(mb2_ctx* ctx, int p, const int32_t* ncp, const double* const* knots, const int32_t* el_ijk,
                             int quad_order, double density);
/* Viscosity (damping) matrix on the device: A += S, S = what visc->AddDomainIntegrator(new
 * mfem::VectorDiffusionIntegrator(viscosity)); visc->Assemble(0) leaves in the BilinearForm
 * (src/mimi/py/py_nonlinear_solid.cpp:174-191; third_party/mfem/fem/bilininteg.cpp:2854-2934), bit for bit: rule
 * IntRules.Get(geom, 2p + dim - 1) (quad_order < 0), per point w = (ip.weight / Trans.Weight()) * viscosity,
 * dshapedxt = dshape adj(J), pelmat = w * dshapedxt dshapedxt^t (Mult_a_AAt) on every diagonal component block. */
int mb2_assemble_viscosity_device(mb2_ctx* ctx, int p, const int32_t* ncp, const double* const* knots, const int32_t* el_ijk,
                                  int quad_order, double viscosity);
/* Body-force linear form on the device: r += b on the resident vector, b = what
 * rhs->AddDomainIntegrator(new mfem::VectorDomainLFIntegrator(body_force)); rhs->Assemble() produces for a constant
 * force (src/mimi/py/py_nonlinear_solid.cpp:221-241; third_party/mfem/fem/lininteg.cpp:268-308), bit for bit: rule
 * IntRules.Get(geom, 2p) (quad_order < 0), elvect(nd k + s) += ip.weight * (Tr.Weight() * f_k) * shape(s), elements added
 * in ascending order. force: dim values. Essential entries are zeroed by the caller (rhs->SetSubVector(zero_dofs, 0.0))
 * or by mb2_newton_rhs_from_r. */
int mb2_assemble_body_force_device(mb2_ctx* ctx, int p, const int32_t* ncp, const double* const* knots,
                                   const int32_t* el_ijk, int quad_order, const double* force);
/* Traction linear form on the device: r = ((r + f_1) + f_2) + ... over the loaded faces in ascending order, f = what
 * mfem::VectorBoundaryLFIntegrator contributes per boundary element (src/mimi/py/py_nonlinear_solid.cpp:243-276;
 * third_party/mfem/fem/lininteg.cpp:375-408): elvect(ndf k + s) += (t_k * (Tr.Weight() * ip.weight)) * shape(s), bit for
 * bit. Face tables as the reference precomputes boundary data at the integrator's rule (order 2p,
 * PrecomputedData::PrecomputeElementQuadData, precomputed.cpp:264-330): face_vdofs NF x dim*ndf (component-major),
 * N NF x nqf x ndf, w = integration_weight and detW = det_dX_dxi NF x nqf. traction: dim values (one loaded boundary per
 * call). LinearForm::Assemble adds domain integrators first: call mb2_assemble_body_force_device before this one. */
int mb2_assemble_traction_device(mb2_ctx* ctx, int64_t NF, int ndf, int nqf, const int32_t* face_vdofs, const double* N,
                                 const double* w, const double* detW, const double* traction);
/* Shape values N: NE x nq x nd (tests; contact). */
int mb2_download_shape_values(mb2_ctx* ctx, double* N);

/* Copies the device tables back (tests). Any pointer may be NULL. */
int mb2_download_quad_tables(mb2_ctx* ctx, double* dN_dX, double* w_detJ);

/* Sparsity pattern built ON DEVICE from the connectivity (replaces PrepareSparsity,
 * precomputed.cpp:151-218): rows = vdofs, columns ascending. Returns nnz. */
int mb2_build_pattern(mb2_ctx* ctx, int64_t* nnz);
/* Pattern hand-back. I: NV+1 row offsets, J: nnz columns. The 32-bit form fails with
 * MB2_ERR_UNSUPPORTED when nnz does not fit int (the reference's own limit, precomputed.cpp:196-199). */
int mb2_get_pattern_i32(mb2_ctx* ctx, int32_t* I, int32_t* J);
int mb2_get_pattern_i64(mb2_ctx* ctx, int64_t* I, int32_t* J);
/* Element -> CSR value index map in the reference's layout (ElementData::A_ids, precomputed.cpp:183-199),
 * expanded on device from the factored map; for parity tests. A_ids: NE x tdof^2 int32. */
int mb2_get_A_ids_i32(mb2_ctx* ctx, int32_t* A_ids);

/* MaterialBase parameters (materials.hpp:24-40). */
int mb2_set_material(mb2_ctx* ctx, int kind, const double* params, int nparams);

/* History variables (QuadData::material_state, precomputed.hpp:58-69; MaterialBase::CreateState). They live on the
 * device as [NE][nq][mb2_state_size] doubles -- J2Linear: plastic_strain[9] (column-major dim x dim in the first
 * dim*dim slots), beta[9], accumulated_plastic_strain -- allocated and initialised by mb2_set_material (call it after
 * mb2_set_connectivity). mb2_set_state(NULL) re-initialises; get/set are for checkpointing and tests. */
int mb2_state_size(mb2_ctx* ctx);
/* NonlinearBase::dt_ (pushed into MaterialBase::dt_ by every assembly, nonlinear_solid.cpp:154,167): the time step the
 * rate-dependent hardening laws divide by. */
int mb2_set_time_step(mb2_ctx* ctx, double dt);
/* Device failure flag: MB2_ERR_NUMERIC + a message per cause if raised since the last check -- the return-mapping scalar
 * solve (the reference throws, src/mimi/solvers/newton.hpp:81-93, 119-132: "root not bracketed" / "failed to converge"),
 * EliminateRowCol on a non-symmetric pattern, DiagScale without a diagonal. The host-buffer entries check it themselves;
 * of the device-resident entries mb2_newton_solve_gmres_device and mb2_newton_residual_device (when it returns the norm)
 * do, because they synchronise anyway. The asynchronous ones (mb2_assemble_*_device,
 * mb2_newton_residual_and_grad_device) do NOT: a caller that consumes their output without going through the solve
 * must call mb2_check_numeric once per iteration. */
int mb2_check_numeric(mb2_ctx* ctx);
/* Measurement counters of the J2 family (on with mb2_set_option(ctx, "j2_counters", 1); they cost atomics, keep them off
 * in timed runs): out2[0] = return mappings run since the last reset (yielded stress evaluations; an assembly makes
 * NE * nq * (tdof + 1) stress evaluations in all), out2[1] = hardening-law evaluations inside their ScalarSolve calls
 * (src/mimi/solvers/newton.hpp:53-169). bench.py uses them to turn the data-dependent J2 work into executed flops. */
int mb2_get_j2_counters(mb2_ctx* ctx, uint64_t* out2, int reset);
int mb2_get_state(mb2_ctx* ctx, double* state);
int mb2_set_state(mb2_ctx* ctx, const double* state);
/* Commits the history variables at the converged u. Replaces NonlinearSolid::DomainPostTimeAdvance
 * (nonlinear_solid.cpp:179-199) + MaterialBase::Accumulate. No-op for stateless materials. */
int mb2_post_time_advance(mb2_ctx* ctx, const double* u);
int mb2_post_time_advance_device(mb2_ctx* ctx); /* on the resident u */

/* ---- the hot path, host-buffer form (what the integrator virtuals forward to) ------------------- */
/* r += R(u).  Replaces NonlinearSolid::AddDomainResidual (nonlinear_solid.cpp:151-160). */
int mb2_add_domain_residual(mb2_ctx* ctx, const double* u, double* r);
/* r += R(u); A_values += grad_factor * K_fd(u), in the reference's accumulation order.
 * Replaces NonlinearSolid::AddDomainResidualAndGrad (nonlinear_solid.cpp:162-177). A_values holds
 * the nnz values of the CSR whose pattern is mb2_build_pattern's (already containing the mass). */
int mb2_add_domain_residual_and_grad(mb2_ctx* ctx, const double* u, double grad_factor, double* r,
                                     double* A_values);

/* ---- the hot path, device-resident form (solver-on-device callers, bench) ------------------------ */
int mb2_set_u(mb2_ctx* ctx, const double* u_host);                 /* H2D into the resident u */
int mb2_set_r(mb2_ctx* ctx, const double* r_host);                 /* NULL = zero */
int mb2_set_A(mb2_ctx* ctx, const double* A_host);                 /* NULL = zero */
int mb2_get_r(mb2_ctx* ctx, double* r_host);
int mb2_get_A(mb2_ctx* ctx, double* A_host);
int mb2_assemble_residual_device(mb2_ctx* ctx);                    /* r_dev += R(u_dev) */
int mb2_assemble_residual_and_grad_device(mb2_ctx* ctx, double grad_factor);
/* Raw device pointers of the resident arrays (u: NV, r: NV, A: nnz doubles). */
int mb2_device_pointers(mb2_ctx* ctx, void** u, void** r, void** A);

/* Pins a caller-owned host array (e.g. the mfem::SparseMatrix value array) so that the host-buffer entries
 * can stream it chunk by chunk with the copy engines; without it the copies fall back to staged pageable copies. */
int mb2_host_register(mb2_ctx* ctx, void* ptr, size_t bytes);
int mb2_host_unregister(mb2_ctx* ctx, void* ptr);
/* Chunk pipeline of the residual+tangent assembly: elements per chunk and number of chunks. */
int mb2_plan_info(mb2_ctx* ctx, int64_t* chunk_elems, int* nchunks);

/* Milliseconds of the kernels of the last assemble call, from CUDA events on ctx's stream:
 * out[0] = element kernels (summed over chunks), out[1] = CSR/residual gather-scatter (summed), out[2] = start to
 * end of the call; the gather of chunk c overlaps the element kernel of chunk c+1, so [0]+[1] may exceed [2]. */
int mb2_last_timing_ms(mb2_ctx* ctx, float* out3);
/* Kernel launches issued by the last assemble call. */
int mb2_last_launch_count(mb2_ctx* ctx);
/* Synchronise ctx's stream. */
int mb2_sync(mb2_ctx* ctx);

/* ---- rigid-body mortar contact: mimi::integrators::MortarContact (src/mimi/integrators/mortar_contact.{hpp,cpp}) ----
 * One contact boundary per context. Call order: mb2_set_connectivity, mb2_set_nodes (the faces' x_ref comes from the
 * node coordinates, MortarContact::Prepare mortar_contact.cpp:96-107), mb2_build_pattern, mb2_contact_set_faces,
 * mb2_contact_set_rigid_body. */
enum {
  /* closest point on the affine patch o + s t0 (+ t t1), s, t clamped to [lo, hi] (a degree-1 spline);
   * params = o[3], t0[3], t1[3], lo, hi  (unused components 0; 2D uses o[0..1], t0[0..1]) */
  MB2_RIGID_PLANE = 0,
  /* closest point on the circle (2D) / sphere (3D): params = c[3], R; the outward normal is the contact normal */
  MB2_RIGID_SPHERE = 1,
  /* Bezier curve (2D) / tensor-product surface (3D), degree 1..5, what NearestDistanceToSplines::AddSpline holds
   * (coefficients/nearest_distance.hpp:225-244): params = deg_u, deg_v (0 for curves), n_seed, control points
   * [deg_v+1][deg_u+1][dim] (u fastest). Closest point: nearest of (n_seed+1)^(dim-1) samples, then bound-clipped
   * Newton. The reference delegates this search to splinepy: converged results agree to rounding, not bitwise
   * (parity unpinned). The spline's orientation defines the contact normal as in the reference. */
  MB2_RIGID_BEZIER = 2,
  /* B-spline curve / surface with interior knots (open knot vectors; what toy_problem.py's extracted boundaries are):
   * params = deg_u, deg_v, n_seed, ncp_u, ncp_v (1 for curves), knots_u[ncp_u + deg_u + 1], knots_v[ncp_v + deg_v + 1]
   * (absent for curves), control points [ncp_v][ncp_u][dim]. The library splits it into Bezier patches by knot insertion
   * on the host and queries every patch (the nearest wins). MB2_RIGID_BEZIER also accepts several patches of one degree
   * back to back after n_seed. Parity unpinned like MB2_RIGID_BEZIER. */
  MB2_RIGID_BSPLINE = 3,
  /* Rational Bezier patches (splinepy RationalBezier bodies: circular arcs, spherical caps ...): the MB2_RIGID_BEZIER
   * list with dim + 1 doubles per control point, {x[dim], weight} (weight > 0, x not pre-multiplied). */
  MB2_RIGID_RATIONAL_BEZIER = 4,
  /* NURBS curve / surface: the MB2_RIGID_BSPLINE list with {x[dim], weight} per control point. Knot insertion runs on
   * the homogeneous points {w x, w}; every patch is evaluated as numerator / denominator with the quotient rule for
   * the first and second derivatives (oracle: bezier_eval in mimi_oracle_contact.c). Parity unpinned like
   * MB2_RIGID_BEZIER; a rational quarter circle / sphere octant reproduces MB2_RIGID_SPHERE to rounding (tests). */
  MB2_RIGID_NURBS = 5
};
/* The marked boundary faces and their quadrature tables as PrecomputedData holds them for the integrator's name
 * (CreateBoundaryElementQuadData + PrecomputeElementQuadData(name, order, false), precomputed.cpp:243-330):
 * v_dofs NF x (dim*nd_f) component-major byVDIM ids; N NF x nq_f x nd_f (QuadData::N); dN_dxi NF x nq_f x
 * (nd_f x (dim-1)) column-major (QuadData::dN_dxi); w NF x nq_f (integration_weight). Performs MortarContact::Prepare
 * (mortar_contact.cpp:19-133): sorted unique marked vdofs and the reduced nodal numbering. */
int mb2_contact_set_faces(mb2_ctx* ctx, int64_t NF, int nd_f, int nq_f, const int32_t* v_dofs, const double* N,
                          const double* dN_dxi, const double* w);
/* The rigid body (replaces the NearestDistanceBase scene handed to the MortarContact constructor,
 * mortar_contact.hpp:73-79) and NearestDistanceBase::coefficient_ (the penalty). May be called again every time step
 * (moving body). */
int mb2_contact_set_rigid_body(mb2_ctx* ctx, int kind, const double* params, int nparams, double penalty);
int mb2_contact_n_marked_nodes(mb2_ctx* ctx, int64_t* n);
/* MortarContact::marked_boundary_v_dofs_ (sorted, unique): dim * n_marked_nodes entries. */
int mb2_contact_get_marked_v_dofs(mb2_ctx* ctx, int32_t* out);
/* Boundary ElementData::A_ids (precomputed.cpp:202-217), NF x (dim*nd_f)^2, for parity tests. */
int mb2_contact_get_A_ids_i32(mb2_ctx* ctx, int32_t* A_ids);
/* r += R_contact(u). Replaces MortarContact::AddBoundaryResidual (mortar_contact.cpp:297-352). */
int mb2_add_boundary_residual(mb2_ctx* ctx, const double* u, double* r);
/* r += R_contact(u); A_values += K_contact * grad_factor (FD tangent with frozen pressure). Replaces
 * MortarContact::AddBoundaryResidualAndGrad (mortar_contact.cpp:354-421). Only the CSR rows of the marked nodes are
 * copied to the device and back. */
int mb2_add_boundary_residual_and_grad(mb2_ctx* ctx, const double* u, double grad_factor, double* r, double* A_values);
/* Device-resident forms: act on the resident u, r, A (after mb2_assemble_*_device for domain + contact in one go). */
int mb2_contact_residual_device(mb2_ctx* ctx);
int mb2_contact_residual_and_grad_device(mb2_ctx* ctx, double grad_factor);
/* average_gap_, area_, average_pressure_ of the last assembly (n_marked_nodes each; any pointer may be NULL). */
int mb2_contact_get_nodal(mb2_ctx* ctx, double* gap, double* area, double* pressure);
/* out5 = {last_area_, last_pressure_, last_force_[0..2]} of the last assembly (what BoundaryPostTimeAdvance records,
 * mortar_contact.cpp:469-488). Summed per face, then across faces in ascending id: 1e-12 relative to the reference. */
int mb2_contact_last_scalars(mb2_ctx* ctx, double* out5);
/* sqrt(sum of g_n^2 over penetrating quadrature points). Replaces MortarContact::GapNorm (mortar_contact.cpp:423-467). */
int mb2_contact_gap_norm(mb2_ctx* ctx, const double* u, double* out);

/* ---- Newton-iteration sweeps on the resident CSR ------------------------------------
 * What mimi::operators::NonlinearSolid::ResidualAndGrad (src/mimi/operators/nonlinear_solid.cpp:240-283) does around
 * the integrators, with forms::Nonlinear::AddMultGrad's essential-dof handling (src/mimi/forms/nonlinear.hpp:111-115),
 * so that an implicit step keeps y and J on the device:
 *   u := x + fac0 a;  y := M a (+ S (v + fac1 a));  J := M;  domain integrators (y += R(u), J += fac0 K(u));
 *   y[ess] := 0, J.EliminateRowCol(ess) (row/column zeroed, diagonal 1);  contact integrators if a rigid body is set,
 *   then the same elimination again;  J += fac1 S;  y -= rhs;  y[ess] := 0.
 * Results stay resident (mb2_get_r / mb2_get_A / mb2_device_pointers) and are bit-identical to the reference's. */
int mb2_set_mass(mb2_ctx* ctx, const double* M_values);          /* nnz values on mb2_build_pattern's pattern */
/* The host-CSR consumer without the per-step CSR upload: the mass stays resident (mb2_set_mass / mb2_newton_mass_from_A)
 * and the call does what operators::NonlinearSolid::ResidualAndGrad does around the domain integrator
 * (src/mimi/operators/nonlinear_solid.cpp:257: J.data <- mass.data; then AddDomainResidualAndGrad): r += R(u),
 * A_out = M + grad_factor K_fd(u), bit-identical to mb2_add_domain_residual_and_grad on an A preloaded with M.
 * Host traffic per call: u, r in (2 NV doubles), r and A_out out (NV + nnz doubles). */
int mb2_domain_residual_and_grad_on_mass(mb2_ctx* ctx, const double* u, double grad_factor, double* r, double* A_out);
/* The resident CSR values (e.g. after mb2_set_A(NULL) + mb2_assemble_mass_device) become the mass of the Newton
 * sweeps without leaving the device; form_system_matrix != 0 applies BilinearForm::FormSystemMatrix(zero_dofs, .)
 * (py_nonlinear_solid.cpp:173) for the dofs of mb2_set_essential_dofs: rows / columns zeroed, diagonal one. */
int mb2_newton_mass_from_A(mb2_ctx* ctx, int form_system_matrix);
int mb2_newton_get_mass(mb2_ctx* ctx, double* M_host);
int mb2_newton_viscosity_from_A(mb2_ctx* ctx, int form_system_matrix);   /* same for the viscosity matrix */
int mb2_newton_get_viscosity(mb2_ctx* ctx, double* S_host);
/* The resident vector r (e.g. after mb2_set_r(NULL) + mb2_assemble_body_force_device) becomes the rhs of the Newton
 * sweeps; zero_essential != 0 applies rhs->SetSubVector(zero_dofs, 0.0) (py_nonlinear_solid.cpp:281). */
int mb2_newton_rhs_from_r(mb2_ctx* ctx, int zero_essential);
int mb2_set_viscosity(mb2_ctx* ctx, const double* S_values);     /* nnz values, NULL = none */
int mb2_set_rhs(mb2_ctx* ctx, const double* rhs);                /* NV, NULL = none (linear forms: body force, traction) */
int mb2_set_essential_dofs(mb2_ctx* ctx, const int32_t* tdofs, int64_t n);
int mb2_newton_residual_and_grad_device(mb2_ctx* ctx, const double* x, const double* v, const double* a, double fac0,
                                        double fac1);
/* The residual-only evaluation of the line search (mimi::operators::NonlinearSolid::Mult(a, y),
 * src/mimi/operators/nonlinear_solid.cpp:172-205; two per Newton iteration, src/mimi/solvers/newton.cpp:138-157) on
 * the resident arrays, and its l2 norm (NULL = not wanted): y bit-identical to the reference's, the norm a
 * deterministic tree reduction. */
int mb2_newton_residual_device(mb2_ctx* ctx, const double* x, const double* v, const double* a, double fac0, double fac1,
                               double* norm_out);
/* the reference's iterative linear solver on the resident arrays, J c = y with J = the
 * resident CSR values and y = the resident residual (what mb2_newton_residual_and_grad_device leaves behind):
 * mfem::GMRESSolver (restart kdim; the reference uses 50, rel 1e-8, abs 1e-12, 300 iterations) left-preconditioned by
 * mfem::DSmoother, iterative_mode == false (src/mimi/py/py_nonlinear_solid.cpp:330-339,
 * third_party/mfem/linalg/solvers.cpp:931-1160). The solution stays on the device (mb2_newton_get_solution copies it out).
 * Floating-point parity: SpMV and the preconditioner keep the reference's expressions; the dot products are
 * deterministic tree reductions, so iterates agree with the reference to rounding (solution within the solver
 * tolerance, iteration count within a few), not bitwise. */
int mb2_newton_solve_gmres_device(mb2_ctx* ctx, double rel_tol, double abs_tol, int max_iter, int kdim, int* iterations,
                                  double* final_norm, int* converged);
/* The reference's DEFAULT linear solver is mfem::UMFPackSolver (src/mimi/py/py_nonlinear_solid.cpp:340-343; Newton drives it
 * as prec->SetOperator(J); prec->Mult(r, c), src/mimi/solvers/newton.cpp:125-132). Direct option on the resident arrays:
 * J c = y by cuSOLVER's sparse QR (cusolverSpDcsrlsvqr with symrcm reordering, loaded with dlopen), solution resident like the
 * iterative option's (mb2_newton_get_solution). A library call, not a kernel of this path: meant for the mesh sizes a direct
 * solve is used at (cuSOLVER indexes with 32 bits); agrees with an LU solve to rounding (tests: 1e-10 relative). A singular
 * Jacobian returns MB2_ERR_NUMERIC and, if singular_row != NULL, the row. */
int mb2_newton_solve_direct_device(mb2_ctx* ctx, int* singular_row);
int mb2_newton_get_solution(mb2_ctx* ctx, double* c_host);

/* ---- multi-GPU: element partition with raw element-matrix exchange ------------------
 * One context per GPU / rank. The rank's connectivity (mb2_set_connectivity) lists its own elements first
 * (global ascending order) and then its GHOST elements: elements computed by higher ranks that touch a node this
 * rank owns. Node ids stay global. mb2_set_partition (before mb2_build_pattern): elements [0, n_compute) are
 * computed here; owned_nodes[Nn] != 0 marks the nodes whose CSR rows / residual entries this rank assembles
 * (owner = rank of the node's lowest element, mimi_b200/partition.py). Replaces the reference's thread-private
 * accumulators + ordered reduction (src/mimi/integrators/nonlinear_base.hpp:90-151) across devices; results stay
 * bit-identical to its single-thread order because raw K_e / R_e travel, not partial sums. */
int mb2_set_partition(mb2_ctx* ctx, int64_t n_compute, const uint8_t* owned_nodes);
/* Elements (local indices, ascending, < n_compute) whose K_e / R_e go to rank `peer`. */
int mb2_mg_set_send(mb2_ctx* ctx, int peer, const int64_t* local_elems, int64_t n);
/* Contiguous block of ghost elements [first_local_elem, +n) that rank `peer` fills. */
int mb2_mg_set_recv(mb2_ctx* ctx, int peer, int64_t first_local_elem, int64_t n);
/* NCCL transport (libnccl is loaded with dlopen): rank 0 creates the 128-byte id, every rank joins. After this the
 * ordinary entries (mb2_assemble_*_device, mb2_add_domain_*) perform the exchange inside the call. */
int mb2_mg_unique_id(void* out128);
int mb2_mg_init_nccl(mb2_ctx* ctx, int rank, int world, const void* unique_id128);
/* Manual transport: the host moves the packed buffers (tests emulating ranks on one GPU; foreign transports).
 * Sequence per assembly: mb2_mg_begin -> mb2_mg_export_send on the sender / mb2_mg_import_recv on the owner ->
 * mb2_mg_finish. Ke: [n][mb2_mg_ke_doubles] (NULL in residual-only mode), Re: [n][tdof]. */
int mb2_mg_set_manual(mb2_ctx* ctx, int on);
int mb2_mg_begin(mb2_ctx* ctx, double grad_factor, int grad);
int mb2_mg_export_send(mb2_ctx* ctx, int peer, double* Ke_host, double* Re_host);
int mb2_mg_import_recv(mb2_ctx* ctx, int peer, const double* Ke_host, const double* Re_host);
int mb2_mg_finish(mb2_ctx* ctx, double grad_factor, int grad);
int64_t mb2_mg_ke_doubles(mb2_ctx* ctx);

/* ---- partition planner (pure host code; mimi_b200/partition.py is its test oracle) --------------------------------
 * Plans rank `rank` of `world` for a GLOBAL connectivity (v_dofs NE x dim*nd as in mb2_set_connectivity): balanced
 * contiguous element ranges, node owner = rank of the node's lowest element, ghosts = elements of higher ranks touching an
 * owned node, send lists = own elements touching a lower rank's node. Replaces the reference's thread chunks
 * (src/mimi/utils/n_thread_exe.hpp:36-97) across devices. */
typedef struct mb2_plan mb2_plan;
int mb2_plan_partition(const int32_t* v_dofs, int dim, int nd, int64_t NE, int64_t NV, int world, int rank, mb2_plan** out);
void mb2_plan_free(mb2_plan* plan);
/* what: 0 elements computed here, 1 local elements (own + ghosts), 2 global id of the first own element, 3 peers, 4 nodes,
 * 5 = 1 if some rank's outgoing elements lie in the second half of its range (mb2_plan_apply then sets "mg_exchange_late") */
int64_t mb2_plan_size(const mb2_plan* plan, int what);
int mb2_plan_local_elems(const mb2_plan* plan, int64_t* global_ids);          /* own elements, then ghosts */
int mb2_plan_owned_nodes(const mb2_plan* plan, uint8_t* owned);               /* one byte per node */
int mb2_plan_peer(const mb2_plan* plan, int k, int* peer, int64_t* n_send, int64_t* recv_first, int64_t* n_recv);
int mb2_plan_peer_send(const mb2_plan* plan, int k, int64_t* send_local_elems);
/* mb2_set_connectivity (the plan's local elements, global node ids) + mb2_set_partition + mb2_mg_set_send / _recv. */
int mb2_plan_apply(mb2_ctx* ctx, const mb2_plan* plan, const int32_t* v_dofs_global, int nq);
/* One process driving several GPUs (ctxs[r] = rank r's context, each after mb2_plan_apply ... mb2_build_pattern): the
 * packed element matrices travel by cudaMemcpyPeerAsync (NVLink peer-to-peer where available). The C++ route to more than
 * one GPU for a caller without MPI / torchrun (adapter/b200_multigpu.hpp). grad != 0: r += R(u), A += grad_factor K(u) on
 * every rank's resident arrays (mb2_set_u / _r / _A per context before, mb2_get_r / _A after); 0: residual only. */
int mb2_mg_group_init(mb2_ctx** ctxs, int n);
int mb2_mg_group_assemble(mb2_ctx** ctxs, int n, double grad_factor, int grad);

/* No counterpart in the reference: added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid allocation). d_out[0] = sum_i d_a[i] d_b[i]
 * over n doubles, all three DEVICE pointers of the context's device, with the deterministic reduction the library's own
 * Newton norm (mb2_newton_residual_device) and GMRES inner products use -- the same bits as those. The call allocates
 * nothing (no cudaMalloc / cudaFree, no allocating library routine): the block partials live in a file-scope device array
 * of MB2_NOALLOC_MAX_N / 4096 = 16384 doubles (128 KiB held by every GPU that loads the library until exit), so
 * n > MB2_NOALLOC_MAX_N is rejected with MB2_ERR_ARG and a message. It only enqueues two kernels on the context's stream
 * and does not synchronise; the first call on a device also resolves the array's device address (cudaGetSymbolAddress: no
 * allocation, but make that first call outside a graph capture). All calls share the array: nothing has to be cleared
 * between calls (every partial that is read has been rewritten), but calls that overlap on one GPU (two streams, two
 * contexts) race on it -- serialise them. */
#define MB2_NOALLOC_MAX_N (64 * 1024 * 1024)
int mb2_dot_noalloc(mb2_ctx* ctx, const double* d_a, const double* d_b, int64_t n, double* d_out);

/* Tuning / A-B switches (never change results): "elem_variant" 0 = default kernels, 1 = one-thread-per-dof FD
 * kernel; "chunk_elems" = elements per pipeline chunk (0 = derive from "chunk_mb", the K_e buffer budget in MiB);
 * "park_acc" 0 = FD kernels keep their accumulators in registers / local memory instead of tensor memory; "res_variant"
 * 1 = first form of the residual-only kernel; "mg_exchange_late" 1 = the NCCL exchange of a partitioned assembly is
 * enqueued behind the rank's last element kernel instead of behind its pack kernels (set it on EVERY rank when interface
 * elements are spread through the element ranges: NCCL's kernels hold SMs while they wait for a peer);
 * "tangent_mode", "fast_gemm", "j2_counters": see MB2_TANGENT_ANALYTIC and mb2_get_j2_counters. */
int mb2_set_option(mb2_ctx* ctx, const char* name, int value);

/* FP64 pipe micro-benchmark used for the roofline denominator: runs `iters` dependent-chain-free
 * DMUL+DADD pairs (fma = 0), DFMA (fma = 1) or FP64 tensor-core mma.sync.m8n8k4 (fma = 2) per thread on a full grid;
 * returns TFLOP/s. */
int mb2_measure_fp64_peak(mb2_ctx* ctx, int fma, double* tflops);

#ifdef __cplusplus
}
#endif
#endif /* MIMI_B200_H */
