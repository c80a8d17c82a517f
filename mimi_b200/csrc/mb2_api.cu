// C ABI of libmimi_b200 (include/mimi_b200.h): context lifetime, uploads, hand-backs and the
// host-buffer forms of the integrator entry points.
#include <cmath>
#include <cstring>
#include <new>

#include "mb2_ctx.h"

static thread_local std::string g_create_err;

#define CHECK_CTX(c)          \
  if (!(c)) return MB2_ERR_ARG
#define REQUIRE(cond, msg)    \
  if (!(cond)) {              \
    c->err = (msg);           \
    return MB2_ERR_ARG;       \
  }

template <typename T>
static void dfree(T*& p) {
  if (p) cudaFree(p);
  p = nullptr;
}

static void free_problem(mb2_ctx* c) {
  dfree(c->d_vdofs); dfree(c->d_enodes); dfree(c->d_dNdX); dfree(c->d_wdetJ); dfree(c->d_N); dfree(c->d_xref); dfree(c->d_nodew);
  dfree(c->d_n2e_ptr); dfree(c->d_n2e_elem); dfree(c->d_n2e_loc); dfree(c->d_nnbr); dfree(c->d_nodeptr);
  dfree(c->d_nbr); dfree(c->d_rank); dfree(c->d_I); dfree(c->d_J);
  dfree(c->d_u); dfree(c->d_r); dfree(c->d_A); dfree(c->d_Re); dfree(c->d_state);
  mb2::free_plan(c);
  mb2::free_mg(c);
  mb2::free_contact(c);
  mb2::free_newton(c);
  dfree(c->d_owned);
  c->have_pattern = false;
  c->nnz = 0;
}

extern "C" {

int mb2_version(void) { return 100; }

const char* mb2_last_error(const mb2_ctx* ctx) { return ctx ? ctx->err.c_str() : g_create_err.c_str(); }

int mb2_create(mb2_ctx** out, int device, void* stream) {
  if (!out) return MB2_ERR_ARG;
  *out = nullptr;
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev <= device || device < 0) {
    g_create_err = std::string("no usable CUDA device: ") + (e != cudaSuccess ? cudaGetErrorString(e) : "bad ordinal") +
                   " (libmimi_b200 has no CPU fallback)";
    return MB2_ERR_CUDA;
  }
  cudaDeviceProp prop;
  e = cudaGetDeviceProperties(&prop, device);
  if (e != cudaSuccess || prop.major < 10) {
    g_create_err = "device is not sm_100-class; this library is built for sm_100a only";
    return MB2_ERR_CUDA;
  }
  mb2_ctx* c = new (std::nothrow) mb2_ctx();
  if (!c) return MB2_ERR_ARG;
  c->device = device;
  c->sm_count = prop.multiProcessorCount;
  cudaSetDevice(device);
  if (stream) {
    c->stream = (cudaStream_t)stream;
  } else {
    if (cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking) != cudaSuccess) {
      g_create_err = "cudaStreamCreate failed";
      delete c;
      return MB2_ERR_CUDA;
    }
    c->own_stream = true;
  }
  for (auto& ev : c->ev) cudaEventCreate(&ev);
  cudaMalloc(&c->d_flag, sizeof(int));
  *out = c;
  return MB2_OK;
}

int mb2_destroy(mb2_ctx* c) {
  CHECK_CTX(c);
  cudaSetDevice(c->device);
  cudaStreamSynchronize(c->stream);
  free_problem(c);
  dfree(c->d_flag);
  dfree(c->d_cnt);
  for (auto& ev : c->ev) if (ev) cudaEventDestroy(ev);
  if (c->own_stream) cudaStreamDestroy(c->stream);
  delete c;
  return MB2_OK;
}

int mb2_set_connectivity(mb2_ctx* c, int dim, int nd, int nq, int64_t NE, int64_t NV, const int32_t* v_dofs) {
  CHECK_CTX(c);
  REQUIRE(dim == 2 || dim == 3, "dim must be 2 or 3");
  REQUIRE(nd > 0 && nq > 0 && NE > 0 && NV > 0 && v_dofs, "bad connectivity arguments");
  REQUIRE(NV % dim == 0, "NV must be dim * number of nodes");
  MB2_CUDA(cudaSetDevice(c->device));
  free_problem(c);
  c->dim = dim; c->nd = nd; c->nq = nq; c->tdof = dim * nd; c->NE = NE; c->NV = NV; c->Nn = NV / dim; c->n_compute = NE;
  MB2_CUDA(cudaMalloc(&c->d_vdofs, sizeof(int32_t) * NE * c->tdof));
  MB2_CUDA(cudaMemcpyAsync(c->d_vdofs, v_dofs, sizeof(int32_t) * NE * c->tdof, cudaMemcpyHostToDevice, c->stream));
  MB2_CUDA(cudaMalloc(&c->d_u, sizeof(double) * NV));
  MB2_CUDA(cudaMalloc(&c->d_r, sizeof(double) * NV));
  MB2_CUDA(cudaMalloc(&c->d_Re, sizeof(double) * NE * c->tdof));
  MB2_CUDA(cudaMemsetAsync(c->d_u, 0, sizeof(double) * NV, c->stream));
  MB2_CUDA(cudaMemsetAsync(c->d_r, 0, sizeof(double) * NV, c->stream));
  MB2_CUDA(cudaStreamSynchronize(c->stream));
  return MB2_OK;
}

__global__ void k_wdetj(int64_t n, const double* __restrict__ w, const double* __restrict__ dj, double* __restrict__ o) {
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i < n) o[i] = w[i] * dj[i];  // q.integration_weight * q.det_dX_dxi, nonlinear_solid.hpp:79
}

int mb2_upload_quad_tables(mb2_ctx* c, const double* dN_dX, const double* w, const double* detJ) {
  CHECK_CTX(c);
  REQUIRE(c->NE > 0, "call mb2_set_connectivity first");
  REQUIRE(dN_dX && w && detJ, "null table pointer");
  MB2_CUDA(cudaSetDevice(c->device));
  const int64_t nqp = c->NE * c->nq;
  if (!c->d_dNdX) MB2_CUDA(cudaMalloc(&c->d_dNdX, sizeof(double) * nqp * c->tdof));
  if (!c->d_wdetJ) MB2_CUDA(cudaMalloc(&c->d_wdetJ, sizeof(double) * nqp));
  double *d_w = nullptr, *d_dj = nullptr;
  MB2_CUDA(cudaMalloc(&d_w, sizeof(double) * nqp));
  MB2_CUDA(cudaMalloc(&d_dj, sizeof(double) * nqp));
  MB2_CUDA(cudaMemcpyAsync(c->d_dNdX, dN_dX, sizeof(double) * nqp * c->tdof, cudaMemcpyHostToDevice, c->stream));
  MB2_CUDA(cudaMemcpyAsync(d_w, w, sizeof(double) * nqp, cudaMemcpyHostToDevice, c->stream));
  MB2_CUDA(cudaMemcpyAsync(d_dj, detJ, sizeof(double) * nqp, cudaMemcpyHostToDevice, c->stream));
  k_wdetj<<<(unsigned)((nqp + 255) / 256), 256, 0, c->stream>>>(nqp, d_w, d_dj, c->d_wdetJ);
  MB2_CUDA(cudaGetLastError());
  MB2_CUDA(cudaStreamSynchronize(c->stream));
  cudaFree(d_w);
  cudaFree(d_dj);
  return MB2_OK;
}

int mb2_download_quad_tables(mb2_ctx* c, double* dN_dX, double* w_detJ) {
  CHECK_CTX(c);
  REQUIRE(c->d_dNdX && c->d_wdetJ, "no tables on device");
  const int64_t nqp = c->NE * c->nq;
  if (dN_dX) MB2_CUDA(cudaMemcpyAsync(dN_dX, c->d_dNdX, sizeof(double) * nqp * c->tdof, cudaMemcpyDeviceToHost, c->stream));
  if (w_detJ) MB2_CUDA(cudaMemcpyAsync(w_detJ, c->d_wdetJ, sizeof(double) * nqp, cudaMemcpyDeviceToHost, c->stream));
  MB2_CUDA(cudaStreamSynchronize(c->stream));
  return MB2_OK;
}

__global__ void k_fill_strided(int64_t n, int stride, int off, double v, double* __restrict__ a) {
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i < n) a[i * stride + off] = v;
}

// History variables as MaterialBase::CreateState initialises them (materials.cpp:120-133): all zero for J2Linear.
static int alloc_state(mb2_ctx* c) {
  dfree(c->d_state);
  c->ns = mb2::mat_state_size(c->mat_kind);
  if (c->ns == 0 || c->NE == 0) return MB2_OK;
  const size_t bytes = sizeof(double) * (size_t)c->NE * c->nq * c->ns;
  MB2_CUDA(cudaMalloc(&c->d_state, bytes));
  MB2_CUDA(cudaMemsetAsync(c->d_state, 0, bytes, c->stream));
  const int64_t n = c->NE * c->nq;
  const unsigned gb = (unsigned)((n + 255) / 256);
  if (c->mat_kind == MB2_MAT_J2) {  // J2::CreateState (materials.cpp:151-166): temperature = initial_temperature_
    k_fill_strided<<<gb, 256, 0, c->stream>>>(n, c->ns, 10, c->mp.T0, c->d_state);
    MB2_CUDA(cudaGetLastError());
  } else if (c->mat_kind == MB2_MAT_J2SIMO) {  // J2Simo::CreateState (materials.cpp:184-203): be_old = F_old = I
    for (int k = 0; k < c->dim; k++) {
      k_fill_strided<<<gb, 256, 0, c->stream>>>(n, c->ns, k * (c->dim + 1), 1.0, c->d_state);
      k_fill_strided<<<gb, 256, 0, c->stream>>>(n, c->ns, 9 + k * (c->dim + 1), 1.0, c->d_state);
    }
    k_fill_strided<<<gb, 256, 0, c->stream>>>(n, c->ns, 19, c->mp.T0, c->d_state);
    MB2_CUDA(cudaGetLastError());
  } else if (c->mat_kind == MB2_MAT_J2LOG) {  // J2Log::CreateState (materials.cpp:221-240): Fp_inv = I
    for (int k = 0; k < c->dim; k++)
      k_fill_strided<<<gb, 256, 0, c->stream>>>(n, c->ns, k * (c->dim + 1), 1.0, c->d_state);
    k_fill_strided<<<gb, 256, 0, c->stream>>>(n, c->ns, 10, c->mp.T0, c->d_state);
    MB2_CUDA(cudaGetLastError());
  }
  return MB2_OK;
}

int mb2_set_material(mb2_ctx* c, int kind, const double* params, int nparams) {
  CHECK_CTX(c);
  REQUIRE(kind >= MB2_MAT_NEOHOOKEAN && kind <= MB2_MAT_J2LOG, "unknown material kind");
  REQUIRE(params && nparams >= 2, "material needs {lambda, mu}");
  mb2::MatP m;
  m.lambda = params[0];
  m.mu = params[1];
  if (kind == MB2_MAT_J2LINEAR) {
    REQUIRE(nparams >= 7, "J2Linear needs {lambda, mu, K, G, isotropic_hardening, kinematic_hardening, sigma_y}");
    REQUIRE(c->NE > 0, "call mb2_set_connectivity before setting a material with history variables");
    m.K = params[2]; m.G = params[3]; m.iso = params[4]; m.kin = params[5]; m.sigma_y = params[6];
    // materials.hpp:163-166: const members initialised with std::sqrt
    m.sqrt_6 = std::sqrt(6.0);
    m.sqrt_3_2 = std::sqrt(3.0 / 2.0);
    m.sqrt_2_3 = std::sqrt(2.0 / 3.0);
  }
  if (kind == MB2_MAT_J2 || kind == MB2_MAT_J2SIMO || kind == MB2_MAT_J2LOG) {
    REQUIRE(nparams >= 21, "J2 needs {lambda, mu, K, G, hardening kind, hp[10], dt, heat_fraction, density, specific_heat, "
                           "initial_temperature, melting_temperature}");
    REQUIRE(c->NE > 0, "call mb2_set_connectivity before setting a material with history variables");
    m.K = params[2]; m.G = params[3];
    m.hard_kind = (int)params[4];
    REQUIRE(m.hard_kind >= 0 && m.hard_kind <= 5, "unknown hardening law");
    for (int i = 0; i < 10; i++) m.hp[i] = params[5 + i];
    m.dt = params[15]; m.heat_fraction = params[16]; m.density = params[17]; m.specific_heat = params[18];
    m.T0 = params[19];
    const double Tmelt = params[20];
    m.sqrt_3_2 = std::sqrt(3.0 / 2.0);  // materials.hpp:290
    if (m.hard_kind >= mb2::HARD_JC_THERMO_RATE) {
      // J2::Setup -> InitializeTemperature / Validate (materials.cpp:135-149, material_hardening.hpp:236-257,303-321)
      m.hp[6] = Tmelt;
      REQUIRE(!(m.hp[5] > m.hp[6]), "reference temperature can't be bigger than melting temperature");
      if (m.hard_kind == mb2::HARD_JC_CONST_T) {
        m.hp[8] = 1.0 - std::pow((m.T0 - m.hp[5]) / (m.hp[6] - m.hp[5]), m.hp[7]);
        REQUIRE(m.hp[8] > 0.0, "Invalid temperature contribution");
      }
    }
    m.flag = c->d_flag;
  }
  const bool realloc_state = (kind != c->mat_kind) || (mb2::mat_state_size(kind) && !c->d_state);
  c->mat_kind = kind;
  { unsigned long long* keep = c->mp.cnt; c->mp = m; c->mp.cnt = keep; }
  if (realloc_state) {
    MB2_CUDA(cudaSetDevice(c->device));
    return alloc_state(c);
  }
  return MB2_OK;
}

int mb2_set_time_step(mb2_ctx* c, double dt) {
  CHECK_CTX(c);
  c->mp.dt = dt;
  return MB2_OK;
}

// Device failure flag: every code has its own message, and the flag is cleared once reported.
//   1, 2  return-mapping scalar solve (newton.hpp:81-93,119-132 throw in the reference)
//   3     EliminateRowCol on a non-symmetric pattern (sparsemat.cpp "EliminateRowCol() #2")
//   4     DSmoother / DiagScale without a diagonal entry
extern "C++" {
namespace mb2 {
int consume_flag(mb2_ctx* c) {
  int flag = 0;
  MB2_CUDA(cudaMemcpyAsync(&flag, c->d_flag, sizeof(int), cudaMemcpyDeviceToHost, c->stream));
  MB2_CUDA(cudaStreamSynchronize(c->stream));
  if (!flag) return MB2_OK;
  MB2_CUDA(cudaMemsetAsync(c->d_flag, 0, sizeof(int), c->stream));
  switch (flag) {
    case 1: c->err = "ScalarSolve: root not bracketed by input bounds"; break;
    case 2: c->err = "ScalarSolve: failed to converge in allotted iterations"; break;
    case 3: c->err = "EliminateRowCol: the sparsity pattern is not symmetric (column entry without its row twin)"; break;
    case 4: c->err = "Diagonal not found in SparseMatrix::DiagScale"; break;
    default: c->err = "device failure flag " + std::to_string(flag); break;
  }
  return MB2_ERR_NUMERIC;
}
}  // namespace mb2
}  // extern "C++"
int mb2_check_numeric(mb2_ctx* c) {
  CHECK_CTX(c);
  return mb2::consume_flag(c);
}

int mb2_get_j2_counters(mb2_ctx* c, uint64_t* out2, int reset) {
  CHECK_CTX(c);
  REQUIRE(out2 && c->d_cnt, "counters are off: mb2_set_option(ctx, \"j2_counters\", 1)");
  MB2_CUDA(cudaStreamSynchronize(c->stream));
  MB2_CUDA(cudaMemcpy(out2, c->d_cnt, 2 * sizeof(unsigned long long), cudaMemcpyDeviceToHost));
  if (reset) MB2_CUDA(cudaMemset(c->d_cnt, 0, 2 * sizeof(unsigned long long)));
  return MB2_OK;
}
int mb2_state_size(mb2_ctx* c) { return c ? c->ns : 0; }
int mb2_get_state(mb2_ctx* c, double* state) {
  CHECK_CTX(c);
  REQUIRE(c->d_state && state, "material has no history variables");
  MB2_CUDA(cudaMemcpyAsync(state, c->d_state, sizeof(double) * (size_t)c->NE * c->nq * c->ns, cudaMemcpyDeviceToHost, c->stream));
  MB2_CUDA(cudaStreamSynchronize(c->stream));
  return MB2_OK;
}
int mb2_set_state(mb2_ctx* c, const double* state) {
  CHECK_CTX(c);
  REQUIRE(c->d_state, "material has no history variables");
  if (state) MB2_CUDA(cudaMemcpyAsync(c->d_state, state, sizeof(double) * (size_t)c->NE * c->nq * c->ns, cudaMemcpyHostToDevice, c->stream));
  else return alloc_state(c);
  return MB2_OK;
}

int mb2_post_time_advance_device(mb2_ctx* c) {
  CHECK_CTX(c);
  REQUIRE(c->d_dNdX, "quadrature tables missing");
  REQUIRE(c->mat_kind >= 0, "material not set");
  MB2_CUDA(cudaSetDevice(c->device));
  return mb2::launch_post_time_advance(c);
}
int mb2_post_time_advance(mb2_ctx* c, const double* u) {
  CHECK_CTX(c);
  REQUIRE(u, "null vector");
  int rc;
  if ((rc = mb2_set_u(c, u))) return rc;
  if ((rc = mb2_post_time_advance_device(c))) return rc;
  return mb2_sync(c);
}

int mb2_build_pattern(mb2_ctx* c, int64_t* nnz) {
  CHECK_CTX(c);
  REQUIRE(c->NE > 0, "call mb2_set_connectivity first");
  MB2_CUDA(cudaSetDevice(c->device));
  if (!c->have_pattern) {
    int rc = mb2::build_pattern(c);
    if (rc) return rc;
    MB2_CUDA(cudaMalloc(&c->d_A, sizeof(double) * c->nnz));
    MB2_CUDA(cudaMemsetAsync(c->d_A, 0, sizeof(double) * c->nnz, c->stream));
  }
  if (nnz) *nnz = c->nnz;
  return MB2_OK;
}

__global__ void k_i64_to_i32(int64_t n, const int64_t* __restrict__ in, int32_t* __restrict__ out) {
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i < n) out[i] = (int32_t)in[i];
}

int mb2_get_pattern_i64(mb2_ctx* c, int64_t* I, int32_t* J) {
  CHECK_CTX(c);
  REQUIRE(c->have_pattern, "call mb2_build_pattern first");
  if (I) MB2_CUDA(cudaMemcpyAsync(I, c->d_I, sizeof(int64_t) * (c->NV + 1), cudaMemcpyDeviceToHost, c->stream));
  if (J) MB2_CUDA(cudaMemcpyAsync(J, c->d_J, sizeof(int32_t) * c->nnz, cudaMemcpyDeviceToHost, c->stream));
  MB2_CUDA(cudaStreamSynchronize(c->stream));
  return MB2_OK;
}

int mb2_get_pattern_i32(mb2_ctx* c, int32_t* I, int32_t* J) {
  CHECK_CTX(c);
  REQUIRE(c->have_pattern, "call mb2_build_pattern first");
  if (c->nnz > 2147483647LL) {
    c->err = "nnz does not fit a 32-bit int (the reference's own limit); use mb2_get_pattern_i64";
    return MB2_ERR_UNSUPPORTED;
  }
  if (I) {
    int32_t* d_tmp = nullptr;
    MB2_CUDA(cudaMalloc(&d_tmp, sizeof(int32_t) * (c->NV + 1)));
    k_i64_to_i32<<<(unsigned)((c->NV + 256) / 256), 256, 0, c->stream>>>(c->NV + 1, c->d_I, d_tmp);
    MB2_CUDA(cudaMemcpyAsync(I, d_tmp, sizeof(int32_t) * (c->NV + 1), cudaMemcpyDeviceToHost, c->stream));
    MB2_CUDA(cudaStreamSynchronize(c->stream));
    cudaFree(d_tmp);
  }
  if (J) MB2_CUDA(cudaMemcpyAsync(J, c->d_J, sizeof(int32_t) * c->nnz, cudaMemcpyDeviceToHost, c->stream));
  MB2_CUDA(cudaStreamSynchronize(c->stream));
  return MB2_OK;
}

int mb2_get_A_ids_i32(mb2_ctx* c, int32_t* A_ids) {
  CHECK_CTX(c);
  REQUIRE(c->have_pattern && A_ids, "call mb2_build_pattern first");
  REQUIRE(c->nnz <= 2147483647LL, "nnz does not fit int");
  const int64_t n = c->NE * c->tdof * c->tdof;
  int32_t* d_tmp = nullptr;
  MB2_CUDA(cudaMalloc(&d_tmp, sizeof(int32_t) * n));
  int rc = mb2::expand_A_ids(c, d_tmp);
  if (rc) return rc;
  MB2_CUDA(cudaMemcpyAsync(A_ids, d_tmp, sizeof(int32_t) * n, cudaMemcpyDeviceToHost, c->stream));
  MB2_CUDA(cudaStreamSynchronize(c->stream));
  cudaFree(d_tmp);
  return MB2_OK;
}

// ---- resident state ----
int mb2_set_u(mb2_ctx* c, const double* u) {
  CHECK_CTX(c);
  REQUIRE(c->d_u && u, "no problem set");
  MB2_CUDA(cudaMemcpyAsync(c->d_u, u, sizeof(double) * c->NV, cudaMemcpyHostToDevice, c->stream));
  return MB2_OK;
}
int mb2_set_r(mb2_ctx* c, const double* r) {
  CHECK_CTX(c);
  REQUIRE(c->d_r, "no problem set");
  if (r) MB2_CUDA(cudaMemcpyAsync(c->d_r, r, sizeof(double) * c->NV, cudaMemcpyHostToDevice, c->stream));
  else MB2_CUDA(cudaMemsetAsync(c->d_r, 0, sizeof(double) * c->NV, c->stream));
  return MB2_OK;
}
int mb2_set_A(mb2_ctx* c, const double* A) {
  CHECK_CTX(c);
  REQUIRE(c->d_A, "call mb2_build_pattern first");
  if (A) MB2_CUDA(cudaMemcpyAsync(c->d_A, A, sizeof(double) * c->nnz, cudaMemcpyHostToDevice, c->stream));
  else MB2_CUDA(cudaMemsetAsync(c->d_A, 0, sizeof(double) * c->nnz, c->stream));
  return MB2_OK;
}
int mb2_get_r(mb2_ctx* c, double* r) {
  CHECK_CTX(c);
  REQUIRE(c->d_r && r, "no problem set");
  MB2_CUDA(cudaMemcpyAsync(r, c->d_r, sizeof(double) * c->NV, cudaMemcpyDeviceToHost, c->stream));
  MB2_CUDA(cudaStreamSynchronize(c->stream));
  return MB2_OK;
}
int mb2_get_A(mb2_ctx* c, double* A) {
  CHECK_CTX(c);
  REQUIRE(c->d_A && A, "call mb2_build_pattern first");
  MB2_CUDA(cudaMemcpyAsync(A, c->d_A, sizeof(double) * c->nnz, cudaMemcpyDeviceToHost, c->stream));
  MB2_CUDA(cudaStreamSynchronize(c->stream));
  return MB2_OK;
}
int mb2_device_pointers(mb2_ctx* c, void** u, void** r, void** A) {
  CHECK_CTX(c);
  if (u) *u = c->d_u;
  if (r) *r = c->d_r;
  if (A) *A = c->d_A;
  return MB2_OK;
}

static int ready(mb2_ctx* c, bool grad) {
  REQUIRE(c->d_dNdX && c->d_wdetJ, "quadrature tables missing (mb2_upload_quad_tables / mb2_generate_patch_tables)");
  REQUIRE(c->mat_kind >= 0, "material not set");
  REQUIRE(c->have_pattern, "call mb2_build_pattern first");
  (void)grad;
  return MB2_OK;
}

int mb2_assemble_residual_device(mb2_ctx* c) {
  CHECK_CTX(c);
  int rc = ready(c, false);
  if (rc) return rc;
  MB2_CUDA(cudaSetDevice(c->device));
  return mb2::launch_residual(c);
}
int mb2_assemble_residual_and_grad_device(mb2_ctx* c, double gf) {
  CHECK_CTX(c);
  int rc = ready(c, true);
  if (rc) return rc;
  MB2_CUDA(cudaSetDevice(c->device));
  return mb2::launch_residual_and_grad(c, gf, nullptr, nullptr);
}

int mb2_add_domain_residual(mb2_ctx* c, const double* u, double* r) {
  CHECK_CTX(c);
  REQUIRE(u && r, "null vector");
  int rc;
  if ((rc = mb2_set_u(c, u))) return rc;
  if ((rc = mb2_set_r(c, r))) return rc;
  if ((rc = mb2_assemble_residual_device(c))) return rc;
  if ((rc = mb2_get_r(c, r))) return rc;
  return c->ns ? mb2_check_numeric(c) : MB2_OK;
}
int mb2_add_domain_residual_and_grad(mb2_ctx* c, const double* u, double gf, double* r, double* A) {
  CHECK_CTX(c);
  REQUIRE(u && r && A, "null vector");
  int rc = ready(c, true);
  if (rc) return rc;
  MB2_CUDA(cudaSetDevice(c->device));
  if ((rc = mb2_set_u(c, u))) return rc;
  if ((rc = mb2_set_r(c, r))) return rc;
  // A travels chunk by chunk inside the pipeline: rows are uploaded just before the chunk that completes them
  // is gathered and downloaded right after (copy engines, overlapped with the element kernels).
  if ((rc = mb2::launch_residual_and_grad(c, gf, A, A))) return rc;
  if ((rc = mb2_get_r(c, r))) return rc;
  return c->ns ? mb2_check_numeric(c) : MB2_OK;
}

int mb2_host_register(mb2_ctx* c, void* ptr, size_t bytes) {
  CHECK_CTX(c);
  REQUIRE(ptr && bytes, "null");
  MB2_CUDA(cudaHostRegister(ptr, bytes, cudaHostRegisterDefault));
  return MB2_OK;
}
int mb2_host_unregister(mb2_ctx* c, void* ptr) {
  CHECK_CTX(c);
  MB2_CUDA(cudaHostUnregister(ptr));
  return MB2_OK;
}

int mb2_last_timing_ms(mb2_ctx* c, float* out3) {
  CHECK_CTX(c);
  REQUIRE(out3, "null");
  return mb2::collect_timing(c, out3);
}
int mb2_plan_info(mb2_ctx* c, int64_t* chunk_elems, int* nchunks) {
  CHECK_CTX(c);
  REQUIRE(c->have_pattern, "call mb2_build_pattern first");
  return mb2::plan_info(c, chunk_elems, nchunks);
}
int mb2_last_launch_count(mb2_ctx* c) { return c ? c->last_launches : 0; }
int mb2_sync(mb2_ctx* c) {
  CHECK_CTX(c);
  MB2_CUDA(cudaStreamSynchronize(c->stream));
  return MB2_OK;
}
int mb2_measure_fp64_peak(mb2_ctx* c, int fma, double* tflops) {
  CHECK_CTX(c);
  REQUIRE(tflops, "null");
  MB2_CUDA(cudaSetDevice(c->device));
  if (fma == 2) return mb2::launch_dmma_peak(c, tflops);
  return mb2::launch_fp64_peak(c, fma, tflops);
}

int mb2_set_option(mb2_ctx* c, const char* name, int value) {
  CHECK_CTX(c);
  REQUIRE(name, "null option name");
  if (!std::strcmp(name, "elem_variant")) { c->elem_variant = value; return MB2_OK; }
  if (!std::strcmp(name, "park_acc")) { c->park_acc = value; return MB2_OK; }
  if (!std::strcmp(name, "res_variant")) { c->res_variant = value; return MB2_OK; }
  if (!std::strcmp(name, "mg_exchange_late")) { c->mg_late = value; return MB2_OK; }
  if (!std::strcmp(name, "tangent_mode")) {
    if (value != MB2_TANGENT_FD_PARITY && value != MB2_TANGENT_ANALYTIC) { c->err = "unknown tangent mode"; return MB2_ERR_ARG; }
    c->tangent_mode = value;
    return MB2_OK;
  }
  if (!std::strcmp(name, "fast_gemm")) { c->fast_gemm = value; return MB2_OK; }
  if (!std::strcmp(name, "j2_counters")) {  // measurement only: atomics in the return mapping, never on in timed runs
    if (value && !c->d_cnt) {
      MB2_CUDA(cudaMalloc(&c->d_cnt, 2 * sizeof(unsigned long long)));
      MB2_CUDA(cudaMemset(c->d_cnt, 0, 2 * sizeof(unsigned long long)));
    }
    c->mp.cnt = value ? c->d_cnt : nullptr;
    return MB2_OK;
  }
  if (!std::strcmp(name, "chunk_elems")) { c->chunk_elems = value; mb2::free_plan(c); return MB2_OK; }
  if (!std::strcmp(name, "chunk_mb")) { c->chunk_mb = value > 0 ? value : 2048; mb2::free_plan(c); return MB2_OK; }
  c->err = std::string("unknown option ") + name;
  return MB2_ERR_ARG;
}

int mb2_set_nodes(mb2_ctx* c, const double* x_ref, const double* weights) {
  CHECK_CTX(c);
  REQUIRE(c->NE > 0, "call mb2_set_connectivity first");
  REQUIRE(x_ref, "null x_ref");
  MB2_CUDA(cudaSetDevice(c->device));
  return mb2::set_nodes(c, x_ref, weights);
}

int mb2_assemble_mass_device(mb2_ctx* c, int p, const int32_t* ncp, const double* const* knots, const int32_t* el_ijk,
                             int quad_order, double density) {
  CHECK_CTX(c);
  REQUIRE(c->NE > 0, "call mb2_set_connectivity first");
  REQUIRE(ncp && knots && el_ijk, "null argument");
  MB2_CUDA(cudaSetDevice(c->device));
  return mb2::assemble_mass(c, p, ncp, knots, el_ijk, quad_order, density);
}

int mb2_assemble_traction_device(mb2_ctx* c, int64_t NF, int ndf, int nqf, const int32_t* face_vdofs, const double* N,
                                 const double* w, const double* detW, const double* traction) {
  CHECK_CTX(c);
  REQUIRE(c->NE > 0, "call mb2_set_connectivity first");
  REQUIRE(face_vdofs && N && w && detW && traction, "null argument");
  MB2_CUDA(cudaSetDevice(c->device));
  return mb2::assemble_traction(c, NF, ndf, nqf, face_vdofs, N, w, detW, traction);
}

int mb2_assemble_viscosity_device(mb2_ctx* c, int p, const int32_t* ncp, const double* const* knots, const int32_t* el_ijk,
                                  int quad_order, double viscosity) {
  CHECK_CTX(c);
  REQUIRE(c->NE > 0, "call mb2_set_connectivity first");
  REQUIRE(ncp && knots && el_ijk, "null argument");
  MB2_CUDA(cudaSetDevice(c->device));
  return mb2::assemble_viscosity(c, p, ncp, knots, el_ijk, quad_order, viscosity);
}

int mb2_assemble_body_force_device(mb2_ctx* c, int p, const int32_t* ncp, const double* const* knots, const int32_t* el_ijk,
                                   int quad_order, const double* force) {
  CHECK_CTX(c);
  REQUIRE(c->NE > 0, "call mb2_set_connectivity first");
  REQUIRE(ncp && knots && el_ijk && force, "null argument");
  MB2_CUDA(cudaSetDevice(c->device));
  return mb2::assemble_body_force(c, p, ncp, knots, el_ijk, quad_order, force);
}

int mb2_generate_patch_tables(mb2_ctx* c, int64_t elem_begin, int64_t elem_count, int p, const int32_t* ncp,
                              const double* const* knots, const int32_t* el_ijk, int quad_order, int want_N) {
  CHECK_CTX(c);
  REQUIRE(c->NE > 0, "call mb2_set_connectivity first");
  REQUIRE(ncp && knots && el_ijk, "null argument");
  MB2_CUDA(cudaSetDevice(c->device));
  return mb2::generate_patch_tables(c, elem_begin, elem_count, p, ncp, knots, el_ijk, quad_order, want_N != 0);
}

int mb2_download_shape_values(mb2_ctx* c, double* N) {
  CHECK_CTX(c);
  REQUIRE(c->d_N && N, "no shape values on device (mb2_generate_patch_tables with want_N)");
  MB2_CUDA(cudaMemcpyAsync(N, c->d_N, sizeof(double) * c->NE * c->nq * c->nd, cudaMemcpyDeviceToHost, c->stream));
  MB2_CUDA(cudaStreamSynchronize(c->stream));
  return MB2_OK;
}

}  // extern "C"
