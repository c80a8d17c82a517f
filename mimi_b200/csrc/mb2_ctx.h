// Internal context of libmimi_b200 (not part of the ABI).
#pragma once
#include <cuda_runtime.h>

#include <cstdint>
#include <string>
#include <vector>

#include "../../include/mimi_b200.h"
#include "mb2_math.cuh"

// Multi-GPU exchange state: which element matrices leave to / arrive from which peer (mb2_multigpu.cu).
struct MgPeer {
  int peer = -1;
  // send: local element indices (ascending, < n_compute) whose K_e / R_e go to `peer`
  std::vector<int64_t> send_elems;
  int64_t* d_send_elems = nullptr;
  double* d_send = nullptr;       // [n_send][nd*KP] K_e blocks, then [n_send][tdof] residuals
  // recv: contiguous block of ghost elements [recv_first, recv_first + n_recv) filled by `peer`
  int64_t recv_first = 0, n_recv = 0;
};
struct MgState {
  std::vector<MgPeer> peers;
  int rank = 0, world = 1;
  void* nccl_lib = nullptr;   // dlopen handle
  void* comm = nullptr;       // ncclComm_t
  cudaStream_t s_comm = nullptr;
  cudaEvent_t ev_pack = nullptr, ev_comm = nullptr, ev_sent = nullptr;
  bool manual = false;        // exchange driven by the host through mb2_mg_export/import (tests, foreign transports)
};

struct mb2_ctx {
  int device = 0;
  cudaStream_t stream = nullptr;
  bool own_stream = false;
  std::string err;
  int sm_count = 148;

  // problem shape
  int dim = 0, nd = 0, nq = 0, tdof = 0;
  int64_t NE = 0, NV = 0, Nn = 0, nnz = 0;

  // material
  int chunk_mb = 2048;      // K_e buffer budget per chunk (MiB), two buffers
  int64_t chunk_elems = 0;  // explicit chunk size (0 = from chunk_mb)
  void* plan = nullptr;     // chunk pipeline state (mb2_assembly.cu)
  bool last_grad = false;
  int elem_variant = 0;  // 0 = default kernels, 1 = force the one-thread-per-dof FD kernel (A/B tests)
  int park_acc = 1;      // FD kernels park their accumulators in tensor memory during phase 1 (p=2: plastic materials, p=3: all); 0 = never (A/B)
  int mg_late = 0;       // NCCL exchange enqueued behind the rank's last element kernel instead of behind its pack (see mb2_set_option)
  int res_variant = 0;   // residual-only kernel: 0 = element_res2_kernel, 1 = the first form (A/B)
  int tangent_mode = 0;  // MB2_TANGENT_FD_PARITY | MB2_TANGENT_ANALYTIC (mb2_fast.cu)
  int fast_gemm = 0;     // analytic mode, p = 2 hexahedra: 0 = FP64 tensor cores (DMMA), 1 = CUDA-core DFMA tiles
  int mat_kind = -1;
  mb2::MatP mp;
  int ns = 0;                // history doubles per quadrature point (0 = stateless material)
  double* d_state = nullptr; // [NE][nq][ns]

  // partition (multi-GPU): elements [0, n_compute) are computed here, [n_compute, NE) are ghosts whose
  // element matrices arrive from other ranks; only rows of owned nodes are assembled
  int64_t n_compute = 0;
  uint8_t* d_owned = nullptr;  // Nn, nullptr = every node
  MgState* mg = nullptr;       // exchange state (mb2_multigpu.cu)
  void* contact = nullptr;     // mortar contact state (mb2_contact.cu)
  void* newton = nullptr;      // Newton-iteration sweeps: mass / viscosity / rhs / essential dofs (mb2_newton.cu)
  void* mass_tab = nullptr;    // non-null only inside mb2_assemble_mass_device: the element stage writes M_e (mb2_tables.cu)

  // connectivity
  int32_t* d_vdofs = nullptr;   // NE x tdof (reference layout)
  int32_t* d_enodes = nullptr;  // NE x nd   node ids

  // tables
  double* d_dNdX = nullptr;  // NE x nq x (nd x dim col-major)
  double* d_wdetJ = nullptr; // NE x nq   (integration_weight * det_dX_dxi)
  double* d_N = nullptr;     // NE x nq x nd shape values (optional; contact / tests)
  double* d_xref = nullptr;  // NV  reference node coordinates, byVDIM
  double* d_nodew = nullptr; // Nn  NURBS weights (nullptr = all 1)

  // node -> element adjacency (ascending element id) and pattern
  int64_t* d_n2e_ptr = nullptr;  // Nn+1
  int32_t* d_n2e_elem = nullptr; // sum
  uint8_t* d_n2e_loc = nullptr;  // local node index inside that element
  int32_t* d_nnbr = nullptr;     // Nn   neighbours per node
  int64_t* d_nodeptr = nullptr;  // Nn+1 prefix sum of nnbr
  int32_t* d_nbr = nullptr;      // nodeptr[Nn] sorted neighbour nodes
  uint16_t* d_rank = nullptr;    // NE x nd x nd  rank of node b in nbr list of node a
  int64_t* d_I = nullptr;        // NV+1
  int32_t* d_J = nullptr;        // nnz
  bool have_pattern = false;

  // resident state
  double* d_u = nullptr;  // NV
  double* d_r = nullptr;  // NV
  double* d_A = nullptr;  // nnz

  // per-element results
  double* d_Re = nullptr;  // NE x tdof
  double* d_KeT = nullptr; // NE x [row node][perturbed dof][row component]  = K_e(node + comp*nd, dof)
  int* d_flag = nullptr;   // device failure flag
  unsigned long long* d_cnt = nullptr;  // J2 measurement counters (2), allocated by mb2_set_option("j2_counters", 1)

  cudaEvent_t ev[4] = {nullptr, nullptr, nullptr, nullptr};
  float last_ms[3] = {0, 0, 0};
  int last_launches = 0;
};

namespace mb2 {
// arguments of every element kernel (mb2_assembly.cu, mb2_fast.cu)
struct ElemArgs {
  int64_t e0, e1;  // element range of this launch; K_e is stored relative to e0
  const int32_t* __restrict__ vdofs;
  const double* __restrict__ dNdX;
  const double* __restrict__ wdetJ;
  const double* __restrict__ u;
  double* __restrict__ Re;
  double* __restrict__ KeT;
  MatP mp;
  const double* __restrict__ state;  // [NE][nq][NS] history variables (nullptr for stateless materials)
};
// mb2_fast.cu: MB2_TANGENT_ANALYTIC element kernels (closed-form tangent, FMA allowed), FP64 tensor-core peak
int launch_element_analytic(mb2_ctx* c, const ElemArgs& a, cudaStream_t st);
int launch_dmma_peak(mb2_ctx* c, double* tflops);
// mb2_pattern.cu
int build_pattern(mb2_ctx* c);
int expand_A_ids(mb2_ctx* c, int32_t* d_out);
// mb2_api.cu
int consume_flag(mb2_ctx* c);  // reads + clears the device failure flag; MB2_ERR_NUMERIC with a per-code message
// mb2_assembly.cu
int launch_residual(mb2_ctx* c);
int launch_post_time_advance(mb2_ctx* c);
int launch_residual_and_grad(mb2_ctx* c, double gf, const double* A_in, double* A_out);
void free_plan(mb2_ctx* c);
int collect_timing(mb2_ctx* c, float* out3);
int plan_info(mb2_ctx* c, int64_t* chunk_elems, int* nchunks);
int assemble_begin(mb2_ctx* c, double gf, bool grad, const double* A_in, double* A_out);
int assemble_finish(mb2_ctx* c, double gf, bool grad, double* A_out);
double* ghost_ke_buffer(mb2_ctx* c);  // [NE - n_compute][nd][KP]
// mb2_contact.cu
void free_contact(mb2_ctx* c);
// mb2_newton.cu
void free_newton(mb2_ctx* c);
// mb2_multigpu.cu
void free_mg(mb2_ctx* c);
int mg_pack_chunk(mb2_ctx* c, int64_t e0, int64_t e1, const double* d_KeT_chunk, bool grad, cudaStream_t st);
int mg_exchange(mb2_ctx* c, bool grad, cudaStream_t after, cudaStream_t before);
int launch_fp64_peak(mb2_ctx* c, int fma, double* tflops);
// mb2_tables.cu
int set_nodes(mb2_ctx* c, const double* x_ref, const double* weights);
int generate_patch_tables(mb2_ctx* c, int64_t e0, int64_t ne, int p, const int32_t* ncp, const double* const* knots,
                          const int32_t* el_ijk, int quad_order, bool want_N);
int assemble_mass(mb2_ctx* c, int p, const int32_t* ncp, const double* const* knots, const int32_t* el_ijk, int quad_order,
                  double rho);
int launch_mass_element(mb2_ctx* c, int64_t e0, int64_t e1, double* KeT, double* Re, cudaStream_t st);
int assemble_body_force(mb2_ctx* c, int p, const int32_t* ncp, const double* const* knots, const int32_t* el_ijk,
                        int quad_order, const double* f);
int launch_lform_element(mb2_ctx* c, int64_t e0, int64_t e1, double* Re, cudaStream_t st);
int assemble_viscosity(mb2_ctx* c, int p, const int32_t* ncp, const double* const* knots, const int32_t* el_ijk,
                       int quad_order, double coef);
int assemble_traction(mb2_ctx* c, int64_t NF, int ndf, int nqf, const int32_t* face_vdofs, const double* N, const double* w,
                      const double* detW, const double* traction);
int form_mode(const mb2_ctx* c);  // 0 mass, 1 body force, 2 viscosity (valid while mass_tab is set)
}  // namespace mb2

// Device temporaries of a set-up routine: freed on scope exit, also on the early returns of MB2_CUDA.
struct DevTmp {
  std::vector<void**> slots;
  template <class T>
  void own(T** p) { slots.push_back(reinterpret_cast<void**>(p)); }
  ~DevTmp() {
    for (void** s : slots) {
      if (*s) cudaFree(*s);
      *s = nullptr;
    }
  }
};

#define MB2_CUDA(call)                                                                      \
  do {                                                                                      \
    cudaError_t e_ = (call);                                                                \
    if (e_ != cudaSuccess) {                                                                \
      c->err = std::string(#call) + ": " + cudaGetErrorString(e_) + " (" + __FILE__ + ":" + \
               std::to_string(__LINE__) + ")";                                              \
      return MB2_ERR_CUDA;                                                                  \
    }                                                                                       \
  } while (0)
