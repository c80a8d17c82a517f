// Multi-GPU exchange of raw element matrices.
//
// The reference adds element contributions in ascending element id (nonlinear_solid.cpp:126-143). To keep that
// order when elements are spread over ranks, a rank does not send partial row sums: it sends the element
// matrices K_e / residuals R_e of the elements that touch rows owned by a lower rank, and the owner appends
// them after its own elements (all of which have smaller ids) before its ordered gather.
//
// Transport: NCCL point-to-point (ncclSend / ncclRecv in one group) on a high-priority stream, loaded with
// dlopen so that the library itself has no link-time dependency on NCCL; or "manual" mode, in which the host
// moves the packed buffers itself (single-GPU tests emulating several ranks, foreign transports).
#include <dlfcn.h>

#include <cstring>

#include "mb2_ctx.h"

namespace mb2 {

// minimal NCCL surface (nccl.h, NCCL 2.x ABI)
typedef struct ncclComm* ncclComm_t;
typedef struct { char internal[128]; } ncclUniqueId;
enum { ncclSuccess = 0 };
enum { ncclDouble = 8 };  // ncclFloat64
struct NcclApi {
  int (*GetUniqueId)(ncclUniqueId*) = nullptr;
  int (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
  int (*CommDestroy)(ncclComm_t) = nullptr;
  int (*Send)(const void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
  int (*Recv)(void*, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
  int (*GroupStart)() = nullptr;
  int (*GroupEnd)() = nullptr;
  const char* (*GetErrorString)(int) = nullptr;
  void* lib = nullptr;
};
static NcclApi g_nccl;

static int load_nccl(std::string& err) {
  if (g_nccl.lib) return MB2_OK;
  // an already loaded NCCL (e.g. the one PyTorch ships) wins: same SONAME
  const char* names[] = {"libnccl.so.2", "libnccl.so"};
  for (const char* n : names)
    if ((g_nccl.lib = dlopen(n, RTLD_NOW | RTLD_GLOBAL))) break;
  if (!g_nccl.lib) { err = std::string("cannot load libnccl: ") + dlerror(); return MB2_ERR_UNSUPPORTED; }
#define MB2_SYM(field, name)                                                       \
  g_nccl.field = reinterpret_cast<decltype(g_nccl.field)>(dlsym(g_nccl.lib, name)); \
  if (!g_nccl.field) { err = std::string("libnccl lacks ") + name; return MB2_ERR_UNSUPPORTED; }
  MB2_SYM(GetUniqueId, "ncclGetUniqueId")
  MB2_SYM(CommInitRank, "ncclCommInitRank")
  MB2_SYM(CommDestroy, "ncclCommDestroy")
  MB2_SYM(Send, "ncclSend")
  MB2_SYM(Recv, "ncclRecv")
  MB2_SYM(GroupStart, "ncclGroupStart")
  MB2_SYM(GroupEnd, "ncclGroupEnd")
  MB2_SYM(GetErrorString, "ncclGetErrorString")
#undef MB2_SYM
  return MB2_OK;
}

#define MB2_NCCL(call)                                                                   \
  do {                                                                                   \
    int r_ = (call);                                                                     \
    if (r_ != ncclSuccess) {                                                             \
      c->err = std::string(#call) + ": " + g_nccl.GetErrorString(r_);                    \
      return MB2_ERR_CUDA;                                                               \
    }                                                                                    \
  } while (0)

static inline size_t ke_doubles(const mb2_ctx* c) { return (size_t)c->nd * ((size_t)(c->tdof * c->dim + 3) / 4 * 4); }

static MgState* state(mb2_ctx* c) {
  if (!c->mg) {
    c->mg = new MgState();
    int lo = 0, hi = 0;
    cudaDeviceGetStreamPriorityRange(&lo, &hi);
    cudaStreamCreateWithPriority(&c->mg->s_comm, cudaStreamNonBlocking, hi);
    cudaEventCreateWithFlags(&c->mg->ev_pack, cudaEventDisableTiming);
    cudaEventCreateWithFlags(&c->mg->ev_comm, cudaEventDisableTiming);
    cudaEventCreateWithFlags(&c->mg->ev_sent, cudaEventDisableTiming);
  }
  return c->mg;
}

static MgPeer& peer_of(MgState* mg, int peer) {
  for (auto& p : mg->peers)
    if (p.peer == peer) return p;
  mg->peers.emplace_back();
  mg->peers.back().peer = peer;
  return mg->peers.back();
}

void free_mg(mb2_ctx* c) {
  MgState* mg = c->mg;
  if (!mg) return;
  for (auto& p : mg->peers) {
    cudaFree(p.d_send_elems);
    cudaFree(p.d_send);
  }
  if (mg->comm && g_nccl.CommDestroy) g_nccl.CommDestroy((ncclComm_t)mg->comm);
  if (mg->s_comm) cudaStreamDestroy(mg->s_comm);
  if (mg->ev_pack) cudaEventDestroy(mg->ev_pack);
  if (mg->ev_comm) cudaEventDestroy(mg->ev_comm);
  if (mg->ev_sent) cudaEventDestroy(mg->ev_sent);
  delete mg;
  c->mg = nullptr;
}

// One CTA per outgoing element inside [e0, e1): copies its K_e block (grad) and its residual row.
__global__ void k_pack(const int64_t* __restrict__ elems, int64_t n, int64_t n_total, int64_t e0, int64_t e1, int ked,
                       int tdof, const double* __restrict__ KeT_chunk, const double* __restrict__ Re,
                       double* __restrict__ out) {
  const int64_t x = blockIdx.x;
  if (x >= n) return;
  const int64_t e = elems[x];
  if (e < e0 || e >= e1) return;
  if (KeT_chunk) {
    const double2* src = reinterpret_cast<const double2*>(KeT_chunk + (e - e0) * (int64_t)ked);
    double2* dst = reinterpret_cast<double2*>(out + x * (int64_t)ked);
    for (int i = threadIdx.x; i < ked / 2; i += blockDim.x) dst[i] = src[i];
  }
  double* r = out + n_total * (int64_t)ked + x * tdof;
  for (int i = threadIdx.x; i < tdof; i += blockDim.x) r[i] = Re[e * tdof + i];
}

int mg_pack_chunk(mb2_ctx* c, int64_t e0, int64_t e1, const double* d_KeT_chunk, bool grad, cudaStream_t st) {
  MgState* mg = c->mg;
  const int ked = (int)ke_doubles(c);
  for (auto& p : mg->peers) {
    const int64_t n = (int64_t)p.send_elems.size();
    if (!n || p.send_elems.front() >= e1 || p.send_elems.back() < e0) continue;
    k_pack<<<(unsigned)n, 256, 0, st>>>(p.d_send_elems, n, n, e0, e1, ked, c->tdof, grad ? d_KeT_chunk : nullptr, c->d_Re,
                                        p.d_send);
    MB2_CUDA(cudaGetLastError());
  }
  return MB2_OK;
}

// Enqueues the point-to-point exchange: waits for the packs, then one NCCL group with every send and receive.
int mg_exchange(mb2_ctx* c, bool grad, cudaStream_t, cudaStream_t) {
  MgState* mg = c->mg;
  if (!mg->comm) { c->err = "partitioned assembly without a communicator: call mb2_mg_init_nccl (or use the manual exchange)"; return MB2_ERR_ARG; }
  const size_t ked = ke_doubles(c);
  MB2_CUDA(cudaStreamWaitEvent(mg->s_comm, mg->ev_pack, 0));
  double* ghost = ghost_ke_buffer(c);
  MB2_NCCL(g_nccl.GroupStart());
  for (auto& p : mg->peers) {
    const size_t ns = p.send_elems.size();
    if (ns) {
      if (grad) MB2_NCCL(g_nccl.Send(p.d_send, ns * ked, ncclDouble, p.peer, (ncclComm_t)mg->comm, mg->s_comm));
      MB2_NCCL(g_nccl.Send(p.d_send + ns * ked, ns * c->tdof, ncclDouble, p.peer, (ncclComm_t)mg->comm, mg->s_comm));
    }
    if (p.n_recv) {
      if (grad)
        MB2_NCCL(g_nccl.Recv(ghost + (size_t)(p.recv_first - c->n_compute) * ked, (size_t)p.n_recv * ked, ncclDouble, p.peer,
                             (ncclComm_t)mg->comm, mg->s_comm));
      MB2_NCCL(g_nccl.Recv(c->d_Re + (size_t)p.recv_first * c->tdof, (size_t)p.n_recv * c->tdof, ncclDouble, p.peer,
                           (ncclComm_t)mg->comm, mg->s_comm));
    }
  }
  MB2_NCCL(g_nccl.GroupEnd());
  MB2_CUDA(cudaEventRecord(mg->ev_comm, mg->s_comm));
  return MB2_OK;
}

}  // namespace mb2

// ---------------------------------------------------------------------------------------------
// C ABI
// ---------------------------------------------------------------------------------------------
#define MG_CHECK_CTX(c) \
  if (!(c)) return MB2_ERR_ARG
#define MG_REQUIRE(cond, msg) \
  if (!(cond)) {              \
    c->err = (msg);           \
    return MB2_ERR_ARG;       \
  }
using namespace mb2;

extern "C" {

int mb2_set_partition(mb2_ctx* c, int64_t n_compute, const uint8_t* owned_nodes) {
  MG_CHECK_CTX(c);
  MG_REQUIRE(c->NE > 0, "call mb2_set_connectivity first");
  MG_REQUIRE(!c->have_pattern, "mb2_set_partition must precede mb2_build_pattern");
  MG_REQUIRE(n_compute > 0 && n_compute <= c->NE && owned_nodes, "bad partition arguments");
  MB2_CUDA(cudaSetDevice(c->device));
  c->n_compute = n_compute;
  if (!c->d_owned) MB2_CUDA(cudaMalloc(&c->d_owned, c->Nn));
  MB2_CUDA(cudaMemcpy(c->d_owned, owned_nodes, c->Nn, cudaMemcpyHostToDevice));
  state(c);
  return MB2_OK;
}

int mb2_mg_set_send(mb2_ctx* c, int peer, const int64_t* local_elems, int64_t n) {
  MG_CHECK_CTX(c);
  MG_REQUIRE(c->mg, "call mb2_set_partition first");
  MG_REQUIRE(local_elems && n > 0, "empty send list");
  for (int64_t i = 0; i < n; i++)
    MG_REQUIRE(local_elems[i] >= 0 && local_elems[i] < c->n_compute && (i == 0 || local_elems[i] > local_elems[i - 1]),
               "send elements must be ascending computed elements");
  MB2_CUDA(cudaSetDevice(c->device));
  MgPeer& p = peer_of(c->mg, peer);
  p.send_elems.assign(local_elems, local_elems + n);
  cudaFree(p.d_send_elems);
  cudaFree(p.d_send);
  MB2_CUDA(cudaMalloc(&p.d_send_elems, sizeof(int64_t) * n));
  MB2_CUDA(cudaMemcpy(p.d_send_elems, local_elems, sizeof(int64_t) * n, cudaMemcpyHostToDevice));
  MB2_CUDA(cudaMalloc(&p.d_send, sizeof(double) * (size_t)n * (ke_doubles(c) + c->tdof)));
  return MB2_OK;
}

int mb2_mg_set_recv(mb2_ctx* c, int peer, int64_t first_local_elem, int64_t n) {
  MG_CHECK_CTX(c);
  MG_REQUIRE(c->mg, "call mb2_set_partition first");
  MG_REQUIRE(n > 0 && first_local_elem >= c->n_compute && first_local_elem + n <= c->NE, "ghost block out of range");
  MgPeer& p = peer_of(c->mg, peer);
  p.recv_first = first_local_elem;
  p.n_recv = n;
  return MB2_OK;
}

int mb2_mg_unique_id(void* out128) {
  std::string err;
  if (!out128 || load_nccl(err)) return MB2_ERR_UNSUPPORTED;
  ncclUniqueId id;
  if (g_nccl.GetUniqueId(&id) != ncclSuccess) return MB2_ERR_CUDA;
  std::memcpy(out128, &id, sizeof(id));
  return MB2_OK;
}

int mb2_mg_init_nccl(mb2_ctx* c, int rank, int world, const void* unique_id128) {
  MG_CHECK_CTX(c);
  MG_REQUIRE(c->mg, "call mb2_set_partition first");
  MG_REQUIRE(unique_id128 && rank >= 0 && rank < world, "bad communicator arguments");
  int rc = load_nccl(c->err);
  if (rc) return rc;
  MB2_CUDA(cudaSetDevice(c->device));
  ncclUniqueId id;
  std::memcpy(&id, unique_id128, sizeof(id));
  ncclComm_t comm = nullptr;
  MB2_NCCL(g_nccl.CommInitRank(&comm, world, id, rank));
  c->mg->comm = comm;
  c->mg->rank = rank;
  c->mg->world = world;
  c->mg->manual = false;
  return MB2_OK;
}

int mb2_mg_set_manual(mb2_ctx* c, int on) {
  MG_CHECK_CTX(c);
  MG_REQUIRE(c->mg, "call mb2_set_partition first");
  c->mg->manual = on != 0;
  return MB2_OK;
}

static int mg_ready(mb2_ctx* c) {
  MG_REQUIRE(c->mg && c->mg->manual, "manual exchange is not enabled (mb2_mg_set_manual)");
  MG_REQUIRE(c->have_pattern && c->d_dNdX && c->mat_kind >= 0, "problem not set up");
  return MB2_OK;
}

/* manual exchange, stage 1: own elements (grad != 0: residual + tangent; 0: residual only) */
int mb2_mg_begin(mb2_ctx* c, double grad_factor, int grad) {
  MG_CHECK_CTX(c);
  int rc = mg_ready(c);
  if (rc) return rc;
  MB2_CUDA(cudaSetDevice(c->device));
  return assemble_begin(c, grad_factor, grad != 0, nullptr, nullptr);
}

/* packed outgoing matrices for `peer`: Ke [n_send][nd*KP] (may be NULL), Re [n_send][tdof] */
int mb2_mg_export_send(mb2_ctx* c, int peer, double* Ke_host, double* Re_host) {
  MG_CHECK_CTX(c);
  int rc = mg_ready(c);
  if (rc) return rc;
  MgPeer& p = peer_of(c->mg, peer);
  const size_t ns = p.send_elems.size(), ked = ke_doubles(c);
  MG_REQUIRE(ns > 0, "nothing to send to this peer");
  MB2_CUDA(cudaEventSynchronize(c->mg->ev_pack));
  if (Ke_host) MB2_CUDA(cudaMemcpy(Ke_host, p.d_send, sizeof(double) * ns * ked, cudaMemcpyDeviceToHost));
  if (Re_host) MB2_CUDA(cudaMemcpy(Re_host, p.d_send + ns * ked, sizeof(double) * ns * c->tdof, cudaMemcpyDeviceToHost));
  return MB2_OK;
}

int mb2_mg_import_recv(mb2_ctx* c, int peer, const double* Ke_host, const double* Re_host) {
  MG_CHECK_CTX(c);
  int rc = mg_ready(c);
  if (rc) return rc;
  MgPeer& p = peer_of(c->mg, peer);
  const size_t ked = ke_doubles(c);
  MG_REQUIRE(p.n_recv > 0, "nothing to receive from this peer");
  if (Ke_host) {
    double* ghost = ghost_ke_buffer(c);
    MG_REQUIRE(ghost, "no ghost buffer (call mb2_mg_begin first)");
    MB2_CUDA(cudaMemcpy(ghost + (size_t)(p.recv_first - c->n_compute) * ked, Ke_host, sizeof(double) * p.n_recv * ked,
                        cudaMemcpyHostToDevice));
  }
  if (Re_host)
    MB2_CUDA(cudaMemcpy(c->d_Re + (size_t)p.recv_first * c->tdof, Re_host, sizeof(double) * p.n_recv * c->tdof,
                        cudaMemcpyHostToDevice));
  return MB2_OK;
}

int mb2_mg_finish(mb2_ctx* c, double grad_factor, int grad) {
  MG_CHECK_CTX(c);
  int rc = mg_ready(c);
  if (rc) return rc;
  MB2_CUDA(cudaSetDevice(c->device));
  return assemble_finish(c, grad_factor, grad != 0, nullptr);
}

/* ---- one process, several GPUs: the contexts of all ranks are driven together and the packed element matrices travel by
 * cudaMemcpyPeerAsync (NVLink peer-to-peer when the devices allow it; the same device works too, which is how the
 * single-GPU test suite covers this path). ctxs[r] = the context of rank r, set up with mb2_plan_apply / mb2_set_partition.
 * grad != 0: r += R(u), A += gf K(u) on every rank's resident arrays; 0: residual only. */
int mb2_mg_group_init(mb2_ctx** ctxs, int n) {
  if (!ctxs || n < 1) return MB2_ERR_ARG;
  for (int a = 0; a < n; a++) {
    mb2_ctx* c = ctxs[a];
    MG_CHECK_CTX(c);
    MG_REQUIRE(n == 1 || c->mg, "call mb2_set_partition (mb2_plan_apply) on every context first");
    if (c->mg) { c->mg->manual = false; c->mg->rank = a; c->mg->world = n; }
    for (int b = 0; b < n; b++) {
      if (ctxs[b]->device == c->device) continue;
      int can = 0;
      MB2_CUDA(cudaDeviceCanAccessPeer(&can, c->device, ctxs[b]->device));
      if (!can) continue;                      // cudaMemcpyPeerAsync then stages through the host
      MB2_CUDA(cudaSetDevice(c->device));
      cudaError_t e = cudaDeviceEnablePeerAccess(ctxs[b]->device, 0);
      if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) MB2_CUDA(e);
      (void)cudaGetLastError();
    }
  }
  return MB2_OK;
}

int mb2_mg_group_assemble(mb2_ctx** ctxs, int n, double grad_factor, int grad) {
  if (!ctxs || n < 1) return MB2_ERR_ARG;
  const bool g = grad != 0;
  for (int a = 0; a < n; a++) {  // stage 1 on every rank: own elements, packs
    mb2_ctx* c = ctxs[a];
    MG_CHECK_CTX(c);
    MB2_CUDA(cudaSetDevice(c->device));
    int rc = assemble_begin(c, grad_factor, g, nullptr, nullptr);
    if (rc) return rc;
  }
  if (n > 1) {
    // sends, issued on the sender's communication stream after its packs
    for (int a = 0; a < n; a++) {
      mb2_ctx* c = ctxs[a];
      MgState* mg = c->mg;
      MB2_CUDA(cudaSetDevice(c->device));
      MB2_CUDA(cudaStreamWaitEvent(mg->s_comm, mg->ev_pack, 0));
      const size_t ked = ke_doubles(c);
      for (auto& p : mg->peers) {
        const size_t ns = p.send_elems.size();
        if (!ns) continue;
        if (p.peer < 0 || p.peer >= n) { c->err = "send peer outside the group"; return MB2_ERR_ARG; }
        mb2_ctx* d = ctxs[p.peer];
        const MgPeer* q = nullptr;
        for (auto& x : d->mg->peers)
          if (x.peer == a) q = &x;
        if (!q || (size_t)q->n_recv != ns) { c->err = "send / receive lists of the group do not match"; return MB2_ERR_ARG; }
        // the receiver's buffers are free once ITS packs of this step are queued (its previous ghost gather precedes them)
        MB2_CUDA(cudaStreamWaitEvent(mg->s_comm, d->mg->ev_pack, 0));
        if (g) {
          double* ghost = ghost_ke_buffer(d);
          if (!ghost) { c->err = "receiver has no ghost buffer"; return MB2_ERR_ARG; }
          MB2_CUDA(cudaMemcpyPeerAsync(ghost + (size_t)(q->recv_first - d->n_compute) * ked, d->device, p.d_send, c->device,
                                       sizeof(double) * ns * ked, mg->s_comm));
        }
        MB2_CUDA(cudaMemcpyPeerAsync(d->d_Re + (size_t)q->recv_first * d->tdof, d->device, p.d_send + ns * ked, c->device,
                                     sizeof(double) * ns * c->tdof, mg->s_comm));
      }
      MB2_CUDA(cudaEventRecord(mg->ev_sent, mg->s_comm));  // everything this rank sends has left
    }
    // every rank: wait for what it receives, then (ev_comm) both its sends and its receives are complete
    for (int a = 0; a < n; a++) {
      mb2_ctx* c = ctxs[a];
      MB2_CUDA(cudaSetDevice(c->device));
      for (auto& p : c->mg->peers)
        if (p.n_recv) MB2_CUDA(cudaStreamWaitEvent(c->mg->s_comm, ctxs[p.peer]->mg->ev_sent, 0));
      MB2_CUDA(cudaEventRecord(c->mg->ev_comm, c->mg->s_comm));
    }
  }
  for (int a = 0; a < n; a++) {  // stage 2: ghost gather, residual gather, join
    mb2_ctx* c = ctxs[a];
    MB2_CUDA(cudaSetDevice(c->device));
    int rc = assemble_finish(c, grad_factor, g, nullptr);
    if (rc) return rc;
  }
  return MB2_OK;
}

/* doubles per packed element matrix (nd * KP) */
int64_t mb2_mg_ke_doubles(mb2_ctx* c) { return c ? (int64_t)ke_doubles(c) : 0; }

}  // extern "C"
