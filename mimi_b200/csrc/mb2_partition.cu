// Element partition planner and the single-process multi-GPU exchange.
//
// mb2_plan_partition is the C++ statement of mimi_b200/partition.py::rank_problem (kept as the test oracle): contiguous,
// balanced element ranges per rank; a node is owned by the rank of its LOWEST element; elements of higher ranks that touch
// an owned node are the rank's ghosts (their raw K_e / R_e arrive and are appended after the rank's own elements, so the
// ordered gather still adds in ascending global element id = the reference's single-thread order,
// src/mimi/integrators/nonlinear_solid.cpp:126-143). Pure host code: usable without a GPU.
//
// mb2_mg_group_*: one process driving several GPUs (one context each) -- what a reference maintainer reaches from
// PyNonlinearSolid::Setup without MPI / torchrun: the packed element matrices travel by cudaMemcpyPeerAsync over NVLink.
#include <algorithm>
#include <cstring>
#include <limits>

#include "mb2_ctx.h"

struct mb2_plan {
  int dim = 0, nd = 0, world = 1, rank = 0;
  int64_t NE = 0, NV = 0, e0 = 0, e1 = 0;
  std::vector<int64_t> local;        // global ids: own elements, then ghosts (ascending)
  std::vector<uint8_t> owned;        // per node
  struct Peer {
    int peer;
    std::vector<int64_t> send;       // local indices (ascending) of own elements the peer needs
    int64_t recv_first = 0, n_recv = 0;
  };
  std::vector<Peer> peers;
  // some rank's outgoing elements lie in the second half of its element range: a receiver's NCCL kernel launched at the start
  // of the step would spin on SMs until that rank has packed them (option "mg_exchange_late", mb2_assembly.cu)
  bool late_exchange = false;
  Peer& peer_of(int r) {
    for (auto& p : peers)
      if (p.peer == r) return p;
    peers.push_back(Peer{r, {}, 0, 0});
    return peers.back();
  }
};

extern "C" {

int mb2_plan_partition(const int32_t* v_dofs, int dim, int nd, int64_t NE, int64_t NV, int world, int rank, mb2_plan** out) {
  if (!v_dofs || !out || dim < 2 || dim > 3 || nd < 1 || NE < 1 || NV % dim || world < 1 || rank < 0 || rank >= world) return MB2_ERR_ARG;
  mb2_plan* P = new mb2_plan();
  P->dim = dim; P->nd = nd; P->world = world; P->rank = rank; P->NE = NE; P->NV = NV;
  const int tdof = dim * nd;
  const int64_t Nn = NV / dim;
  auto range = [&](int r) { return (NE * r) / world; };
  P->e0 = range(rank);
  P->e1 = range(rank + 1);
  // first (lowest) element of every node -> owner rank
  std::vector<int64_t> first((size_t)Nn, std::numeric_limits<int64_t>::max());
  for (int64_t e = NE - 1; e >= 0; --e)
    for (int a = 0; a < nd; ++a) first[(size_t)(v_dofs[e * tdof + a] / dim)] = e;
  auto rank_of_elem = [&](int64_t e) {
    int r = (int)((e * world) / NE);               // first guess, then correct for the integer ranges
    while (r + 1 < world && range(r + 1) <= e) ++r;
    while (r > 0 && range(r) > e) --r;
    return r;
  };
  std::vector<int32_t> owner((size_t)Nn, -1);
  P->owned.assign((size_t)Nn, 0);
  for (int64_t n = 0; n < Nn; ++n)
    if (first[(size_t)n] != std::numeric_limits<int64_t>::max()) {
      owner[(size_t)n] = rank_of_elem(first[(size_t)n]);
      P->owned[(size_t)n] = owner[(size_t)n] == rank;
    }
  for (int64_t e = 0; e < NE && !P->late_exchange; ++e) {
    const int re = rank_of_elem(e);
    if (2 * (e - range(re)) < range(re + 1) - range(re)) continue;  // first half of its rank's range
    for (int a = 0; a < nd; ++a) {
      const int r = owner[(size_t)(v_dofs[e * tdof + a] / dim)];
      if (r >= 0 && r < re) { P->late_exchange = true; break; }
    }
  }
  for (int64_t e = P->e0; e < P->e1; ++e) P->local.push_back(e);
  // sends: own elements touching a node of a lower rank (one ascending list per such rank)
  for (int64_t e = P->e0; e < P->e1; ++e) {
    std::vector<int> seen;
    for (int a = 0; a < nd; ++a) {
      const int r = owner[(size_t)(v_dofs[e * tdof + a] / dim)];
      if (r == rank || r < 0 || std::find(seen.begin(), seen.end(), r) != seen.end()) continue;
      seen.push_back(r);
      P->peer_of(r).send.push_back(e - P->e0);
    }
  }
  // ghosts: elements of higher ranks touching an owned node; grouped by source rank they are contiguous blocks
  for (int64_t e = P->e1; e < NE; ++e) {
    bool mine = false;
    for (int a = 0; a < nd && !mine; ++a) mine = owner[(size_t)(v_dofs[e * tdof + a] / dim)] == rank;
    if (!mine) continue;
    const int src = rank_of_elem(e);
    mb2_plan::Peer& p = P->peer_of(src);
    if (p.n_recv == 0) p.recv_first = (int64_t)P->local.size();
    p.n_recv++;
    P->local.push_back(e);
  }
  std::sort(P->peers.begin(), P->peers.end(), [](const mb2_plan::Peer& a, const mb2_plan::Peer& b) { return a.peer < b.peer; });
  *out = P;
  return MB2_OK;
}

void mb2_plan_free(mb2_plan* p) { delete p; }

int64_t mb2_plan_size(const mb2_plan* p, int what) {
  if (!p) return -1;
  switch (what) {
    case 0: return p->e1 - p->e0;               // elements computed here
    case 1: return (int64_t)p->local.size();    // + ghosts
    case 2: return p->e0;                       // global id of the first own element
    case 3: return (int64_t)p->peers.size();
    case 4: return p->NV / p->dim;              // nodes
    case 5: return p->late_exchange ? 1 : 0;    // the NCCL exchange should wait for the element kernels (see mb2_set_option)
  }
  return -1;
}
int mb2_plan_local_elems(const mb2_plan* p, int64_t* out) {
  if (!p || !out) return MB2_ERR_ARG;
  std::memcpy(out, p->local.data(), sizeof(int64_t) * p->local.size());
  return MB2_OK;
}
int mb2_plan_owned_nodes(const mb2_plan* p, uint8_t* out) {
  if (!p || !out) return MB2_ERR_ARG;
  std::memcpy(out, p->owned.data(), p->owned.size());
  return MB2_OK;
}
int mb2_plan_peer(const mb2_plan* p, int k, int* peer, int64_t* n_send, int64_t* recv_first, int64_t* n_recv) {
  if (!p || k < 0 || k >= (int)p->peers.size()) return MB2_ERR_ARG;
  const auto& q = p->peers[(size_t)k];
  if (peer) *peer = q.peer;
  if (n_send) *n_send = (int64_t)q.send.size();
  if (recv_first) *recv_first = q.recv_first;
  if (n_recv) *n_recv = q.n_recv;
  return MB2_OK;
}
int mb2_plan_peer_send(const mb2_plan* p, int k, int64_t* send_local) {
  if (!p || k < 0 || k >= (int)p->peers.size() || !send_local) return MB2_ERR_ARG;
  std::memcpy(send_local, p->peers[(size_t)k].send.data(), sizeof(int64_t) * p->peers[(size_t)k].send.size());
  return MB2_OK;
}

// mb2_set_connectivity (own + ghost elements, global node ids) + mb2_set_partition + mb2_mg_set_send / _recv
int mb2_plan_apply(mb2_ctx* c, const mb2_plan* p, const int32_t* v_dofs_global, int nq) {
  if (!c) return MB2_ERR_ARG;
  if (!p || !v_dofs_global) { c->err = "null plan / connectivity"; return MB2_ERR_ARG; }
  const int tdof = p->dim * p->nd;
  std::vector<int32_t> vd(p->local.size() * (size_t)tdof);
  for (size_t x = 0; x < p->local.size(); ++x)
    std::memcpy(&vd[x * tdof], v_dofs_global + p->local[x] * tdof, sizeof(int32_t) * tdof);
  int rc = mb2_set_connectivity(c, p->dim, p->nd, nq, (int64_t)p->local.size(), p->NV, vd.data());
  if (rc) return rc;
  if (p->world == 1) return MB2_OK;
  if ((rc = mb2_set_partition(c, p->e1 - p->e0, p->owned.data()))) return rc;
  if ((rc = mb2_set_option(c, "mg_exchange_late", p->late_exchange ? 1 : 0))) return rc;
  for (const auto& q : p->peers) {
    if (!q.send.empty() && (rc = mb2_mg_set_send(c, q.peer, q.send.data(), (int64_t)q.send.size()))) return rc;
    if (q.n_recv && (rc = mb2_mg_set_recv(c, q.peer, q.recv_first, q.n_recv))) return rc;
  }
  return MB2_OK;
}

}  // extern "C"
