// Sparsity pattern and element->CSR index map, built on the device from the element connectivity.
//
// Replaces PrecomputedData::PrepareSparsity (src/mimi/utils/precomputed.cpp:151-218), which inserts
// every v_dofs x v_dofs block into a linked-list mfem::SparseMatrix, finalises, sorts the columns and
// reads the value positions back through GetSubMatrix. With byVDIM ordering the pattern is a block
// pattern over nodes: row dim*I+ci holds, for every neighbour node J of I in ascending order, the dim
// columns dim*J+cj. The reference's 6561-int A_ids per element (p=2) therefore factor into
//   A_id = rowptr[dim*I+ci] + dim*rank(I,J) + cj
// and only rank (nd*nd uint16 per element) is stored.
#include <vector>

#include "mb2_ctx.h"

namespace mb2 {

__global__ void k_enodes(int dim, int nd, int64_t NE, const int32_t* __restrict__ vdofs, int32_t* __restrict__ enodes,
                         int* __restrict__ flag) {
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= NE * nd) return;
  const int64_t e = i / nd;
  const int a = (int)(i % nd);
  const int32_t* vd = vdofs + e * dim * nd;
  const int32_t node = vd[a] / dim;
  for (int c = 0; c < dim; c++)
    if (vd[c * nd + a] != node * dim + c) atomicExch(flag, 1);  // not byVDIM / not component-major
  enodes[i] = node;
}

__global__ void k_count(int64_t n, const int32_t* __restrict__ enodes, int32_t* __restrict__ cnt) {
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i < n) atomicAdd(&cnt[enodes[i]], 1);
}

// ---- exclusive scan int32 -> int64 (three small kernels; setup only) ----
constexpr int SCAN_B = 1024;
__global__ void k_scan_block(int64_t n, const int32_t* __restrict__ in, int64_t* __restrict__ out,
                             int64_t* __restrict__ bsum) {
  __shared__ int64_t s[SCAN_B];
  const int64_t i = blockIdx.x * (int64_t)SCAN_B + threadIdx.x;
  const int64_t v = (i < n) ? in[i] : 0;
  s[threadIdx.x] = v;
  __syncthreads();
  for (int off = 1; off < SCAN_B; off <<= 1) {
    int64_t x = (threadIdx.x >= off) ? s[threadIdx.x - off] : 0;
    __syncthreads();
    s[threadIdx.x] += x;
    __syncthreads();
  }
  if (i < n) out[i] = s[threadIdx.x] - v;
  if (threadIdx.x == SCAN_B - 1) bsum[blockIdx.x] = s[threadIdx.x];
}
__global__ void k_scan_bsum(int nb, int64_t* bsum) {  // one thread: nb <= a few thousand
  int64_t run = 0;
  for (int i = 0; i < nb; i++) {
    const int64_t v = bsum[i];
    bsum[i] = run;
    run += v;
  }
  bsum[nb] = run;
}
__global__ void k_scan_add(int64_t n, int64_t* __restrict__ out, const int64_t* __restrict__ bsum, int nb) {
  const int64_t i = blockIdx.x * (int64_t)SCAN_B + threadIdx.x;
  if (i < n) out[i] += bsum[blockIdx.x];
  if (i == n) out[n] = bsum[nb];
}

static int exclusive_scan(mb2_ctx* c, int64_t n, const int32_t* d_in, int64_t* d_out /* n+1 */) {
  const int nb = (int)((n + SCAN_B) / SCAN_B);  // covers index n too
  int64_t* d_bsum = nullptr;
  MB2_CUDA(cudaMalloc(&d_bsum, sizeof(int64_t) * (nb + 1)));
  k_scan_block<<<nb, SCAN_B, 0, c->stream>>>(n, d_in, d_out, d_bsum);
  k_scan_bsum<<<1, 1, 0, c->stream>>>(nb, d_bsum);
  k_scan_add<<<nb, SCAN_B, 0, c->stream>>>(n, d_out, d_bsum, nb);
  MB2_CUDA(cudaGetLastError());
  MB2_CUDA(cudaStreamSynchronize(c->stream));
  MB2_CUDA(cudaFree(d_bsum));
  return MB2_OK;
}

__global__ void k_fill_n2e(int64_t n, const int32_t* __restrict__ enodes, const int64_t* __restrict__ ptr,
                           int32_t* __restrict__ fill, int32_t* __restrict__ n2e_elem, int nd) {
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int32_t node = enodes[i];
  const int pos = atomicAdd(&fill[node], 1);
  n2e_elem[ptr[node] + pos] = (int32_t)(i / nd);
}

// ascending element id per node (the slot order above is arbitrary), then the local index of the node
__global__ void k_sort_n2e(int64_t Nn, int nd, const int64_t* __restrict__ ptr, int32_t* __restrict__ n2e_elem,
                           uint8_t* __restrict__ n2e_loc, const int32_t* __restrict__ enodes) {
  const int64_t I = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (I >= Nn) return;
  const int64_t b = ptr[I], e = ptr[I + 1];
  for (int64_t i = b + 1; i < e; i++) {
    const int32_t v = n2e_elem[i];
    int64_t j = i - 1;
    while (j >= b && n2e_elem[j] > v) {
      n2e_elem[j + 1] = n2e_elem[j];
      j--;
    }
    n2e_elem[j + 1] = v;
  }
  for (int64_t i = b; i < e; i++) {
    const int32_t* en = enodes + (int64_t)n2e_elem[i] * nd;
    int loc = 0;
    for (int a = 0; a < nd; a++)
      if (en[a] == (int32_t)I) loc = a;
    n2e_loc[i] = (uint8_t)loc;
  }
}

// One CTA per node: gather the nodes of all its elements, bitonic-sort in shared memory, unique.
// out == nullptr: only count (nnbr). Otherwise write the sorted unique list at nodeptr[I].
__global__ void k_neighbours(int64_t Nn, int nd, int cap /* pow2 */, const int64_t* __restrict__ ptr,
                             const int32_t* __restrict__ n2e_elem, const int32_t* __restrict__ enodes,
                             int32_t* __restrict__ nnbr, const int64_t* __restrict__ nodeptr,
                             int32_t* __restrict__ out, const uint8_t* __restrict__ owned) {
  extern __shared__ int32_t s[];  // cap candidates + 1 counter
  __shared__ int s_cnt;
  for (int64_t I = blockIdx.x; I < Nn; I += gridDim.x) {
    // rows of nodes owned by another rank are not assembled here: empty block row
    if (owned && !owned[I]) {
      if (out == nullptr && threadIdx.x == 0) nnbr[I] = 0;
      continue;
    }
    const int64_t b = ptr[I];
    const int ne = (int)(ptr[I + 1] - b);
    const int n = ne * nd;
    for (int i = threadIdx.x; i < cap; i += blockDim.x)
      s[i] = (i < n) ? enodes[(int64_t)n2e_elem[b + i / nd] * nd + i % nd] : 0x7fffffff;
    if (threadIdx.x == 0) s_cnt = 0;
    __syncthreads();
    for (int k = 2; k <= cap; k <<= 1)
      for (int j = k >> 1; j > 0; j >>= 1) {
        for (int i = threadIdx.x; i < cap; i += blockDim.x) {
          const int ixj = i ^ j;
          if (ixj > i) {
            const int32_t x = s[i], y = s[ixj];
            const bool up = ((i & k) == 0);
            if ((x > y) == up) {
              s[i] = y;
              s[ixj] = x;
            }
          }
        }
        __syncthreads();
      }
    // unique: position = number of "first occurrences" before i (small lists: serial prefix per thread is fine)
    if (out == nullptr) {
      int local = 0;
      for (int i = threadIdx.x; i < n; i += blockDim.x)
        if (i == 0 || s[i] != s[i - 1]) local++;
      atomicAdd(&s_cnt, local);
      __syncthreads();
      if (threadIdx.x == 0) nnbr[I] = s_cnt;
    } else {
      if (threadIdx.x == 0) {
        int32_t* o = out + nodeptr[I];
        int m = 0;
        for (int i = 0; i < n; i++)
          if (i == 0 || s[i] != s[i - 1]) o[m++] = s[i];
      }
    }
    __syncthreads();
  }
}

// Row offsets (64-bit) and column indices.
__global__ void k_rows(int dim, int64_t Nn, const int32_t* __restrict__ nnbr, const int64_t* __restrict__ nodeptr,
                       int64_t* __restrict__ I) {
  const int64_t n = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (n > Nn) return;
  if (n == Nn) {
    I[dim * Nn] = (int64_t)dim * dim * nodeptr[Nn];
    return;
  }
  for (int c = 0; c < dim; c++) I[dim * n + c] = (int64_t)dim * dim * nodeptr[n] + (int64_t)c * dim * nnbr[n];
}
__global__ void k_cols(int dim, int64_t Nn, const int32_t* __restrict__ nnbr, const int64_t* __restrict__ nodeptr,
                       const int32_t* __restrict__ nbr, int32_t* __restrict__ J) {
  const int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;  // over nodeptr[Nn] (node, slot) pairs
  if (t >= nodeptr[Nn]) return;
  // find node by binary search in nodeptr
  int64_t lo = 0, hi = Nn;
  while (hi - lo > 1) {
    const int64_t mid = (lo + hi) >> 1;
    if (nodeptr[mid] <= t) lo = mid; else hi = mid;
  }
  const int64_t n = lo;
  const int s = (int)(t - nodeptr[n]);
  const int nn = nnbr[n];
  const int32_t col0 = nbr[t] * dim;
  for (int ci = 0; ci < dim; ci++)
    for (int cj = 0; cj < dim; cj++)
      J[(int64_t)dim * dim * nodeptr[n] + (int64_t)ci * dim * nn + dim * s + cj] = col0 + cj;
}

__global__ void k_rank(int nd, int64_t NE, const int32_t* __restrict__ enodes, const int32_t* __restrict__ nnbr,
                       const int64_t* __restrict__ nodeptr, const int32_t* __restrict__ nbr,
                       uint16_t* __restrict__ rank) {
  const int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (t >= NE * nd * nd) return;
  const int b = (int)(t % nd);
  const int a = (int)((t / nd) % nd);
  const int64_t e = t / ((int64_t)nd * nd);
  const int32_t A = enodes[e * nd + a], B = enodes[e * nd + b];
  const int32_t* list = nbr + nodeptr[A];
  int lo = 0, hi = nnbr[A] - 1, pos = 0;
  while (lo <= hi) {
    const int mid = (lo + hi) >> 1;
    if (list[mid] == B) { pos = mid; break; }
    if (list[mid] < B) lo = mid + 1; else hi = mid - 1;
  }
  rank[t] = (uint16_t)pos;
}

// Expansion of the factored map to the reference's A_ids layout (tests): A_ids[e][j*tdof + i] is the
// position of (row v_dofs[i], col v_dofs[j]).
__global__ void k_expand_A_ids(int dim, int nd, int64_t NE, const int32_t* __restrict__ enodes,
                               const uint16_t* __restrict__ rank, const int64_t* __restrict__ I,
                               int32_t* __restrict__ out) {
  const int tdof = dim * nd;
  const int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (t >= NE * tdof * tdof) return;
  const int i = (int)(t % tdof), j = (int)((t / tdof) % tdof);
  const int64_t e = t / ((int64_t)tdof * tdof);
  const int a = i % nd, ci = i / nd, b = j % nd, cj = j / nd;
  const int32_t A = enodes[e * nd + a];
  out[t] = (int32_t)(I[(int64_t)dim * A + ci] + dim * rank[(e * nd + a) * nd + b] + cj);
}

static inline unsigned nblk(int64_t n, int b) { return (unsigned)((n + b - 1) / b); }

int build_pattern(mb2_ctx* c) {
  const int dim = c->dim, nd = c->nd;
  const int64_t NE = c->NE, Nn = c->Nn;
  cudaStream_t st = c->stream;
  if (nd > 255) { c->err = "nd > 255 unsupported"; return MB2_ERR_UNSUPPORTED; }

  // node ids per element + ordering check
  MB2_CUDA(cudaMalloc(&c->d_enodes, sizeof(int32_t) * NE * nd));
  MB2_CUDA(cudaMemsetAsync(c->d_flag, 0, sizeof(int), st));
  k_enodes<<<nblk(NE * nd, 256), 256, 0, st>>>(dim, nd, NE, c->d_vdofs, c->d_enodes, c->d_flag);
  int flag = 0;
  MB2_CUDA(cudaMemcpyAsync(&flag, c->d_flag, sizeof(int), cudaMemcpyDeviceToHost, st));
  MB2_CUDA(cudaStreamSynchronize(st));
  if (flag) { c->err = "v_dofs are not component-major byVDIM (dim*node + c)"; return MB2_ERR_UNSUPPORTED; }

  // node -> elements
  int32_t* d_cnt = nullptr;
  MB2_CUDA(cudaMalloc(&d_cnt, sizeof(int32_t) * (Nn + 1)));
  MB2_CUDA(cudaMemsetAsync(d_cnt, 0, sizeof(int32_t) * (Nn + 1), st));
  k_count<<<nblk(NE * nd, 256), 256, 0, st>>>(NE * nd, c->d_enodes, d_cnt);
  MB2_CUDA(cudaMalloc(&c->d_n2e_ptr, sizeof(int64_t) * (Nn + 1)));
  int rc = exclusive_scan(c, Nn, d_cnt, c->d_n2e_ptr);
  if (rc) return rc;
  MB2_CUDA(cudaMalloc(&c->d_n2e_elem, sizeof(int32_t) * NE * nd));
  MB2_CUDA(cudaMalloc(&c->d_n2e_loc, sizeof(uint8_t) * NE * nd));
  MB2_CUDA(cudaMemsetAsync(d_cnt, 0, sizeof(int32_t) * (Nn + 1), st));
  k_fill_n2e<<<nblk(NE * nd, 256), 256, 0, st>>>(NE * nd, c->d_enodes, c->d_n2e_ptr, d_cnt, c->d_n2e_elem, nd);
  k_sort_n2e<<<nblk(Nn, 128), 128, 0, st>>>(Nn, nd, c->d_n2e_ptr, c->d_n2e_elem, c->d_n2e_loc, c->d_enodes);
  MB2_CUDA(cudaGetLastError());

  // max elements per node -> candidate capacity (power of two)
  // (p+1)^dim elements share an interior node of a degree-p patch; take the true max from the counts.
  std::vector<int32_t> h_cnt((size_t)Nn);
  // after k_fill_n2e the fill counters equal the per-node element counts
  MB2_CUDA(cudaMemcpyAsync(h_cnt.data(), d_cnt, sizeof(int32_t) * Nn, cudaMemcpyDeviceToHost, st));
  MB2_CUDA(cudaStreamSynchronize(st));
  int maxe = 1;
  for (int64_t i = 0; i < Nn; i++) maxe = h_cnt[i] > maxe ? h_cnt[i] : maxe;
  int cap = 32;
  while (cap < maxe * nd) cap <<= 1;
  if (cap > 8192) { c->err = "node valence too large for the neighbour sort"; return MB2_ERR_UNSUPPORTED; }
  const size_t smem = sizeof(int32_t) * cap;
  MB2_CUDA(cudaFuncSetAttribute(k_neighbours, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  const unsigned grid = (unsigned)(Nn < (int64_t)c->sm_count * 16 ? Nn : (int64_t)c->sm_count * 16);

  MB2_CUDA(cudaMalloc(&c->d_nnbr, sizeof(int32_t) * (Nn + 1)));
  MB2_CUDA(cudaMemsetAsync(c->d_nnbr, 0, sizeof(int32_t) * (Nn + 1), st));
  k_neighbours<<<grid, 256, smem, st>>>(Nn, nd, cap, c->d_n2e_ptr, c->d_n2e_elem, c->d_enodes, c->d_nnbr, nullptr,
                                        nullptr, c->d_owned);
  MB2_CUDA(cudaGetLastError());
  MB2_CUDA(cudaMalloc(&c->d_nodeptr, sizeof(int64_t) * (Nn + 1)));
  rc = exclusive_scan(c, Nn, c->d_nnbr, c->d_nodeptr);
  if (rc) return rc;
  int64_t nblocks = 0;
  MB2_CUDA(cudaMemcpy(&nblocks, c->d_nodeptr + Nn, sizeof(int64_t), cudaMemcpyDeviceToHost));
  c->nnz = nblocks * dim * dim;
  MB2_CUDA(cudaMalloc(&c->d_nbr, sizeof(int32_t) * nblocks));
  k_neighbours<<<grid, 256, smem, st>>>(Nn, nd, cap, c->d_n2e_ptr, c->d_n2e_elem, c->d_enodes, c->d_nnbr,
                                        c->d_nodeptr, c->d_nbr, c->d_owned);
  MB2_CUDA(cudaGetLastError());

  MB2_CUDA(cudaMalloc(&c->d_I, sizeof(int64_t) * (c->NV + 1)));
  MB2_CUDA(cudaMalloc(&c->d_J, sizeof(int32_t) * c->nnz));
  k_rows<<<nblk(Nn + 1, 256), 256, 0, st>>>(dim, Nn, c->d_nnbr, c->d_nodeptr, c->d_I);
  k_cols<<<nblk(nblocks, 256), 256, 0, st>>>(dim, Nn, c->d_nnbr, c->d_nodeptr, c->d_nbr, c->d_J);
  MB2_CUDA(cudaMalloc(&c->d_rank, sizeof(uint16_t) * NE * nd * nd));
  k_rank<<<nblk(NE * nd * nd, 256), 256, 0, st>>>(nd, NE, c->d_enodes, c->d_nnbr, c->d_nodeptr, c->d_nbr, c->d_rank);
  MB2_CUDA(cudaGetLastError());
  MB2_CUDA(cudaStreamSynchronize(st));
  MB2_CUDA(cudaFree(d_cnt));
  c->have_pattern = true;
  return MB2_OK;
}

int expand_A_ids(mb2_ctx* c, int32_t* d_out) {
  const int64_t n = c->NE * c->tdof * c->tdof;
  k_expand_A_ids<<<nblk(n, 256), 256, 0, c->stream>>>(c->dim, c->nd, c->NE, c->d_enodes, c->d_rank, c->d_I, d_out);
  MB2_CUDA(cudaGetLastError());
  return MB2_OK;
}

}  // namespace mb2
