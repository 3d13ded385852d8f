// Helpers shared by the two deferred-update trainers (train_lazy.cu: any number of ranks, column
// tasks; train_lazy1.cu: one GPU, row panels): the 8-factor chunks and their dot products.
#pragma once
#include "kp_common.cuh"

namespace hpnn {

// rows of layer l owned by rank q (the balanced split of kp_common.cuh), tabulated once per launch:
// the consumers of per-(rank, CTA) slots ask "did CTA b of rank q produce anything" for hundreds of
// slots per phase, and row0_of() is two 64-bit divisions (measured at 8 ranks: 3 us per gather).
static __shared__ int s_rows_of[HPNN_MAX_LAYERS][HPNN_MAX_WORLD];
__device__ __forceinline__ void rows_table_init(const hpnn_dev_net &n) {
  for (int i = threadIdx.x; i < n.n_layers * n.world; i += blockDim.x) {
    const int l = i / n.world, q = i % n.world;
    s_rows_of[l][q] = row0_of(q + 1, n.dims[l + 1], n.world) - row0_of(q, n.dims[l + 1], n.world);
  }
  __syncthreads();
}

// block-wide sum of kLazyChunk values per thread; coef[j] (shared) = total for j < m, 0 otherwise.
// Fixed order: lanes by butterfly, warps in index order.
__device__ __forceinline__ void chunk_block_sum(double (&acc)[kLazyChunk], int m, double *coef, double *red) {
#pragma unroll
  for (int j = 0; j < kLazyChunk; ++j)
#pragma unroll
    for (int o = 16; o; o >>= 1) acc[j] += __shfl_xor_sync(0xffffffffu, acc[j], o);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  __syncthreads();  // `red` / `coef` may still be in use
  if (lane == 0) {
#pragma unroll
    for (int j = 0; j < kLazyChunk; ++j) red[warp * kLazyChunk + j] = acc[j];
  }
  __syncthreads();
  if (threadIdx.x < kLazyChunk) {
    double s = 0.0;
    for (int w = 0; w < kWarps; ++w) s += red[w * kLazyChunk + threadIdx.x];
    coef[threadIdx.x] = (threadIdx.x < m) ? s : 0.0;  // slots >= m may hold stale factors
  }
  __syncthreads();
}

// acc[j] += chunk[j] * x for j < m (chunk: 8 doubles in global memory, loaded as 4 x 16 bytes)
__device__ __forceinline__ void chunk_fma(const double *chunk, double x, int m, double (&acc)[kLazyChunk]) {
  const double2 *c = reinterpret_cast<const double2 *>(chunk);
  double2 q[kLazyChunk / 2];
#pragma unroll
  for (int k = 0; k < kLazyChunk / 2; ++k) q[k] = (2 * k < m) ? ldg_cg2(c + k) : make_double2(0.0, 0.0);
#pragma unroll
  for (int k = 0; k < kLazyChunk / 2; ++k) {
    acc[2 * k] = fma(q[k].x, x, acc[2 * k]);
    acc[2 * k + 1] = fma(q[k].y, x, acc[2 * k + 1]);
  }
}

// coef[j] = sum_i H[i*8 + j] * v[i], j < m: the m dot products in one sweep (H: chunks in global
// memory, v: shared memory).  Only used for short vectors (the output layer's deltas).
__device__ __forceinline__ void chunk_dots(const double *H, const double *v, int len, int m, double *coef, double *red) {
  double acc[kLazyChunk];
#pragma unroll
  for (int j = 0; j < kLazyChunk; ++j) acc[j] = 0.0;
#pragma unroll 2
  for (int i = threadIdx.x; i < len; i += kThreads) chunk_fma(H + (size_t)i * kLazyChunk, v[i], m, acc);
  chunk_block_sum(acc, m, coef, red);
}

// sum_j chunk[j] * coef[j], j < m (chunk: 8 doubles in global memory)
__device__ __forceinline__ double chunk_apply(const double *chunk, const double *coef, int m) {
  const double2 *c = reinterpret_cast<const double2 *>(chunk);
  double s = 0.0;
#pragma unroll
  for (int k = 0; k < kLazyChunk / 2; ++k) {
    if (2 * k < m) {
      const double2 q = ldg_cg2(c + k);
      s = fma(q.x, coef[2 * k], s);
      if (2 * k + 1 < m) s = fma(q.y, coef[2 * k + 1], s);  // slot m may be half written by another CTA
    }
  }
  return s;
}


// where the per-CTA partial dot products v_j . y of the NEXT layer go (one chunk per rank)
struct DotOut {
  double *p[HPNN_MAX_WORLD];  // slot of (this rank, this CTA) in every rank's arena; p[0] == nullptr: none
  int n;
  const double *HV;  // factor chunks of the next layer, indexed by the global row of this layer
  int row0;
  unsigned long long tag;  // != 0: publish the chunk as 8 {value, tag} pairs (tagged hand-over)
  double *mc;              // plain chunks only: the slot at the multicast address (one store reaches all ranks)
};


// coef[j] = sum over the producers' chunks (rank-major, CTA order): v_j . x for the layer whose
// input x was produced by the previous row phase.  rows_prev(q) = rows of that phase on rank q.
__device__ __forceinline__ void gather_dots(const hpnn_dev_net &n, int l, int m, double *coef, double *red, int grid) {
  const double *slots = n.arena[n.rank] + n.off_hist_dot[l];
  double acc[kLazyChunk];
#pragma unroll
  for (int j = 0; j < kLazyChunk; ++j) acc[j] = 0.0;
  for (int idx = threadIdx.x; idx < n.world * kLazyDotSlots; idx += kThreads) {
    const int q = idx / kLazyDotSlots, b = idx % kLazyDotSlots;
    const int rows_q = s_rows_of[l - 1][q];  // rows of layer l-1 on rank q
    if (b < grid && b < rows_q) chunk_fma(slots + (size_t)idx * kLazyChunk, 1.0, m, acc);
  }
  chunk_block_sum(acc, m, coef, red);
}

}  // namespace hpnn
