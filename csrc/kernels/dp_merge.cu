// K-M: sample-parallel (data-parallel) gradient merge -- an EXTENSION (BASELINE.json lists an "MPI
// sample-parallel gradient merge"; the reference itself never splits samples; the serial
// loop being parallelised is src/libhpnn.c:1221-1288).  Every rank holds a full replica and one sample.
//
// A BP update is a rank-1 matrix per layer, lr*delta_l (x) h_(l-1).  The merged update of P ranks is a
// rank-P matrix, so what has to cross NVLink is not the P gradient matrices (O(N*M) each) but their
// FACTORS (O(N+M) each):
//   k_dp_publish : copies this rank's factors into slot [rank] of EVERY rank's factor buffer -- one
//                  multimem.st per element when the buffer has an NVLS multicast mapping, a store per
//                  peer otherwise -- then fence.sys and a step flag on every rank;
//   k_dp_apply   : waits (on the device, bounded) until the flags of all ranks show this step, then
//                  streams its own replica once: W[i][j] += sum_r u_r[i] * v_r[j], ranks in index order,
//                  so all replicas stay bit-identical without any broadcast.
// No host barrier, no all-reduce of matrices.  For comparison the matrix form is kept too:
//   k_grad_merge (ops.cu)  peer loads of P gradient matrices;
//   k_grad_merge_nvls      rank r reduces slice r of the P matrices with multimem.ld_reduce (the switch
//                          adds) and writes the updated weights into every replica with multimem.st.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "dev_common.cuh"

namespace hpnn {
void set_error(const char *what, cudaError_t e);
void set_error_msg(const char *what);
extern long long g_launches;

struct DpPtrs {
  double *dst[HPNN_MAX_WORLD];               // factor slot of THIS rank in every rank's buffer
  unsigned long long *flag[HPNN_MAX_WORLD];  // step flag of THIS rank in every rank's buffer
  double *dst_mc;                            // the same slot / flag at the multicast address, or nullptr
  unsigned long long *flag_mc;
};

__device__ __forceinline__ unsigned long long dp_now_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

// one CTA: the factor vector is a few thousand doubles
__global__ void __launch_bounds__(1024, 1) k_dp_publish(const double *src, long long n, DpPtrs p, int world,
                                                        unsigned long long step) {
  for (long long i = threadIdx.x; i < n; i += blockDim.x) {
    const double v = src[i];
    if (p.dst_mc) multimem_st_f64(p.dst_mc + i, v);
    else
      for (int q = 0; q < world; ++q) stg_cg(p.dst[q] + i, v);
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    fence_sys();  // the factors are visible everywhere before the flag
    if (p.flag_mc) multimem_st_u64(p.flag_mc, step);
    else
      for (int q = 0; q < world; ++q) st_relaxed_sys(p.flag[q], step);
  }
}

constexpr int kDpTile = 1024;  // columns per CTA tile (double2 per thread, 512 threads)
constexpr int kDpRows = 64;    // rows per CTA

// W[i][j] += sum_r U[r*us + i] * V[r*vs + j]  (r = 0..world-1 in order), after all flags show `step`
__global__ void __launch_bounds__(512, 2) k_dp_apply(double *W, int ld, int rows, int cols, const double *U, long long us,
                                                    const double *V, long long vs, int world,
                                                    const unsigned long long *flags, unsigned long long step,
                                                    unsigned long long budget_ns, int *status) {
  __shared__ double su[HPNN_MAX_WORLD][kDpRows];
  __shared__ int s_bad;
  if (threadIdx.x == 0) {
    s_bad = 0;
    const unsigned long long t0 = dp_now_ns();
    for (int q = 0; q < world; ++q) {
      unsigned spins = 0;
      while (ld_relaxed_sys(flags + q) < step) {
        if ((++spins & 1023u) == 0 && budget_ns && dp_now_ns() - t0 > budget_ns) {
          s_bad = 1;
          break;
        }
      }
      if (ld_acquire_sys(flags + q) < step) s_bad = 1;
    }
    if (s_bad && status) *status = 1;
  }
  __syncthreads();
  if (s_bad) return;  // a rank never published: leave the replica untouched and report
  const int r0 = blockIdx.y * kDpRows;
  const int nr = min(kDpRows, rows - r0);
  for (int i = threadIdx.x; i < world * kDpRows; i += blockDim.x) {
    const int q = i / kDpRows, r = i % kDpRows;
    su[q][r] = r < nr ? ldg_cg(U + (size_t)q * us + r0 + r) : 0.0;
  }
  const int c = blockIdx.x * kDpTile + 2 * threadIdx.x;
  double2 v[HPNN_MAX_WORLD];
#pragma unroll
  for (int q = 0; q < HPNN_MAX_WORLD; ++q) {
    v[q] = make_double2(0.0, 0.0);
    if (q < world && c < cols) {
      v[q].x = ldg_cg(V + (size_t)q * vs + c);
      if (c + 1 < cols) v[q].y = ldg_cg(V + (size_t)q * vs + c + 1);
    }
  }
  __syncthreads();
  if (c >= ld) return;
  double2 *Wp = reinterpret_cast<double2 *>(W + (size_t)r0 * ld + c);
  const size_t stride2 = (size_t)(ld >> 1);
  for (int r = 0; r < nr; r += 4) {
    double2 w[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) w[k] = (r + k < nr) ? __ldcs(Wp + (size_t)(r + k) * stride2) : make_double2(0.0, 0.0);
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      if (r + k < nr) {
        double ax = 0.0, ay = 0.0;
#pragma unroll
        for (int q = 0; q < HPNN_MAX_WORLD; ++q) {
          if (q < world) {
            const double u = su[q][r + k];
            ax = fma(u, v[q].x, ax);
            ay = fma(u, v[q].y, ay);
          }
        }
        w[k].x += ax;
        w[k].y += ay;
        __stcs(Wp + (size_t)(r + k) * stride2, w[k]);
      }
    }
  }
}

// matrix form through NVLS: rank `rank` owns elements [i0, i1)
__global__ void k_grad_merge_nvls(const double *W_local, double *W_mc, const double *G_mc, long long i0, long long i1,
                                  double scale) {
  for (long long i = i0 + (long long)blockIdx.x * blockDim.x + threadIdx.x; i < i1; i += (long long)gridDim.x * blockDim.x) {
    const double g = multimem_ld_reduce_add_f64(G_mc + i);  // the switch adds the P copies
    multimem_st_f64(W_mc + i, fma(scale, g, __ldcg(W_local + i)));  // and writes every replica
  }
}

}  // namespace hpnn

using namespace hpnn;

extern "C" {

int hpnn_dev_dp_publish(const double *src, long long n, double *const *dst_all, double *dst_mc,
                        unsigned long long *const *flag_all, unsigned long long *flag_mc, int world,
                        unsigned long long step, void *stream) {
  if (!src || n < 1 || !dst_all || !flag_all || world < 1 || world > HPNN_MAX_WORLD) {
    set_error_msg("hpnn_dev_dp_publish: bad arguments");
    return -1;
  }
  DpPtrs p;
  for (int q = 0; q < HPNN_MAX_WORLD; ++q) {
    p.dst[q] = q < world ? dst_all[q] : nullptr;
    p.flag[q] = q < world ? flag_all[q] : nullptr;
  }
  p.dst_mc = dst_mc;
  p.flag_mc = flag_mc;
  k_dp_publish<<<1, 1024, 0, (cudaStream_t)stream>>>(src, n, p, world, step);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_error("k_dp_publish", e);
    return -1;
  }
  ++g_launches;
  return 0;
}

int hpnn_dev_dp_apply(double *W, int ld, int rows, int cols, const double *U, long long u_stride, const double *V,
                      long long v_stride, int world, const unsigned long long *flags, unsigned long long step,
                      long long timeout_ms, int *status_dev, void *stream) {
  if (!W || !U || !V || !flags || rows < 1 || cols < 1 || ld < cols || (ld & 1) || world < 1 || world > HPNN_MAX_WORLD) {
    set_error_msg("hpnn_dev_dp_apply: bad arguments");
    return -1;
  }
  dim3 grid((ld + kDpTile - 1) / kDpTile, (rows + kDpRows - 1) / kDpRows);
  const unsigned long long budget = timeout_ms > 0 ? (unsigned long long)timeout_ms * 1000000ull : 0ull;
  k_dp_apply<<<grid, 512, 0, (cudaStream_t)stream>>>(W, ld, rows, cols, U, u_stride, V, v_stride, world, flags, step,
                                                     budget, status_dev);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_error("k_dp_apply", e);
    return -1;
  }
  ++g_launches;
  return 0;
}

int hpnn_dev_grad_merge_nvls(const double *W_local, double *W_mc, const double *G_mc, long long n, int rank, int world,
                             double scale, void *stream) {
  if (!W_local || !W_mc || !G_mc || n < 1 || world < 1 || rank < 0 || rank >= world) {
    set_error_msg("hpnn_dev_grad_merge_nvls: bad arguments (needs multicast-mapped weights and gradients)");
    return -1;
  }
  const long long i0 = n * rank / world, i1 = n * (rank + 1) / world;
  if (i1 <= i0) return 0;
  long long blocks = (i1 - i0 + 255) / 256;
  if (blocks > 8 * 148) blocks = 8 * 148;
  k_grad_merge_nvls<<<(int)blocks, 256, 0, (cudaStream_t)stream>>>(W_local, W_mc, G_mc, i0, i1, scale);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_error("k_grad_merge_nvls", e);
    return -1;
  }
  ++g_launches;
  return 0;
}

}  // extern "C"
