// Stand-alone phase kernels K-A..K-E (one launch each), batched inference (K-G fp64),
// row softmax, data-parallel gradient merge (K-M) and the small CUDA runtime helpers
// (allocation, copies, IPC) the framework layer and the C shim share.
//
// What each kernel replaces in the reference:
//   k_forward_layer    cublasDgemv(OP_T)+sigmoid / fw_mv_acc / fw_s_acc  cuda_ann.cu:546-551,77-86
//   k_output_stats     amb+Dasum / amb_acc / dsigmoid_mul_diff / softmax_acc+fw_scal /
//                      amb_smax / dsmax_diff                 cuda_ann.cu:1280-1346, cuda_snn.cu:938-1171
//   k_backward_update  cublasDgemv(OP_N)+dsigmoid / dsigmoid_mul_delta_T + cublasDger /
//                      ger_acc / Dger+Daxpy+Dscal / ger_dw_acc   cuda_ann.cu:1447-1453,2139,3155-3165
//   k_update_forward   ger_acc|ger_dw_acc followed by fw_mv_acc  cuda_ann.cu:124-148,77-86
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include "dev_common.cuh"

namespace hpnn {
extern long long g_launches;
void set_error(const char *what, cudaError_t e);
void set_error_msg(const char *what);

static int num_sms() {
  static int sms = 0;
  if (!sms) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    if (sms <= 0) sms = 148;
  }
  return sms;
}

#define HPNN_CHECK_LAUNCH(name)                  \
  do {                                           \
    cudaError_t e__ = cudaGetLastError();        \
    if (e__ != cudaSuccess) {                    \
      set_error(name, e__);                      \
      return -1;                                 \
    }                                            \
    ++g_launches;                                \
  } while (0)

// ------------------------------- K-A ---------------------------------------
template <int UPD>
__global__ void __launch_bounds__(kThreads, 1)
    k_rows(double *W, double *dW, int ld, int rows, const double *x, const double *delta, double *y,
           double lr, double alpha, int act_kind) {
  extern __shared__ double smem[];
  double *xs = smem;
  double *rowpart = smem + ld;
  double *ds = rowpart + row_part_doubles(rows, gridDim.x);
  stage_vec(xs, x, ld);
  if (UPD)
    for (int r = threadIdx.x; r < rows; r += blockDim.x) ds[r] = ldg_cg(delta + r);
  __syncthreads();
  Dst dst;
  dst.n = 1;
  dst.tag = 0;
  dst.mc = nullptr;
  dst.p[0] = y;
  phase_rows<UPD>(W, dW, ld, rows, xs, ds, lr, alpha, act_kind, dst, pick_row_segments(rows, ld, gridDim.x), rowpart, gridDim.x, blockIdx.x);
}

// ------------------------------- K-B ---------------------------------------
__global__ void __launch_bounds__(kThreads, 1)
    k_output_stats(int type, int stable, const double *z, const double *t, int n, double *o, double *delta,
                   double *stats) {
  extern __shared__ double smem[];
  double *ds = smem;          // n doubles
  double *red = smem + ((n + 7) & ~7);
  OutStats st = output_stats(type, stable, z, t, n, 0, n, ds, o, red);
  for (int j = threadIdx.x; j < n; j += blockDim.x)
    if (delta) delta[j] = ds[j];
  if (threadIdx.x == 0 && stats) {
    stats[0] = st.err;
    stats[1] = (double)st.argmax;
    stats[2] = (double)st.target;
  }
}

// ------------------------------- K-C (+K-D/K-E) ------------------------------
template <int UPD>
__global__ void __launch_bounds__(kThreads, 1)
    k_cols(double *W, double *dW, int ld, int rows, int cols, int cb, const double *delta, const double *h,
           double *part, double lr, double alpha) {
  extern __shared__ double smem[];
  double *ds = smem;
  double *red = smem + ((rows + 7) & ~7);
  for (int r = threadIdx.x; r < rows; r += blockDim.x) ds[r] = ldg_cg(delta + r);
  __syncthreads();
  auto make_sink = [=](int c0, int) {
    return [=](int i, double v) {
      if (c0 + i < cols) part[c0 + i] = v;
    };
  };
  phase_cols<UPD>(W, dW, ld, rows, cb, 1, ds, h, false, lr, alpha, red, make_sink, gridDim.x, blockIdx.x);
}

__global__ void k_delta_finalize(const double *part, int n_parts, int stride, const double *h, int n,
                                 double *delta) {
  for (int r = blockIdx.x * blockDim.x + threadIdx.x; r < n; r += gridDim.x * blockDim.x) {
    double s = 0.0;
    for (int g = 0; g < n_parts; ++g) s += part[(size_t)g * stride + r];
    delta[r] = s * dact_f(h[r]);
  }
}

// ------------------------------- K-G (fp64) ---------------------------------
// Y[b, r] = f( sum_c X[b, c] * W[r, c] ) in FP64.  tcgen05 has no FP64 kind, so the FP64 tensor
// path is the warp-level DMMA (mma.sync.m8n8k4.f64): 64x64 output tile per CTA, 4 warps, each
// warp a 32x32 sub-tile = 4x4 DMMA tiles; operands stay K-major in shared memory exactly as they
// are in global memory (both X and W are row-major with K contiguous), staged by a double-
// buffered cp.async pipeline with a 20-double row pitch (conflict-free fragment loads).
// The reference has no batched path (batch = 1); measured against cuBLAS DGEMM.
constexpr int GB = 64, GK = 16, GP = 20;  // tile, k-slab, shared row pitch (doubles)
__device__ __forceinline__ void cp_async16(void *smem, const void *gmem, bool valid) {
  const unsigned s = (unsigned)__cvta_generic_to_shared(smem);
  const int bytes = valid ? 16 : 0;  // src-size 0 => zero fill
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(s), "l"(gmem), "r"(bytes) : "memory");
}
__device__ __forceinline__ void dmma_m8n8k4(double (&c)[2], double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
               : "+d"(c[0]), "+d"(c[1])
               : "d"(a), "d"(b));
}
// where the output tile goes: the same [batch x ldy] matrix on up to HPNN_MAX_WORLD ranks,
// columns shifted by col0 (this rank's row slice of the layer) -- all-gather by peer stores
struct GemmDst {
  double *y[HPNN_MAX_WORLD];
  int n, col0;
  double *y_mc;  // the same matrix at the NVLS multicast address (one store reaches every rank), or nullptr
};
// two adjacent doubles through the multicast mapping in one 16-byte store (bits travel as 4 x b32)
__device__ __forceinline__ void multimem_st_2f64(double *mc, double a, double b) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc),
               "f"(__uint_as_float((unsigned)__double2loint(a))), "f"(__uint_as_float((unsigned)__double2hiint(a))),
               "f"(__uint_as_float((unsigned)__double2loint(b))), "f"(__uint_as_float((unsigned)__double2hiint(b)))
               : "memory");
}
__global__ void __launch_bounds__(128)
    k_forward_batch_f64(const double *__restrict__ W, int ld, int rows, int cols, const double *__restrict__ X,
                        int ldx, int batch, GemmDst dst, int ldy, int act_kind) {
  __shared__ __align__(16) double Xs[2][GB][GP];
  __shared__ __align__(16) double Ws[2][GB][GP];
  const int b0 = blockIdx.y * GB, r0 = blockIdx.x * GB;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int wm = (warp >> 1) * 32, wn = (warp & 1) * 32;  // warp sub-tile origin (batch, rows)
  const int g = lane >> 2, q = lane & 3;                    // DMMA fragment coordinates
  double acc[4][4][2] = {};
  // stage one 64x16 slab of X and W: 64 rows x 8 chunks of 16 B each, 4 chunks per thread per matrix
  auto stage = [&](int buf, int k0) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int e = threadIdx.x + 128 * i;
      const int m = e >> 3, c = (e & 7) * 2;
      const bool kin = k0 + c < cols, kin2 = k0 + c + 1 < cols;
      double *xs = &Xs[buf][m][c], *ws = &Ws[buf][m][c];
      if (kin && !kin2) {  // odd tail column: plain stores, no async copy may touch this chunk
        xs[0] = (b0 + m < batch) ? X[(size_t)(b0 + m) * ldx + k0 + c] : 0.0;
        xs[1] = 0.0;
        ws[0] = (r0 + m < rows) ? W[(size_t)(r0 + m) * ld + k0 + c] : 0.0;
        ws[1] = 0.0;
      } else {             // whole chunk in range, or whole chunk zero-filled
        cp_async16(xs, X + (size_t)(b0 + m) * ldx + k0 + c, kin && (b0 + m < batch));
        cp_async16(ws, W + (size_t)(r0 + m) * ld + k0 + c, kin && (r0 + m < rows));
      }
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
  };
  const int nk = (cols + GK - 1) / GK;
  stage(0, 0);
  for (int kb = 0; kb < nk; ++kb) {
    const int buf = kb & 1;
    if (kb + 1 < nk) {
      stage(buf ^ 1, (kb + 1) * GK);
      asm volatile("cp.async.wait_group 1;" ::: "memory");
    } else {
      asm volatile("cp.async.wait_group 0;" ::: "memory");
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < GK; kk += 4) {
      double a[4], b[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        a[i] = Xs[buf][wm + 8 * i + g][kk + q];  // A[m = g][k = q]   (row-major 8x4)
        b[i] = Ws[buf][wn + 8 * i + g][kk + q];  // B[k = q][n = g]   (col-major 4x8)
      }
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) dmma_m8n8k4(acc[i][j], a[i], b[j]);
    }
    __syncthreads();
  }
  // C fragment: lane holds C[m = g][n = 2q], C[m = g][n = 2q + 1]
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int b = b0 + wm + 8 * i + g;
    if (b >= batch) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int r = r0 + wn + 8 * j + 2 * q;
      const double y0 = act_kind == 0 ? act_f(acc[i][j][0]) : acc[i][j][0];
      const double y1 = act_kind == 0 ? act_f(acc[i][j][1]) : acc[i][j][1];
      const size_t o = (size_t)b * ldy + dst.col0 + r;
      if (dst.y_mc) {  // GEMM fused with the all-gather of its output: the switch replicates the store
        if (r + 1 < rows && !(o & 1)) multimem_st_2f64(dst.y_mc + o, y0, y1);
        else {
          if (r < rows) multimem_st_f64(dst.y_mc + o, y0);
          if (r + 1 < rows) multimem_st_f64(dst.y_mc + o + 1, y1);
        }
      } else {
        for (int p = 0; p < dst.n; ++p) {
          if (r < rows) dst.y[p][o] = y0;
          if (r + 1 < rows) dst.y[p][o + 1] = y1;
        }
      }
    }
  }
}

// one warp per row of Z: exp(z-shift)/(TINY+sum)
__global__ void k_softmax_rows(double *Z, int ldz, int batch, int n, int stable) {
  const int lane = threadIdx.x & 31;
  const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= batch) return;
  double *z = Z + (size_t)row * ldz;
  double shift = 1.0;
  if (stable) {
    double m = -INFINITY;
    for (int j = lane; j < n; j += 32) m = fmax(m, z[j]);
    for (int o = 16; o; o >>= 1) m = fmax(m, __shfl_xor_sync(0xffffffffu, m, o));
    shift = m;
  }
  double s = 0.0;
  for (int j = lane; j < n; j += 32) {
    const double e = exp(z[j] - shift);
    z[j] = e;
    s += e;
  }
  for (int o = 16; o; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  const double dv = s + HPNN_TINY;
  for (int j = lane; j < n; j += 32) z[j] /= dv;
}

// ------------------------------- K-M ----------------------------------------
// W += scale * sum_r G_r, ranks summed in index order (deterministic).  G_r may be
// peer-mapped memory: the loads then travel over NVLink inside this kernel.
struct GradPtrs {
  const double *g[HPNN_MAX_WORLD];
};
__global__ void k_grad_merge(double *W, long long n, GradPtrs gp, int n_ranks, double scale) {
  const long long n2 = n >> 1;
  double2 *W2 = reinterpret_cast<double2 *>(W);
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n2;
       i += (long long)gridDim.x * blockDim.x) {
    double2 s = make_double2(0.0, 0.0);
    for (int r = 0; r < n_ranks; ++r) {
      const double2 g = __ldcg(reinterpret_cast<const double2 *>(gp.g[r]) + i);
      s.x += g.x;
      s.y += g.y;
    }
    double2 w = W2[i];
    w.x = fma(scale, s.x, w.x);
    w.y = fma(scale, s.y, w.y);
    W2[i] = w;
  }
  if ((n & 1) && blockIdx.x == 0 && threadIdx.x == 0) {
    double s = 0.0;
    for (int r = 0; r < n_ranks; ++r) s += gp.g[r][n - 1];
    W[n - 1] = fma(scale, s, W[n - 1]);
  }
}

}  // namespace hpnn

using namespace hpnn;

extern "C" {

int hpnn_dev_forward_layer(const double *W, int ld, int rows, int cols, const double *x, double *y,
                           int act_kind, void *stream) {
  (void)cols;
  if (ld % 8) {
    set_error_msg("hpnn_dev_forward_layer: ld must be a multiple of 8");
    return -1;
  }
  int grid = (rows + kWarps - 1) / kWarps;
  if (grid > num_sms()) grid = num_sms();
  if (grid < 1) grid = 1;
  const size_t smem = sizeof(double) * ((size_t)ld + row_part_doubles(rows, grid) + 8);
  if (smem > 227 * 1024) {
    set_error_msg("hpnn_dev_forward_layer: input vector too long");
    return -2;
  }
  cudaFuncSetAttribute(k_rows<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  k_rows<0><<<grid, kThreads, smem, (cudaStream_t)stream>>>(const_cast<double *>(W), nullptr, ld, rows, x,
                                                            nullptr, y, 0.0, 0.0, act_kind);
  HPNN_CHECK_LAUNCH("k_rows<0>");
  return 0;
}

int hpnn_dev_update_forward(double *W, double *dW, int ld, int rows, int cols, const double *delta,
                            const double *x, double *y, int update, double lr, double alpha, int act_kind,
                            void *stream) {
  (void)cols;
  if (ld % 8 || (update == 2 && !dW)) {
    set_error_msg("hpnn_dev_update_forward: bad arguments");
    return -1;
  }
  int grid = (rows + kWarps - 1) / kWarps;
  if (grid > num_sms()) grid = num_sms();
  if (grid < 1) grid = 1;
  const size_t smem = sizeof(double) * ((size_t)ld + ((rows + 7) & ~7) + row_part_doubles(rows, grid) + 8);
  if (smem > 227 * 1024) {
    set_error_msg("hpnn_dev_update_forward: layer too wide");
    return -2;
  }
  cudaStream_t s = (cudaStream_t)stream;
  if (update == 0) {
    cudaFuncSetAttribute(k_rows<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    k_rows<0><<<grid, kThreads, smem, s>>>(W, dW, ld, rows, x, delta, y, lr, alpha, act_kind);
  } else if (update == 1) {
    cudaFuncSetAttribute(k_rows<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    k_rows<1><<<grid, kThreads, smem, s>>>(W, dW, ld, rows, x, delta, y, lr, alpha, act_kind);
  } else {
    cudaFuncSetAttribute(k_rows<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    k_rows<2><<<grid, kThreads, smem, s>>>(W, dW, ld, rows, x, delta, y, lr, alpha, act_kind);
  }
  HPNN_CHECK_LAUNCH("k_rows<update>");
  return 0;
}

int hpnn_dev_output_stats(int type, int stable, const double *z, const double *target, int n, double *o,
                          double *delta_out, double *stats, void *stream) {
  const size_t smem = sizeof(double) * (((size_t)n + 7) / 8 * 8 + kRedDoubles);
  if (smem > 227 * 1024) {
    set_error_msg("hpnn_dev_output_stats: output layer too wide");
    return -2;
  }
  cudaFuncSetAttribute(k_output_stats, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  k_output_stats<<<1, kThreads, smem, (cudaStream_t)stream>>>(type, stable, z, target, n, o, delta_out, stats);
  HPNN_CHECK_LAUNCH("k_output_stats");
  return 0;
}

int hpnn_dev_backward_update(double *W, double *dW, int ld, int rows, int cols, const double *delta,
                             const double *h, double *part, int update, double lr, double alpha,
                             void *stream) {
  if (ld % 8 || (update == 2 && !dW) || (update && !h)) {
    set_error_msg("hpnn_dev_backward_update: bad arguments");
    return -1;
  }
  const size_t smem = sizeof(double) * (((size_t)rows + 7) / 8 * 8 + kRedDoubles);
  if (smem > 227 * 1024) {
    set_error_msg("hpnn_dev_backward_update: too many rows");
    return -2;
  }
  const int cb = pick_col_block(ld, num_sms());
  int grid = ld / cb;
  if (grid > num_sms()) grid = num_sms();
  cudaStream_t s = (cudaStream_t)stream;
  if (update == 0) {
    cudaFuncSetAttribute(k_cols<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    k_cols<0><<<grid, kThreads, smem, s>>>(W, dW, ld, rows, cols, cb, delta, h, part, lr, alpha);
  } else if (update == 1) {
    cudaFuncSetAttribute(k_cols<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    k_cols<1><<<grid, kThreads, smem, s>>>(W, dW, ld, rows, cols, cb, delta, h, part, lr, alpha);
  } else {
    cudaFuncSetAttribute(k_cols<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    k_cols<2><<<grid, kThreads, smem, s>>>(W, dW, ld, rows, cols, cb, delta, h, part, lr, alpha);
  }
  HPNN_CHECK_LAUNCH("k_cols");
  return 0;
}

int hpnn_dev_delta_finalize(const double *part, int n_parts, int stride, const double *h, int n, double *delta,
                            void *stream) {
  int grid = (n + 255) / 256;
  if (grid > 4 * num_sms()) grid = 4 * num_sms();
  k_delta_finalize<<<grid, 256, 0, (cudaStream_t)stream>>>(part, n_parts, stride, h, n, delta);
  HPNN_CHECK_LAUNCH("k_delta_finalize");
  return 0;
}

int hpnn_dev_forward_batch_f64(const double *W, int ld, int rows, int cols, const double *X, int ldx, int batch,
                               double *Y, int ldy, int act_kind, void *stream) {
  if ((ld & 1) || (ldx & 1) || ((uintptr_t)W & 15) || ((uintptr_t)X & 15)) {
    set_error_msg("hpnn_dev_forward_batch_f64: W and X need 16-byte aligned rows (even leading dimensions)");
    return -1;
  }
  dim3 grid((rows + GB - 1) / GB, (batch + GB - 1) / GB);
  GemmDst dst;
  dst.n = 1;
  dst.col0 = 0;
  dst.y_mc = nullptr;
  dst.y[0] = Y;
  k_forward_batch_f64<<<grid, 128, 0, (cudaStream_t)stream>>>(W, ld, rows, cols, X, ldx, batch, dst, ldy, act_kind);
  HPNN_CHECK_LAUNCH("k_forward_batch_f64");
  return 0;
}

// Row-split layer: this rank computes its `rows` neurons for the whole batch and the epilogue
// stores the tile into every rank's [batch x ldy] activation matrix at column col0 (GEMM fused
// with the all-gather of its result over NVLink peer memory).
int hpnn_dev_forward_batch_f64_tp(const double *W, int ld, int rows, int cols, const double *X, int ldx, int batch,
                                  double *const *Y_all, int n_ranks, int col0, int ldy, int act_kind, void *stream) {
  return hpnn_dev_forward_batch_f64_tp_mc(W, ld, rows, cols, X, ldx, batch, Y_all, nullptr, n_ranks, col0, ldy, act_kind, stream);
}
// Y_mc != NULL: the activation matrices are one symmetric allocation with an NVLS multicast mapping; the
// epilogue then issues ONE 16-byte multimem.st per pair of results instead of a store per rank
int hpnn_dev_forward_batch_f64_tp_mc(const double *W, int ld, int rows, int cols, const double *X, int ldx, int batch,
                                     double *const *Y_all, double *Y_mc, int n_ranks, int col0, int ldy, int act_kind,
                                     void *stream) {
  if ((ld & 1) || (ldx & 1) || ((uintptr_t)W & 15) || ((uintptr_t)X & 15) || n_ranks < 1 || n_ranks > HPNN_MAX_WORLD) {
    set_error_msg("hpnn_dev_forward_batch_f64_tp: bad arguments");
    return -1;
  }
  if (rows <= 0) return 0;
  GemmDst dst;
  dst.n = n_ranks;
  dst.col0 = col0;
  dst.y_mc = Y_mc;
  for (int p = 0; p < HPNN_MAX_WORLD; ++p) dst.y[p] = p < n_ranks ? Y_all[p] : nullptr;
  dim3 grid((rows + GB - 1) / GB, (batch + GB - 1) / GB);
  k_forward_batch_f64<<<grid, 128, 0, (cudaStream_t)stream>>>(W, ld, rows, cols, X, ldx, batch, dst, ldy, act_kind);
  HPNN_CHECK_LAUNCH("k_forward_batch_f64(tp)");
  return 0;
}

// Cross-rank barrier as a one-CTA kernel (every rank launches it on its own GPU): fence.sys,
// publish `epoch` in every peer's flag word, wait for all peers' words.  flags[p] is rank p's
// flag array (HPNN_MAX_WORLD u64) as mapped here.
struct FlagPtrs {
  unsigned long long *p[HPNN_MAX_WORLD];
};
__global__ void k_world_barrier_v(FlagPtrs f, int rank, int world, unsigned long long epoch) {
  if (threadIdx.x == 0) fence_sys();
  __syncthreads();
  if ((int)threadIdx.x < world && (int)threadIdx.x != rank) {
    st_relaxed_sys(f.p[threadIdx.x] + rank, epoch);
    // bounded: a peer that never arrives must not hang this GPU (10 s, then the barrier is given up)
    unsigned long long t0 = 0, now;
    unsigned spins = 0;
    while (ld_relaxed_sys(f.p[rank] + threadIdx.x) < epoch) {
      if ((++spins & 1023u) == 0) {
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
        if (!t0) t0 = now;
        else if (now - t0 > 10000000000ull) break;
      }
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) fence_sys();
}
int hpnn_dev_world_barrier(void *const *flags_all, int rank, int world, unsigned long long epoch, void *stream) {
  if (world < 1 || world > HPNN_MAX_WORLD) return -1;
  if (world == 1) return 0;
  FlagPtrs f;
  for (int p = 0; p < HPNN_MAX_WORLD; ++p) f.p[p] = p < world ? (unsigned long long *)flags_all[p] : nullptr;
  k_world_barrier_v<<<1, 32, 0, (cudaStream_t)stream>>>(f, rank, world, epoch);
  HPNN_CHECK_LAUNCH("k_world_barrier");
  return 0;
}

int hpnn_dev_softmax_rows(double *Z, int ldz, int batch, int n, int stable, void *stream) {
  const int wpb = 8;
  k_softmax_rows<<<(batch + wpb - 1) / wpb, wpb * 32, 0, (cudaStream_t)stream>>>(Z, ldz, batch, n, stable);
  HPNN_CHECK_LAUNCH("k_softmax_rows");
  return 0;
}

int hpnn_dev_grad_merge(double *W, long long n, double *const *grads, int n_ranks, double scale, void *stream) {
  if (n_ranks < 1 || n_ranks > HPNN_MAX_WORLD) {
    set_error_msg("hpnn_dev_grad_merge: bad rank count");
    return -1;
  }
  GradPtrs gp;
  for (int r = 0; r < HPNN_MAX_WORLD; ++r) gp.g[r] = r < n_ranks ? grads[r] : nullptr;
  long long blocks = (n / 2 + 255) / 256;
  const int cap = 8 * num_sms();
  const int grid = (int)(blocks < 1 ? 1 : (blocks > cap ? cap : blocks));
  k_grad_merge<<<grid, 256, 0, (cudaStream_t)stream>>>(W, n, gp, n_ranks, scale);
  HPNN_CHECK_LAUNCH("k_grad_merge");
  return 0;
}

/* ------------------------------ runtime helpers ---------------------------- */
int hpnn_cuda_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) {
    cudaGetLastError();
    return 0;
  }
  return n;
}
int hpnn_cuda_set_device(int dev) {
  cudaError_t e = cudaSetDevice(dev);
  if (e != cudaSuccess) {
    set_error("cudaSetDevice", e);
    return -1;
  }
  return 0;
}
void *hpnn_dev_alloc(size_t bytes) {
  void *p = nullptr;
  cudaError_t e = cudaMalloc(&p, bytes ? bytes : 8);
  if (e != cudaSuccess) {
    set_error("cudaMalloc", e);
    return nullptr;
  }
  cudaMemset(p, 0, bytes ? bytes : 8);
  return p;
}
void hpnn_dev_free(void *p) {
  if (p) cudaFree(p);
}
int hpnn_dev_memset(void *p, int v, size_t bytes, void *stream) {
  return cudaMemsetAsync(p, v, bytes, (cudaStream_t)stream) == cudaSuccess ? 0 : -1;
}
int hpnn_dev_h2d(void *dst, const void *src, size_t bytes, void *stream) {
  cudaError_t e = cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, (cudaStream_t)stream);
  if (e != cudaSuccess) set_error("cudaMemcpyAsync(H2D)", e);
  return e == cudaSuccess ? 0 : -1;
}
int hpnn_dev_d2h(void *dst, const void *src, size_t bytes, void *stream) {
  cudaError_t e = cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, (cudaStream_t)stream);
  if (e != cudaSuccess) set_error("cudaMemcpyAsync(D2H)", e);
  return e == cudaSuccess ? 0 : -1;
}
int hpnn_dev_sync(void *stream) {
  cudaError_t e = stream ? cudaStreamSynchronize((cudaStream_t)stream) : cudaDeviceSynchronize();
  if (e != cudaSuccess) set_error("synchronize", e);
  return e == cudaSuccess ? 0 : -1;
}
void *hpnn_host_alloc_pinned(size_t bytes) {
  void *p = nullptr;
  if (cudaMallocHost(&p, bytes ? bytes : 8) != cudaSuccess) {
    cudaGetLastError();
    return nullptr;
  }
  return p;
}
void hpnn_host_free_pinned(void *p) {
  if (p) cudaFreeHost(p);
}
int hpnn_ipc_export(void *dev_ptr, unsigned char handle[64]) {
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
  cudaIpcMemHandle_t h;
  cudaError_t e = cudaIpcGetMemHandle(&h, dev_ptr);
  if (e != cudaSuccess) {
    set_error("cudaIpcGetMemHandle", e);
    return -1;
  }
  memcpy(handle, &h, 64);
  return 0;
}
void *hpnn_ipc_open(const unsigned char handle[64]) {
  cudaIpcMemHandle_t h;
  memcpy(&h, handle, 64);
  void *p = nullptr;
  cudaError_t e = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
  if (e != cudaSuccess) {
    set_error("cudaIpcOpenMemHandle", e);
    return nullptr;
  }
  return p;
}
int hpnn_ipc_close(void *p) { return cudaIpcCloseMemHandle(p) == cudaSuccess ? 0 : -1; }
int hpnn_enable_peer_access(int dev, int peer) {
  int can = 0;
  cudaDeviceCanAccessPeer(&can, dev, peer);
  if (!can) return -1;
  int cur = 0;
  cudaGetDevice(&cur);
  cudaSetDevice(dev);
  cudaError_t e = cudaDeviceEnablePeerAccess(peer, 0);
  cudaSetDevice(cur);
  if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) {
    set_error("cudaDeviceEnablePeerAccess", e);
    return -1;
  }
  cudaGetLastError();
  return 0;
}

}  // extern "C"
