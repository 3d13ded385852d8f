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
  dst.p[0] = y;
  phase_rows<UPD>(W, dW, ld, rows, xs, ds, lr, alpha, act_kind, dst, pick_row_segments(rows, ld, gridDim.x), rowpart);
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
  phase_cols<UPD>(W, dW, ld, rows, cb, 1, ds, h, lr, alpha, red, make_sink);
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
// Y[b, r] = f( sum_c X[b, c] * W[r, c] ): classic 64x64x16 shared-memory tiling with
// a 4x4 register block per thread (256 threads).  FP64 FMA pipe; the reference has
// no batched path at all (batch = 1), so this is measured against
// B x its single-sample time.
constexpr int GB = 64, GK = 16;
__global__ void __launch_bounds__(256)
    k_forward_batch_f64(const double *__restrict__ W, int ld, int rows, int cols, const double *__restrict__ X,
                        int ldx, int batch, double *__restrict__ Y, int ldy, int act_kind) {
  __shared__ double Xs[GK][GB + 1];
  __shared__ double Ws[GK][GB + 1];
  const int b0 = blockIdx.y * GB, r0 = blockIdx.x * GB;
  const int tx = threadIdx.x % 16, ty = threadIdx.x / 16;  // tx -> rows (N), ty -> batch
  double acc[4][4] = {};
  for (int k0 = 0; k0 < cols; k0 += GK) {
    // 64x16 tiles of X and W: 1024 elements each, 4 per thread
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int e = threadIdx.x + 256 * i;
      const int kk = e % GK, mm = e / GK;
      const int c = k0 + kk;
      Xs[kk][mm] = (b0 + mm < batch && c < cols) ? X[(size_t)(b0 + mm) * ldx + c] : 0.0;
      Ws[kk][mm] = (r0 + mm < rows && c < cols) ? W[(size_t)(r0 + mm) * ld + c] : 0.0;
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < GK; ++kk) {
      double xv[4], wv[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        xv[i] = Xs[kk][ty + 16 * i];
        wv[i] = Ws[kk][tx + 16 * i];
      }
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fma(xv[i], wv[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int b = b0 + ty + 16 * i;
    if (b >= batch) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int r = r0 + tx + 16 * j;
      if (r < rows) Y[(size_t)b * ldy + r] = act_kind == 0 ? act_f(acc[i][j]) : acc[i][j];
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
  dim3 grid((rows + GB - 1) / GB, (batch + GB - 1) / GB);
  k_forward_batch_f64<<<grid, 256, 0, (cudaStream_t)stream>>>(W, ld, rows, cols, X, ldx, batch, Y, ldy, act_kind);
  HPNN_CHECK_LAUNCH("k_forward_batch_f64");
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
