// K-PR: the shared-memory-resident variant of the persistent per-sample trainer.
//
// For the networks libhpnn is actually used with (tutorial shapes: 784-300-10, 784-1000-1000-10,
// 851-230-230) all weights AND momenta fit the 148 x 227 KB of shared memory of a B200.  Every
// CTA then owns a fixed set of rows of EVERY layer (round-robin over the grid, continued from
// layer to layer so that the load is even), loads them once per sample, keeps them on chip for
// the thousands of iterations of the convergence loop, and writes them back at the end.  Per
// iteration only vectors and per-CTA partial sums touch L2:
//
//   top-down  l = NL-1..1 : threads own column pairs; for each owned row: p[c] += w[r][c]*delta[r]
//                           (old w), then the BP/BPM update of w[r][c] in shared memory;
//                           p is written to this CTA's slot of a partial buffer
//                           each CTA sums, in fixed CTA order, the partials of the rows it owns
//                           in layer l-1 and multiplies by dact(h)  -> its own delta values
//   bottom-up l = 0       : warp per owned row: update fused with the dot product
//             l = 1..NL-1 : warp per owned row: dot product with the gathered input
//   There are NO grid barriers: every hand-over is a tagged 16-byte pair (see st_pair below).
//   evaluate  (every CTA, redundantly) as in the streaming kernel.
//
// Same arithmetic per element as the streaming kernel; the only difference is the grouping of
// the partial sums of W^T.delta (per CTA instead of per column block), i.e. round-off level.
// Single rank only: a network this small gains nothing from several GPUs (bench: SNN
// 784-1000-1000-10 is 2-3x slower on 2 GPUs than on 1 because every barrier crosses NVLink).
#include <stdio.h>
#include <stdlib.h>

#include "kp_common.cuh"

namespace hpnn {

void set_error(const char *what, cudaError_t e);
void set_error_msg(const char *what);

struct RP {
  KP kp;
  int shift[HPNN_MAX_LAYERS];   // first row of layer l goes to CTA shift[l]
  int max_own_doubles;          // capacity of the per-CTA weight store (doubles)
  int max_own_rows;             // rows owned by a CTA, all layers
};

// rows of layer l owned by CTA b: first_row + i*grid, i < count
__device__ __host__ inline int rp_first_row(int b, int shift, int grid) { return (b - shift + grid) % grid; }
__device__ __host__ inline int rp_count(int b, int shift, int grid, int rows) {
  const int f = rp_first_row(b, shift, grid);
  return f < rows ? (rows - f + grid - 1) / grid : 0;
}

template <int UPD>
__device__ __forceinline__ double smem_row(double *wrow, double *mrow, const double *xs, int ld, double s, double alpha,
                                           int lane) {
  double2 *w2 = reinterpret_cast<double2 *>(wrow);
  double2 *m2 = reinterpret_cast<double2 *>(mrow);
  const double2 *x2 = reinterpret_cast<const double2 *>(xs);
  double a0 = 0.0, a1 = 0.0;
  for (int i = lane; i < (ld >> 1); i += 32) {
    double2 w = w2[i];
    const double2 x = x2[i];
    if (UPD == 1) {
      w.x = fma(s, x.x, w.x);
      w.y = fma(s, x.y, w.y);
      w2[i] = w;
    } else if (UPD == 2) {
      double2 m = m2[i];
      m.x = fma(s, x.x, m.x);
      m.y = fma(s, x.y, m.y);
      w.x += m.x;
      w.y += m.y;
      w2[i] = w;
      m.x *= alpha;
      m.y *= alpha;
      m2[i] = m;
    }
    a0 = fma(w.x, x.x, a0);
    a1 = fma(w.y, x.y, a1);
  }
  double a = a0 + a1;
#pragma unroll
  for (int o = 16; o; o >>= 1) a += __shfl_xor_sync(0xffffffffu, a, o);
  return a;
}

__global__ void __launch_bounds__(kThreads, 1) k_train_sample_resident(const __grid_constant__ RP rp) {
  extern __shared__ double smem[];
  const KP &p = rp.kp;
  watch_init(p);
  const hpnn_dev_net &n = p.net;
  const int NL = n.n_layers, grid = gridDim.x, b = blockIdx.x;
  const int n_out = n.dims[NL];
  const bool snn = p.type == HPNN_TYPE_SNN;
  const int upd = (p.mode == 0) ? ((p.train == HPNN_TRAIN_BPM) ? 2 : 1) : 0;
  double *A = n.arena[0];

  // ---- shared-memory carve-up
  double *Wsm = smem;                                         // owned rows, all layers
  double *Msm = Wsm + rp.max_own_doubles;                     // momentum (BPM only)
  double *xs = Msm + (upd == 2 ? rp.max_own_doubles : 0);     // staged input of the current layer
  double *x0 = xs + p.max_ld;                                 // the sample (never changes)
  double *dout = x0 + pad8(n.dims[0]);                        // output deltas (all of them)
  double *zout = dout + pad8(n_out);                          // gathered output vector
  double *down = zout + pad8(n_out);                          // deltas of owned rows, per layer
  double *hown = down + rp.max_own_rows;                      // outputs of owned rows, per layer
  double *red = hown + rp.max_own_rows;
  __shared__ int s_first[HPNN_MAX_LAYERS], s_cnt[HPNN_MAX_LAYERS], s_woff[HPNN_MAX_LAYERS], s_doff[HPNN_MAX_LAYERS];
  if (threadIdx.x == 0) {
    int woff = 0, doff = 0;
    for (int l = 0; l < NL; ++l) {
      s_first[l] = rp_first_row(b, rp.shift[l], grid);
      s_cnt[l] = rp_count(b, rp.shift[l], grid, n.rows[l]);
      s_woff[l] = woff;
      s_doff[l] = doff;
      woff += s_cnt[l] * n.ld[l];
      doff += s_cnt[l];
    }
  }
  __syncthreads();

  int trace_n = 0;
  auto trace = [&](int tg) {
    if (p.trace && blockIdx.x == 0 && threadIdx.x == 0 && trace_n < p.trace_cap) {
      p.trace[2 * trace_n] = tg;
      p.trace[2 * trace_n + 1] = clock64();
      ++trace_n;
    }
  };
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  // launch id: bumped by CTA 0 when the kernel ends; every CTA has read it long before that
  // (nobody can finish without every other CTA's data)
  const unsigned long long launch = (unsigned long long)ld_relaxed_gpu(n.local_sync + 160) + 1ull;
  auto tag_of = [&](int iteration) { return (launch << 32) | (unsigned long long)(unsigned)iteration; };
  // tagged vectors live where the streaming kernel keeps its two activation buffers
  auto act_pairs = [&](int i) -> double * { return A + n.off_act[i]; };

  // ---- load the owned rows (and clear the momentum), stage the sample
  for (int l = 0; l < NL; ++l) {
    const int ld2 = n.ld[l] >> 1;
    for (int i = 0; i < s_cnt[l]; ++i) {
      const double2 *src = reinterpret_cast<const double2 *>(n.W[l] + (size_t)(s_first[l] + i * grid) * n.ld[l]);
      double2 *dst = reinterpret_cast<double2 *>(Wsm + s_woff[l] + i * n.ld[l]);
      double2 *mom = reinterpret_cast<double2 *>(Msm + s_woff[l] + i * n.ld[l]);
      for (int c = threadIdx.x; c < ld2; c += blockDim.x) {
        dst[c] = ldg_cg2(src + c);
        if (upd == 2) mom[c] = make_double2(0.0, 0.0);
      }
    }
  }
  stage_vec(x0, A + n.off_act[0], n.ld[0]);
  __syncthreads();

  // forward of layer l for the owned rows; u: 0 plain, 1 BP, 2 BPM (layer 0 only).
  // The result is published as a tagged pair and kept in `hown` (this CTA needs its own
  // outputs again for dact() in the next top-down pass).
  auto forward_layer = [&](int l, const double *xin, unsigned long long tag, int u) {
    const int kind = (snn && l == NL - 1) ? 1 : 0;
    double *y = act_pairs(l + 1);
    const int cnt = s_cnt[l];
    // split every row over `parts` warps (fixed split points, partials added in part order)
    const int parts = cnt > 0 ? min(4, max(1, kWarps / cnt)) : 1;
    const int seg = (((n.ld[l] >> 1) + parts - 1) / parts + 1) & ~1;  // double2 per part, even
    for (int job = warp; job < cnt * parts; job += kWarps) {
      const int i = job / parts, part = job % parts;
      const int c0 = part * seg, c1 = min(n.ld[l] >> 1, c0 + seg);
      double *wrow = Wsm + s_woff[l] + i * n.ld[l] + 2 * c0;
      double *mrow = Msm + s_woff[l] + i * n.ld[l] + 2 * c0;
      const double s = u ? p.lr * down[s_doff[l] + i] : 0.0;
      const int len = max(0, c1 - c0) * 2;
      double z;
      if (u == 0) z = smem_row<0>(wrow, mrow, xin + 2 * c0, len, s, p.alpha, lane);
      else if (u == 1) z = smem_row<1>(wrow, mrow, xin + 2 * c0, len, s, p.alpha, lane);
      else z = smem_row<2>(wrow, mrow, xin + 2 * c0, len, s, p.alpha, lane);
      if (lane == 0) red[64 + job] = z;
    }
    __syncthreads();
    for (int i = threadIdx.x; i < cnt; i += blockDim.x) {
      double z = 0.0;
      for (int q = 0; q < parts; ++q) z += red[64 + i * parts + q];
      const double v = kind == 0 ? act_f(z) : z;
      hown[s_doff[l] + i] = v;
      st_pair<false>(y, s_first[l] + i * grid, v, tag);
    }
  };
  // full forward sweep with tag `tag`; layer 0 optionally fused with its update
  auto forward_all = [&](unsigned long long tag, int u0) {
    forward_layer(0, x0, tag, u0);
    trace(300);
    for (int l = 1; l < NL; ++l) {
      __syncthreads();  // xs is free again
      stage_pairs<false>(xs, act_pairs(l), n.dims[l], n.ld[l], tag);
      __syncthreads();
      trace(400 + l);
      forward_layer(l, xs, tag, 0);
      trace(300 + l);
    }
    __syncthreads();
    stage_pairs<false>(zout, act_pairs(NL), n_out, pad8(n_out), tag);  // the gathered output vector
    __syncthreads();
    trace(500);
  };

  forward_all(tag_of(0), 0);
  const double *target = A + n.off_target;
  const int last = NL - 1;
  OutStats st = output_stats<true>(p.type, p.stable, zout, target, n_out, 0, n_out, dout,
                             (p.mode == 1 && b == 0) ? A + n.off_out : nullptr, red);
  double E = st.err;
  const double init_err = E;
  int iter = 0, ok = (st.argmax == st.target), first_ok = 0;
  double dE = 0.0;

  if (p.mode == 0) {
    for (;;) {
      ++iter;
      trace(1);
      const unsigned long long abort_seen = watch_peek();
      const unsigned long long tag = tag_of(iter);
      // owned deltas of the output layer
      for (int i = threadIdx.x; i < s_cnt[last]; i += blockDim.x) down[s_doff[last] + i] = dout[s_first[last] + i * grid];
      __syncthreads();
      // ---- top-down
      for (int l = last; l >= 1; --l) {
        const int ld2 = n.ld[l] >> 1;
        // h = input vector of layer l as produced by the previous forward sweep (tag iter-1).
        // EVERY thread's h must be in before this CTA publishes a single partial: a fast
        // consumer may otherwise start the next forward sweep and overwrite act[l] under us.
        stage_pairs<false>(xs, act_pairs(l), n.dims[l], n.ld[l], tag - 1);
        __syncthreads();
        const double2 *h2 = reinterpret_cast<const double2 *>(xs);
        double *slot = A + n.off_cta_part[l] + (size_t)b * 2 * n.ld[l];
        for (int c = threadIdx.x; c < ld2; c += blockDim.x) {
          const double2 h = h2[c];
          double2 acc = make_double2(0.0, 0.0);
          for (int i = 0; i < s_cnt[l]; ++i) {
            double2 *wp = reinterpret_cast<double2 *>(Wsm + s_woff[l] + i * n.ld[l]) + c;
            double2 w = *wp;
            const double d = down[s_doff[l] + i];
            acc.x = fma(w.x, d, acc.x);
            acc.y = fma(w.y, d, acc.y);
            const double s = p.lr * d;
            if (upd == 1) {
              w.x = fma(s, h.x, w.x);
              w.y = fma(s, h.y, w.y);
              *wp = w;
            } else {
              double2 *mp = reinterpret_cast<double2 *>(Msm + s_woff[l] + i * n.ld[l]) + c;
              double2 m = *mp;
              m.x = fma(s, h.x, m.x);
              m.y = fma(s, h.y, m.y);
              w.x += m.x;
              w.y += m.y;
              *wp = w;
              m.x *= p.alpha;
              m.y *= p.alpha;
              *mp = m;
            }
          }
          st_pair<false>(slot, 2 * c, acc.x, tag);
          st_pair<false>(slot, 2 * c + 1, acc.y, tag);
        }
        trace(100 + l);
        // deltas of the rows this CTA owns in layer l-1: ordered sum over the CTA slots
        const double *part = A + n.off_cta_part[l];
        for (int i = warp; i < s_cnt[l - 1]; i += kWarps) {
          const int r = s_first[l - 1] + i * grid;
          double s = 0.0;
          for (int q0 = lane; q0 < grid; q0 += 32 * 6) {  // up to 6 slots in flight per lane (192 CTAs)
            const double *addr[6];
            bool want[6];
            double v[6];
#pragma unroll
            for (int u = 0; u < 6; ++u) {
              const int q = q0 + 32 * u;
              want[u] = q < grid;
              addr[u] = part + ((size_t)(want[u] ? q : 0) * n.ld[l] + r) * 2;
            }
            ld_pairs_wait<6, false>(addr, want, tag, v);
#pragma unroll
            for (int u = 0; u < 6; ++u) s += v[u];
          }
#pragma unroll
          for (int o = 16; o; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
          if (lane == 0) down[s_doff[l - 1] + i] = s * dact_f(hown[s_doff[l - 1] + i]);
        }
        __syncthreads();
        trace(200 + l);
      }
      // ---- bottom-up: layer 0 update fused with its forward, then the rest
      forward_all(tag, upd);
      watch_publish(abort_seen);
      st = output_stats<true>(p.type, p.stable, zout, target, n_out, 0, n_out, dout, nullptr, red);
      trace(2);
      if (s_watch.exit_now) break;  // the job was abandoned (a wait ran out of time)
      dE = E - st.err;
      E = st.err;
      ok = (st.argmax == st.target);
      if (iter == 1) first_ok = ok;
      if (iter > p.max_iter) break;
      ok = ok && (iter > p.min_iter);
      if (!(dE > p.thr || !ok)) break;
    }
    if (b == 0) output_stats<true>(p.type, p.stable, zout, target, n_out, 0, 0, nullptr, A + n.off_out, red);
    // ---- write the trained rows back
    __syncthreads();
    for (int l = 0; l < NL; ++l) {
      const int ld2 = n.ld[l] >> 1;
      for (int i = 0; i < s_cnt[l]; ++i) {
        double2 *dst = reinterpret_cast<double2 *>(n.W[l] + (size_t)(s_first[l] + i * grid) * n.ld[l]);
        const double2 *src = reinterpret_cast<const double2 *>(Wsm + s_woff[l] + i * n.ld[l]);
        for (int c = threadIdx.x; c < ld2; c += blockDim.x) stg_cg2(dst + c, src[c]);
      }
    }
  }
  if (b == 0 && threadIdx.x == 0) {
    n.local_sync[160] = (unsigned)launch;
    if (p.result) {
      hpnn_train_result r;
      r.init_err = init_err;
      r.final_de = dE;
      r.last_err = E;
      r.n_iter = iter;
      r.first_ok = first_ok;
      r.success = ok;
      r.status = watch_status();
      *p.result = r;
    }
  }
}

// 0 = launched, < 0 = error, 1 = not eligible
int launch_resident_if_eligible(const KP &p, int grid, void *stream) {
  static const bool disabled = getenv("HPNN_NO_RESIDENT") && atoi(getenv("HPNN_NO_RESIDENT")) != 0;
  const hpnn_dev_net &n = p.net;
  if (disabled || n.world != 1 || n.cta_slots < grid || grid < 1) return 1;
  const int NL = n.n_layers, n_out = n.dims[NL];
  if (n_out > 4096) return 1;
  RP rp;
  rp.kp = p;
  int shift = 0;
  for (int l = 0; l < NL; ++l) {
    rp.shift[l] = shift % grid;
    shift += n.rows[l];
  }
  long long max_doubles = 0;
  int max_rows = 0;
  for (int b = 0; b < grid; ++b) {
    long long d = 0;
    int r = 0;
    for (int l = 0; l < NL; ++l) {
      const int c = rp_count(b, rp.shift[l], grid, n.rows[l]);
      d += (long long)c * n.ld[l];
      r += c;
    }
    max_doubles = d > max_doubles ? d : max_doubles;
    max_rows = r > max_rows ? r : max_rows;
  }
  const bool bpm = p.mode == 0 && p.train == HPNN_TRAIN_BPM;
  rp.max_own_doubles = (int)((max_doubles + 7) & ~7LL);
  rp.max_own_rows = (max_rows + 7) & ~7;
  const long long doubles = (long long)rp.max_own_doubles * (bpm ? 2 : 1) + p.max_ld + ((n.dims[0] + 7) & ~7) +
                            2 * ((n_out + 7) & ~7) + 2 * rp.max_own_rows + kRedDoubles;
  const size_t smem = sizeof(double) * (size_t)doubles;
  if (smem > 226 * 1024) return 1;
  cudaError_t e = cudaFuncSetAttribute(k_train_sample_resident, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) {
    set_error("cudaFuncSetAttribute(k_train_sample_resident)", e);
    return -1;
  }
  void *args[] = {(void *)&rp};
  e = cudaLaunchCooperativeKernel((const void *)k_train_sample_resident, dim3(grid), dim3(kThreads), args, smem,
                                  (cudaStream_t)stream);
  if (e != cudaSuccess) {
    set_error("cudaLaunchCooperativeKernel(k_train_sample_resident)", e);
    return -1;
  }
  return 0;
}

}  // namespace hpnn
