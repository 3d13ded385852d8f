// K-PL: the persistent per-sample trainer with DEFERRED weight updates, for BP on layers that
// are far larger than L2 (the 8192-wide benchmark shape of BASELINE.md).
//
// The plain streaming trainer (train_persistent.cu) reads every weight of a layer l >= 1 twice
// and writes it once per iteration; the write exists only to add the rank-1 term
// lr * delta_l (x) h_{l-1} (src/ann.c:2281-2467, the reference's cublasDger).  Here that term is
// kept in factored form instead: with the pending factors u_j = lr*delta_l (per owned row) and
// v_j = h_{l-1} (per column), j < m,
//
//     W_true = S + sum_j u_j (x) v_j                 (S = what is stored in memory)
//     W_true^T d      = S^T d + sum_j v_j (u_j . d)  (top-down pass, read-only)
//     W_true   x      = S x   + sum_j u_j (v_j . x)  (forward pass, read-only)
//
// so both passes stay read-only, and S is brought up to date every `depth` iterations in the
// top-down pass of that iteration (one read + write instead of `depth` of them).  The factors
// are applied in iteration order with one fma each, exactly like the immediate updates, so the
// stored weights differ from the immediate-update result only through the rounding of the two
// correction sums above (a few ulp in z).  Weight traffic per iteration drops from 3 passes to
// 2 + 1/depth for the layers l >= 1; layer 0 keeps the fused update + forward pass.
//
// Everything else (row partitioning, reduce-scatter / all-gather by peer stores, barriers,
// evaluation, convergence loop) is the protocol of k_train_sample<false>.
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "lazy_common.cuh"

namespace hpnn {

void set_error(const char *what, cudaError_t e);
void set_error_msg(const char *what);


// stage_delta of the barrier protocol (kp_common.cuh) fused with the m dot products u_j . delta:
// the factor chunk of a row travels in the same round trip as its partial sums.
__device__ __forceinline__ void stage_delta_dots(const hpnn_dev_net &n, int l, int rsplit, const double *act_next,
                                                 double *ds, const double *HU, int m, double *coef, double *red) {
  const double *part = n.arena[n.rank] + n.off_part[l];
  const int stride = n.part_stride[l];
  const int rows = n.rows[l];
  double acc[kLazyChunk];
#pragma unroll
  for (int j = 0; j < kLazyChunk; ++j) acc[j] = 0.0;
  constexpr int U = 4;  // rows per thread and round: all their loads are issued before the first use
  for (int r0 = threadIdx.x; r0 < rows; r0 += U * kThreads) {
    double s[U], hn[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int r = r0 + u * kThreads;
      s[u] = 0.0;
      hn[u] = (r < rows) ? ldg_cg(act_next + n.row0[l] + r) : 0.0;
    }
    for (int g = 0; g < n.world; ++g)
      for (int rp = 0; rp < rsplit; ++rp) {
        const double *src = part + (size_t)(g * kMaxRsplit + rp) * stride;
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const int r = r0 + u * kThreads;
          if (r < rows) s[u] += ldg_cg(src + r);
        }
      }
#pragma unroll
    for (int u = 0; u < U; u += 2) {  // two factor chunks (8 x 16 bytes) in flight
      const int ra = r0 + u * kThreads, rb = ra + kThreads;
      const double da = s[u] * dact_f(hn[u]), db = s[u + 1] * dact_f(hn[u + 1]);
      double2 qa[kLazyChunk / 2], qb[kLazyChunk / 2];
      const double2 *ca = reinterpret_cast<const double2 *>(HU + (size_t)(ra < rows ? ra : 0) * kLazyChunk);
      const double2 *cb = reinterpret_cast<const double2 *>(HU + (size_t)(rb < rows ? rb : 0) * kLazyChunk);
#pragma unroll
      for (int k = 0; k < kLazyChunk / 2; ++k) {
        qa[k] = (ra < rows && 2 * k < m) ? ldg_cg2(ca + k) : make_double2(0.0, 0.0);
        qb[k] = (rb < rows && 2 * k < m) ? ldg_cg2(cb + k) : make_double2(0.0, 0.0);
      }
      if (ra < rows) ds[ra] = da;
      if (rb < rows) ds[rb] = db;
#pragma unroll
      for (int k = 0; k < kLazyChunk / 2; ++k) {
        acc[2 * k] = fma(qa[k].x, da, acc[2 * k]);
        acc[2 * k + 1] = fma(qa[k].y, da, acc[2 * k + 1]);
        acc[2 * k] = fma(qb[k].x, db, acc[2 * k]);
        acc[2 * k + 1] = fma(qb[k].y, db, acc[2 * k + 1]);
      }
    }
  }
  if (m > 0) chunk_block_sum(acc, m, coef, red);  // ends with a barrier: ds is complete as well
  else __syncthreads();
}

// Column task that also brings S up to date: partial sums from the OLD stored weights, then
//   w += sum_{j<m} u_j[r] v_j[c]   (pending factors, oldest first)
//   w += (lr*d_r) h[c]             (this iteration's factor; lr == 0 for the final write-back)
// The weights stream through the two-batch software pipeline of col_task; the factor chunks of
// kFlushTile rows at a time are copied to shared memory with cp.async one tile ahead (two buffers
// in `ust`), so a tile boundary costs one barrier and never drains the weight loads in flight.
constexpr int kFlushTile = 512;                                  // rows; multiple of 4 * RL for CB = 8, 16, 32
constexpr int kFlushStageDoubles = 2 * kFlushTile * kLazyChunk;  // 64 KB
template <int CB, typename Sink>
__device__ __forceinline__ void col_task_flush(double *W, int ld, int row_begin, int rows, int c0, const double *ds,
                                               const double *h, double lr, const double *HU, const double *HV, int m,
                                               double *ust, double *red, Sink sink) {
  constexpr int TPR = CB / 2;
  constexpr int RL = kThreads / TPR;
  constexpr int BPT = kFlushTile / (4 * RL);  // batches of 4 row lanes per tile
  static_assert(BPT >= 1 && kFlushTile % (4 * RL) == 0, "tile must hold whole batches");
  const int cp = threadIdx.x % TPR, rl = threadIdx.x / TPR;
  const int c = c0 + 2 * cp;
  const double h0 = ldg_cg(h + c), h1 = ldg_cg(h + c + 1);
  double v0[kLazyChunk], v1[kLazyChunk];
  {
    const double2 *a = reinterpret_cast<const double2 *>(HV + (size_t)c * kLazyChunk);
    const double2 *b = a + kLazyChunk / 2;
#pragma unroll
    for (int k = 0; k < kLazyChunk / 2; ++k) {
      const bool on = 2 * k < m;
      const double2 qa = on ? ldg_cg2(a + k) : make_double2(0.0, 0.0);
      const double2 qb = on ? ldg_cg2(b + k) : make_double2(0.0, 0.0);
      v0[2 * k] = qa.x;
      v0[2 * k + 1] = (2 * k + 1 < m) ? qa.y : 0.0;
      v1[2 * k] = qb.x;
      v1[2 * k + 1] = (2 * k + 1 < m) ? qb.y : 0.0;
    }
  }
  double a0 = 0.0, a1 = 0.0;
  const size_t stride2 = (size_t)(ld >> 1);
  double2 *Wp = reinterpret_cast<double2 *>(W + c);
  const double2 zero2 = make_double2(0.0, 0.0);
  const int mc = (m + 1) >> 1;  // double2 per chunk that hold live factors
  const int nrows = rows - row_begin;
  const int ntiles = (nrows + kFlushTile - 1) / kFlushTile;
  const int nbatch = (nrows + 4 * RL - 1) / (4 * RL);
  const unsigned ust_s = (unsigned)__cvta_generic_to_shared(ust);
  auto stage = [&](int t) {  // factor chunks of tile t -> buffer t & 1 (whole chunks: stale halves are never used)
    const int t0 = row_begin + t * kFlushTile;
    const int n16 = (min(rows, t0 + kFlushTile) - t0) * (kLazyChunk / 2);
    const char *src = reinterpret_cast<const char *>(HU + (size_t)t0 * kLazyChunk);
    const unsigned dst = ust_s + (unsigned)((t & 1) * kFlushTile * kLazyChunk * sizeof(double));
    for (int i = threadIdx.x; i < n16; i += kThreads)
      asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst + 16u * i), "l"(src + 16 * (size_t)i) : "memory");
    asm volatile("cp.async.commit_group;" ::: "memory");
  };
  auto load = [&](double2 (&w)[4], int q) {
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int r = row_begin + q * 4 * RL + rl + u * RL;
      w[u] = (r < rows) ? ldg_cg2(Wp + (size_t)r * stride2) : zero2;
    }
  };
  auto consume = [&](double2 (&w)[4], int q) {
    if (q >= nbatch) return;
    const int t = q / BPT;
    if (q % BPT == 0) {  // first batch of tile t: its factors have landed, the other buffer is free again
      asm volatile("cp.async.wait_group 0;" ::: "memory");
      __syncthreads();
      if (t + 1 < ntiles) stage(t + 1);
    }
    const double2 *ut = reinterpret_cast<const double2 *>(ust) + (size_t)(t & 1) * kFlushTile * (kLazyChunk / 2);
    const int t0 = row_begin + t * kFlushTile;
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int r = row_begin + q * 4 * RL + rl + u * RL;
      if (r < rows) {
        const double d = ds[r];
        a0 = fma(w[u].x, d, a0);
        a1 = fma(w[u].y, d, a1);
        const double2 *uc = ut + (size_t)(r - t0) * (kLazyChunk / 2);
#pragma unroll
        for (int k = 0; k < kLazyChunk / 2; ++k) {
          if (k < mc) {
            const double2 f = uc[k];
            w[u].x = fma(f.x, v0[2 * k], w[u].x);
            w[u].y = fma(f.x, v1[2 * k], w[u].y);
            if (2 * k + 1 < m) {
              w[u].x = fma(f.y, v0[2 * k + 1], w[u].x);
              w[u].y = fma(f.y, v1[2 * k + 1], w[u].y);
            }
          }
        }
        const double s = lr * d;
        w[u].x = fma(s, h0, w[u].x);
        w[u].y = fma(s, h1, w[u].y);
        stg_cg2(Wp + (size_t)r * stride2, w[u]);
      }
    }
  };
  __syncthreads();  // the previous task is done with `ust`
  stage(0);
  double2 wa[4], wb[4];
  load(wa, 0);
  load(wb, 1);
  for (int q = 0; q < nbatch; q += 2) {  // nbatch is CTA-uniform, so are the barriers inside consume
    consume(wa, q);
    load(wa, q + 2);
    consume(wb, q + 1);
    load(wb, q + 3);
  }
#pragma unroll
  for (int o = 16; o >= TPR; o >>= 1) {
    a0 += __shfl_xor_sync(0xffffffffu, a0, o);
    a1 += __shfl_xor_sync(0xffffffffu, a1, o);
  }
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  __syncthreads();
  if (lane < TPR && lane < 32) {
    red[warp * CB + 2 * lane] = a0;
    red[warp * CB + 2 * lane + 1] = a1;
  }
  __syncthreads();
  if (threadIdx.x < CB) {
    double s = 0.0;
#pragma unroll
    for (int w = 0; w < kWarps; ++w) s += red[w * CB + threadIdx.x];
    sink(threadIdx.x, s);
  }
}

// all column tasks of one layer; flush != 0 selects the writing variant
template <typename SinkFactory>
__device__ __forceinline__ void phase_cols_lazy(double *W, int ld, int rows, int cb, int rsplit, const double *ds,
                                                const double *h, double lr, const double *HU, const double *HV, int m,
                                                bool flush, double *ust, double *red, SinkFactory make_sink) {
  const int ncb = ld / cb;
  const int ntasks = ncb * rsplit;
  const int part_rows = (rows + rsplit - 1) / rsplit;
  for (int t = blockIdx.x; t < ntasks; t += gridDim.x) {
    const int c0 = (t % ncb) * cb, rp = t / ncb;
    const int rb = rp * part_rows, re = min(rows, rb + part_rows);
    auto sink = make_sink(c0, rp);
    if (flush) {
      if (cb == 8) col_task_flush<8>(W, ld, rb, re, c0, ds, h, lr, HU, HV, m, ust, red, sink);
      else if (cb == 16) col_task_flush<16>(W, ld, rb, re, c0, ds, h, lr, HU, HV, m, ust, red, sink);
      else col_task_flush<32>(W, ld, rb, re, c0, ds, h, lr, HU, HV, m, ust, red, sink);
    } else {
      if (cb == 8) col_task<8, 0>(W, nullptr, ld, rb, re, c0, ds, h, false, 0.0, 0.0, red, sink);
      else if (cb == 16) col_task<16, 0>(W, nullptr, ld, rb, re, c0, ds, h, false, 0.0, 0.0, red, sink);
      else col_task<32, 0>(W, nullptr, ld, rb, re, c0, ds, h, false, 0.0, 0.0, red, sink);
    }
  }
}

// Row tasks of one layer with the pending factors folded in:
//   UPD 0: z_r = S[r,:].x + sum_j u_j[r] coef[j]          (layers >= 1, read-only)
//   UPD 1: w += lr*delta_r * x ; z_r = w.x                 (layer 0, immediate update)
// and, while the results are at hand, this CTA's share of the next layer's dot products.
template <int UPD>
__device__ __forceinline__ void phase_rows_lazy(double *W, int ld, int rows, const double *xs, const double *ds,
                                                double lr, int act_kind, const Dst &dst, int segs, double *rowpart,
                                                const double *HU, const double *coef, int m, const DotOut &dots,
                                                double *red) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int grid = gridDim.x, b = blockIdx.x;
  const int mine = b < rows ? (rows - b + grid - 1) / grid : 0;
  const int n2 = ld >> 1;
  const int seg2 = (n2 + segs - 1) / segs;
  for (int u = warp; u < mine * segs; u += kWarps) {
    const int i = u / segs, sg = u % segs;
    const int r = b + i * grid;
    const int c0 = sg * seg2, c1 = min(n2, c0 + seg2);
    double *wrow = W + (size_t)r * ld + 2 * c0;
    const double s = UPD ? lr * ds[r] : 0.0;
    const double z = c1 > c0 ? warp_row<UPD>(wrow, nullptr, xs + 2 * c0, 2 * (c1 - c0), s, 0.0, lane) : 0.0;
    if (lane == 0) rowpart[i * 8 + sg] = z;
  }
  __syncthreads();
  const bool want_dots = dots.p[0] != nullptr && m > 0 && mine > 0;
  double acc[kLazyChunk];
#pragma unroll
  for (int j = 0; j < kLazyChunk; ++j) acc[j] = 0.0;
  for (int i = threadIdx.x; i < mine; i += blockDim.x) {
    const int r = b + i * grid;
    double z = 0.0;
    for (int sg = 0; sg < segs; ++sg) z += rowpart[i * 8 + sg];
    if (!UPD && m > 0) z += chunk_apply(HU + (size_t)r * kLazyChunk, coef, m);
    const double y = act_kind == 0 ? act_f(z) : z;
    dst_store(dst, r, y);
    if (want_dots) chunk_fma(dots.HV + (size_t)(dots.row0 + r) * kLazyChunk, y, m, acc);
  }
  if (want_dots) {  // uniform per CTA
    // NOTE: coef is still needed by nobody at this point (all rows of this CTA are finished)
    __shared__ double mine_dots[kLazyChunk];
    chunk_block_sum(acc, m, mine_dots, red);
    if (threadIdx.x < kLazyChunk)
      for (int q = 0; q < dots.n; ++q) stg_cg(dots.p[q] + threadIdx.x, mine_dots[threadIdx.x]);
  }
}

__global__ void __launch_bounds__(kThreads, 1) k_train_sample_lazy(const __grid_constant__ KP p) {
  extern __shared__ double smem[];
  watch_init(p);
  if (p.fault == 1) return;
  rows_table_init(p.net);  // injected fault: this rank never shows up
  // xs doubles as the factor staging area of the write-back pass (the top-down phases read their
  // input vector from global memory), so it is at least kFlushStageDoubles long.  Keeping the
  // footprint at the plain trainer's size matters: 64 KB more shared memory shrank L1 enough to
  // cost the forward phases 13% of their bandwidth.
  double *xs = smem;
  double *ds = xs + max(p.max_ld, kFlushStageDoubles);
  double *red = ds + p.max_rows;
  double *rowpart = red + kRedDoubles;
  double *ust = xs;
  __shared__ double coef[kLazyChunk];

  const hpnn_dev_net &n = p.net;
  const int NL = n.n_layers;
  const int n_out = n.dims[NL];
  double *A = n.arena[n.rank];
  const bool snn = p.type == HPNN_TYPE_SNN;
  const int grid = gridDim.x;
  const int depth = p.lazy_depth;

  Barrier bar;
  bar.init(n, gridDim.x, blockIdx.x);
  int trace_n = 0;
  auto trace = [&](int tag) {
    if (p.trace && blockIdx.x == 0 && threadIdx.x == 0 && trace_n < p.trace_cap) {
      p.trace[2 * trace_n] = tag;
      p.trace[2 * trace_n + 1] = clock64();
      ++trace_n;
    }
  };
  int cur = 0;
  auto act_of = [&](int rank, int i, int buf) -> double * {
    return n.arena[rank] + n.off_act[i] + ((i == 0) ? 0 : (size_t)buf * pad8(n.dims[i]));
  };
  // one forward phase: layer 0 with its immediate update (u == 1) or read-only; m pending factors
  auto forward = [&](int l, int src_buf, int dst_buf, int u, int m) {
    stage_vec(xs, act_of(n.rank, l, src_buf), pad8(n.dims[l]));
    __syncthreads();
    if (l >= 1 && m > 0) gather_dots(n, l, m, coef, red, gridDim.x);  // coef[j] = v_j . x
    Dst dst;
    dst.n = n.world;
    dst.tag = 0ull;
    dst.mc = nullptr;
    for (int q = 0; q < n.world; ++q) dst.p[q] = act_of(q, l + 1, dst_buf) + (size_t)n.row0[l];
    DotOut dots;
    dots.n = n.world;
    dots.tag = 0ull;
    dots.mc = nullptr;
    dots.p[0] = nullptr;
    if (l + 1 < NL && m > 0) {
      for (int q = 0; q < n.world; ++q)
        dots.p[q] = n.arena[q] + n.off_hist_dot[l + 1] + (size_t)(n.rank * kLazyDotSlots + blockIdx.x) * kLazyChunk;
      dots.HV = A + n.off_hist_v[l + 1];
      dots.row0 = n.row0[l];
    }
    const int kind = (snn && l == NL - 1) ? 1 : 0;
    const int segs = pick_row_segments(n.rows[l], n.ld[l], gridDim.x);
    const double *HU = l >= 1 ? A + n.off_hist_u[l] : nullptr;
    if (u) phase_rows_lazy<1>(n.W[l], n.ld[l], n.rows[l], xs, ds, p.lr, kind, dst, segs, rowpart, HU, coef, m, dots, red);
    else phase_rows_lazy<0>(n.W[l], n.ld[l], n.rows[l], xs, ds, p.lr, kind, dst, segs, rowpart, HU, coef, l >= 1 ? m : 0, dots, red);
  };

  for (int l = 0; l < NL; ++l) {
    forward(l, cur, cur, 0, 0);
    bar.sync_all(n, (int)blockIdx.x < n.rows[l]);
  }
  const double *target = A + n.off_target;
  const int last = NL - 1;
  OutStats st = output_stats(p.type, p.stable, act_of(n.rank, NL, cur), target, n_out, n.row0[last], n.rows[last], ds,
                             nullptr, red);
  double E = st.err;
  const double init_err = E;
  int iter = 0, ok = (st.argmax == st.target), first_ok = 0;
  double dE = 0.0;
  int m = 0;  // pending factors of every layer l >= 1 (identical on all CTAs and ranks)

  for (;;) {
    ++iter;
    trace(1);
    const unsigned long long abort_seen = watch_peek();
    const bool flush = (m + 1 >= depth);
    // ---- top-down: partial W_true^T.delta (read-only unless this is a write-back iteration)
    for (int l = last; l >= 1; --l) {
      double *HU = A + n.off_hist_u[l];
      double *HV = A + n.off_hist_v[l];
      // delta_l into ds, coef[j] = u_j . delta_l
      if (l != last)
        stage_delta_dots(n, l, layer_plan(n, l + 1, grid).rsplit, act_of(n.rank, l + 1, cur), ds, HU, m, coef, red);
      else if (m > 0)
        chunk_dots(HU, ds, n.rows[l], m, coef, red);
      const double *h = act_of(n.rank, l, cur);
      if (!flush) {
        // remember this iteration's factors: u_m = lr*delta (owned rows), v_m = h (all columns)
        const int gt = blockIdx.x * kThreads + threadIdx.x;
        for (int r = gt; r < n.rows[l]; r += grid * kThreads) stg_cg(HU + (size_t)r * kLazyChunk + m, p.lr * ds[r]);
        for (int c = gt; c < n.ld[l]; c += grid * kThreads) stg_cg(HV + (size_t)c * kLazyChunk + m, ldg_cg(h + c));
      }
      const int cols = n.dims[l];
      const ColPlan plan = layer_plan(n, l, grid);
      const long long off_prev = n.off_part[l - 1];
      const int stride_prev = n.part_stride[l - 1];
      const int mm = m;
      auto make_sink = [&](int c0, int rp) {
        return [=, &n](int i, double v) {
          const int col = c0 + i;
          if (col < cols) {
            if (mm > 0 && rp == 0) v += chunk_apply(HV + (size_t)col * kLazyChunk, coef, mm);
            const int own = owner_of_row(col, cols, n.world);
            const size_t slot = (size_t)(n.rank * kMaxRsplit + rp) * stride_prev + (col - row0_of(own, cols, n.world));
            stg_cg(n.arena[own] + off_prev + slot, v);
          }
        };
      };
      phase_cols_lazy(n.W[l], n.ld[l], n.rows[l], plan.cb, plan.rsplit, ds, h, p.lr, HU, HV, m, flush, ust, red, make_sink);
      trace(100 + l);
      bar.sync_all(n, (int)blockIdx.x < (n.ld[l] / plan.cb) * plan.rsplit);
      trace(200 + l);
    }
    m = flush ? 0 : m + 1;
    // ---- bottom-up: layer 0 update fused with its forward, then the read-only layers
    stage_delta(n, 0, layer_plan(n, 1, grid).rsplit, act_of(n.rank, 1, cur), ds, false, 0ull);
    forward(0, 0, cur ^ 1, 1, m);
    trace(300);
    bar.sync_all(n, (int)blockIdx.x < n.rows[0]);
    trace(400);
    for (int l = 1; l < NL; ++l) {
      forward(l, cur ^ 1, cur ^ 1, 0, m);
      trace(300 + l);
      bar.sync_all(n, (int)blockIdx.x < n.rows[l]);
      trace(400 + l);
    }
    cur ^= 1;
    watch_publish(abort_seen);
    st = output_stats(p.type, p.stable, act_of(n.rank, NL, cur), target, n_out, n.row0[last], n.rows[last], ds, nullptr,
                      red);
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
  // ---- write the pending factors back so that memory holds W_true again
  if (m > 0) {
    auto no_sink = [](int, int) { return [](int, double) {}; };
    __syncthreads();
    for (int r = threadIdx.x; r < p.max_rows; r += kThreads) ds[r] = 0.0;  // lr == 0 and the sums are dropped
    __syncthreads();
    for (int l = 1; l < NL; ++l) {
      const ColPlan plan = layer_plan(n, l, grid);
      phase_cols_lazy(n.W[l], n.ld[l], n.rows[l], plan.cb, plan.rsplit, ds, act_of(n.rank, l, cur), 0.0,
                      A + n.off_hist_u[l], A + n.off_hist_v[l], m, true, ust, red, no_sink);
    }
  }
  if (blockIdx.x == 0)
    output_stats(p.type, p.stable, act_of(n.rank, NL, cur), target, n_out, 0, 0, nullptr, A + n.off_out, red);
  bar.finish();
  if (blockIdx.x == 0 && threadIdx.x == 0) n.local_sync[160] = ld_relaxed_gpu(n.local_sync + 160) + 1u;
  if (blockIdx.x == 0 && threadIdx.x == 0 && p.result) {
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

// write-back period: -1 = take HPNN_LAZY_DEPTH (default 8), 0 = variant disabled, 2..8
static int g_lazy_depth = -1;
long long g_lazy_min_bytes = -1;  // -1 = kLazyMinBytes; read by hpnn_dev_arena_layout

int lazy_depth_setting() {
  static const int depth_env = [] {
    const char *off = getenv("HPNN_LAZY");
    if (off && !strcmp(off, "0")) return 0;
    const char *e = getenv("HPNN_LAZY_DEPTH");
    const int d = e ? atoi(e) : kLazyChunk;
    return d < 2 ? 2 : (d > kLazyChunk ? kLazyChunk : d);
  }();
  return g_lazy_depth >= 0 ? g_lazy_depth : depth_env;
}

// 0 = launched, < 0 = error, 1 = not eligible.  HPNN_LAZY=0 disables, HPNN_LAZY_DEPTH=2..8 sets
// the write-back period (default 8); hpnn_dev_set_lazy overrides both.
int launch_lazy_if_eligible(const KP &p_in, int grid, size_t smem_base, void *stream) {
  const size_t smem =
      smem_base + sizeof(double) * (p_in.max_ld < kFlushStageDoubles ? kFlushStageDoubles - p_in.max_ld : 0);
  const int depth = lazy_depth_setting();
  const hpnn_dev_net &n = p_in.net;
  if (depth < 2 || smem > 227 * 1024 || grid > kLazyDotSlots || p_in.mode != 0 || p_in.train != HPNN_TRAIN_BP || p_in.tagged || n.hist_depth < 2 || n.n_layers < 2)
    return 1;
  KP p = p_in;
  p.lazy_depth = depth < n.hist_depth ? depth : n.hist_depth;
  cudaError_t e = cudaFuncSetAttribute((const void *)k_train_sample_lazy, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) {
    set_error("cudaFuncSetAttribute(k_train_sample_lazy)", e);
    return -1;
  }
  void *args[] = {(void *)&p};
  e = cudaLaunchCooperativeKernel((const void *)k_train_sample_lazy, dim3(grid), dim3(kThreads), args, smem,
                                  (cudaStream_t)stream);
  if (e != cudaSuccess) {
    set_error("cudaLaunchCooperativeKernel(k_train_sample_lazy)", e);
    return -1;
  }
  return 0;
}

}  // namespace hpnn

extern "C" void hpnn_dev_set_lazy(int depth, long long min_bytes) {
  hpnn::g_lazy_depth = depth < 0 ? -1 : (depth > hpnn::kLazyChunk ? hpnn::kLazyChunk : (depth == 1 ? 2 : depth));
  hpnn::g_lazy_min_bytes = min_bytes;
}
