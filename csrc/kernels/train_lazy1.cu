// K-PL1: deferred-update BP trainer organised around row panels (one GPU, or a row-split job).
//
// Same arithmetic as train_lazy.cu (W_true = S + sum_j u_j (x) v_j, write-back every `depth`
// iterations; src/ann.c:2281-2467 is the loop being reproduced), different work split: CTA b
// owns one contiguous panel of this rank's rows of every layer for the whole sample, in every phase.
//
//   forward   : wide layers: the CTA streams its panel row by row, every thread holding the x
//               values of a few column pairs (panel_z_stream); narrow layers: one warp per
//               (row, segment) as in the other trainers.
//   top-down  : the CTA streams its panel once, every thread owning a few column pairs:
//               P_b[c] = sum_{r in panel} W[r,c] delta[r].  No cross-thread reduction, no CTA-wide
//               barrier inside the pass, every row is read as one contiguous 32 KB run, so this
//               pass runs at 6.9 TB/s (the column tasks of K-P / K-PL reach ~5.2 TB/s because each
//               task ends in a block-wide reduction).  P_b goes to a per-CTA slot in global memory
//               (148 x 64 KB, L2 resident).
//   [barrier]
//   reduce    : one rank: CTA b adds the 148 partial vectors for the columns that are ITS rows of
//               the layer below and turns them into that layer's deltas.  Several ranks: the CTA
//               adds a column slice, sends the sums to the rank owning those rows (peer stores),
//               and after a cross-rank barrier the owning CTA adds one value per rank.
//               Either way deltas and u factors never leave the owning CTA: they live in shared
//               memory, the per-CTA staging of the full delta vector (8192 x 3 loads per CTA and
//               layer) is gone, and the write-back pass reads its u factors as shared-memory
//               broadcasts.
//
// The v factors (one 8-chunk per column), the partial dot products v_j . x and the partial
// u_j . delta travel through global memory exactly as in train_lazy.cu.  Layer 0 is read-only as
// well: its input never changes within a sample, so its pending updates are one factor pair.
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "lazy_common.cuh"

namespace hpnn {

void set_error(const char *what, cudaError_t e);
void set_error_msg(const char *what);
int lazy_depth_setting();

struct Panel {
  int r0, cnt;
};
// contiguous balanced split of `rows` over the first min(grid, rows) CTAs
__host__ __device__ inline Panel panel_of(int rows, int grid, int b) {
  const int nb = grid < rows ? grid : rows;
  Panel p;
  if (b >= nb) {
    p.r0 = rows;
    p.cnt = 0;
    return p;
  }
  p.r0 = (int)((long long)b * rows / nb);
  p.cnt = (int)((long long)(b + 1) * rows / nb) - p.r0;
  return p;
}
__host__ __device__ inline int panel_max(int rows, int grid) {
  const int nb = grid < rows ? grid : rows;
  return nb > 0 ? (rows + nb - 1) / nb : 0;
}

// Bulk L2 prefetch of one weight row.  Panel ownership is static, so while a CTA waits at a grid
// (and cross-rank) barrier it already knows which rows it will stream next; with several ranks the
// barrier lasts about as long as streaming a rank's share of a layer, so the HBM traffic of the
// next phase hides behind the wait.
__device__ __forceinline__ void prefetch_l2(const void *p, unsigned bytes) {
  asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(p), "r"(bytes) : "memory");
}

// Weight accesses of this kernel: a row is only ever touched by the CTA that owns its panel, so
// no other SM's L1 can hold a stale copy and any cache operator is safe.  Measured on the wide
// shape: the evict-first (streaming) operators gain ~6% in the top-down passes and lose ~2% in the
// forward passes against the L1-bypassing ones, so each pass gets the one that suits it.
template <bool STREAM>
__device__ __forceinline__ double2 ldw2(const double2 *p) {
  return STREAM ? __ldcs(p) : __ldcg(p);
}
template <bool STREAM>
__device__ __forceinline__ void stw2(double2 *p, double2 v) {
  if (STREAM) __stcs(p, v);
  else __stcg(p, v);
}

constexpr int kPanelCols = 2 * kThreads;  // columns covered by one pass of the CTA (one double2 per thread)

// Read-only top-down pass over the own panel: four column chunks per thread, rows outermost, two
// rows in flight.  Pout[c] = sum_i W[r0+i, c] * dsl[i].
// Not inlined on purpose: inside the kernel body these loops share the 128 registers of a 512-thread CTA with
// everything that is live around them; on their own they keep DEPTH items of G double2 in flight without spilling
// (measured, profiles/r2_noinline_*.jsonl: 7-row panels 173.6 -> 166.9 us per iteration, 162.7 with DEPTH = 3).
template <bool STREAM, int DEPTH>
__device__ __noinline__ void panel_cols_read(const double *W, int ld, Panel pn, const double *dsl, double *Pout) {
  // The (column group, row) pairs form ONE software-pipelined sequence, two items in flight: with a
  // 7-row panel (8 ranks) a pipeline that restarted for every column group spent more time filling
  // and draining than streaming.
  constexpr int G = 4;  // chunks per group
  const int ng = (ld + G * kPanelCols - 1) / (G * kPanelCols);
  const double2 zero2 = make_double2(0.0, 0.0);
  double2 acc[G];
#pragma unroll
  for (int k = 0; k < G; ++k) acc[k] = zero2;
  int lg = 0, li = 0;  // next item to load
  int cg = 0, ci = 0;  // next item to consume
  auto load_next = [&](double2 (&w)[G]) {
    const bool on = lg < ng && pn.cnt > 0;
    const int cbase = lg * G * kPanelCols + 2 * threadIdx.x;
    const double *row = W + (size_t)(pn.r0 + li) * ld;
#pragma unroll
    for (int k = 0; k < G; ++k) {
      const int c = cbase + k * kPanelCols;
      w[k] = (on && c < ld) ? ldw2<STREAM>(reinterpret_cast<const double2 *>(row + c)) : zero2;
    }
    if (++li >= pn.cnt) {
      li = 0;
      ++lg;
    }
  };
  auto consume = [&](double2 (&w)[G]) {
    if (cg >= ng) return;
    const double d = dsl[ci];
#pragma unroll
    for (int k = 0; k < G; ++k) {
      acc[k].x = fma(w[k].x, d, acc[k].x);
      acc[k].y = fma(w[k].y, d, acc[k].y);
    }
    if (++ci >= pn.cnt) {  // group complete
      const int cbase = cg * G * kPanelCols + 2 * threadIdx.x;
#pragma unroll
      for (int k = 0; k < G; ++k) {
        const int c = cbase + k * kPanelCols;
        if (c < ld) stg_cg2(reinterpret_cast<double2 *>(Pout + c), acc[k]);
        acc[k] = zero2;
      }
      ci = 0;
      ++cg;
    }
  };
  double2 w[DEPTH][G];
  const int total = ng * pn.cnt;
#pragma unroll
  for (int d = 0; d + 1 < DEPTH; ++d) load_next(w[d]);
  for (int t = 0; t < total; t += DEPTH) {
#pragma unroll
    for (int d = 0; d < DEPTH; ++d) {
      load_next(w[(d + DEPTH - 1) % DEPTH]);
      consume(w[d]);
    }
  }
}

// Top-down pass that also brings S up to date (column chunk outermost: the 2 x 8 v factors of
// the thread's column pair stay in registers for the whole panel):
//   P[c] from the OLD stored weights, then w += sum_{j<m} us[i][j] v_j[c] + (lr*d_i) h[c].
// Pout == nullptr: final write-back, no sums wanted.
template <bool HS>  // HS: h lives in shared memory
__device__ __noinline__ void panel_cols_flush(double *W, int ld, Panel pn, const double *dsl, const double *us,
                                                 const double *HV, const double *h, double lr, int m, double *Pout) {
  const double2 zero2 = make_double2(0.0, 0.0);
  const int mc = (m + 1) >> 1;
  for (int c = 2 * threadIdx.x; c < ld; c += kPanelCols) {
    // (lr == 0: final write-back, h is not wanted and may be in another format)
    const double h0 = lr == 0.0 ? 0.0 : (HS ? h[c] : ldg_cg(h + c));
    const double h1 = lr == 0.0 ? 0.0 : (HS ? h[c + 1] : ldg_cg(h + c + 1));
    double v0[kLazyChunk], v1[kLazyChunk];
    {
      const double2 *a = reinterpret_cast<const double2 *>(HV + (size_t)c * kLazyChunk);
      const double2 *b = a + kLazyChunk / 2;
#pragma unroll
      for (int k = 0; k < kLazyChunk / 2; ++k) {
        const bool on = 2 * k < m;
        const double2 qa = on ? ldg_cg2(a + k) : zero2;
        const double2 qb = on ? ldg_cg2(b + k) : zero2;
        v0[2 * k] = qa.x;
        v0[2 * k + 1] = (2 * k + 1 < m) ? qa.y : 0.0;
        v1[2 * k] = qb.x;
        v1[2 * k + 1] = (2 * k + 1 < m) ? qb.y : 0.0;
      }
    }
    double a0 = 0.0, a1 = 0.0;
    double2 *Wp = reinterpret_cast<double2 *>(W + (size_t)pn.r0 * ld + c);
    const size_t stride2 = (size_t)(ld >> 1);
    double2 wa[4], wb[4];
    auto load = [&](double2 (&w)[4], int i0) {
#pragma unroll
      for (int u = 0; u < 4; ++u) w[u] = (i0 + u < pn.cnt) ? ldw2<true>(Wp + (size_t)(i0 + u) * stride2) : zero2;
    };
    auto consume = [&](double2 (&w)[4], int i0) {
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int i = i0 + u;
        if (i < pn.cnt) {
          const double d = dsl[i];
          a0 = fma(w[u].x, d, a0);
          a1 = fma(w[u].y, d, a1);
          const double2 *uc = reinterpret_cast<const double2 *>(us + (size_t)i * kLazyChunk);
#pragma unroll
          for (int k = 0; k < kLazyChunk / 2; ++k) {
            if (k < mc) {
              const double2 f = uc[k];  // same address in every thread: a shared-memory broadcast
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
          stw2<true>(Wp + (size_t)i * stride2, w[u]);
        }
      }
    };
    load(wa, 0);
    for (int i0 = 0; i0 < pn.cnt; i0 += 8) {
      load(wb, i0 + 4);
      consume(wa, i0);
      load(wa, i0 + 8);
      consume(wb, i0 + 4);
    }
    if (Pout) stg_cg2(reinterpret_cast<double2 *>(Pout + c), make_double2(a0, a1));
  }
}

// Momentum (BPM) top-down pass over the own panel, immediate update (src/ann.c:1996-1999,
// cublasDger + Daxpy + Dscal in src/cuda_ann.cu:3147-3166):
//   P[c] from the OLD weights, then  mom += (lr*d_i) h[c];  w += mom;  mom *= alpha.
// Same organisation as the flush pass above: every thread owns column pairs, the panel's rows are
// streamed once with two batches of four rows in flight, no cross-thread reduction.
template <bool HS>
__device__ __forceinline__ void panel_cols_bpm(double *W, double *M, int ld, Panel pn, const double *dsl,
                                               const double *h, double lr, double alpha, double *Pout) {
  const double2 zero2 = make_double2(0.0, 0.0);
  for (int c = 2 * threadIdx.x; c < ld; c += kPanelCols) {
    const double h0 = HS ? h[c] : ldg_cg(h + c);
    const double h1 = HS ? h[c + 1] : ldg_cg(h + c + 1);
    double a0 = 0.0, a1 = 0.0;
    double2 *Wp = reinterpret_cast<double2 *>(W + (size_t)pn.r0 * ld + c);
    double2 *Mp = reinterpret_cast<double2 *>(M + (size_t)pn.r0 * ld + c);
    const size_t stride2 = (size_t)(ld >> 1);
    double2 wa[4], ma[4], wb[4], mb[4];
    auto load = [&](double2 (&w)[4], double2 (&m)[4], int i0) {
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const bool on = i0 + u < pn.cnt;
        w[u] = on ? ldw2<true>(Wp + (size_t)(i0 + u) * stride2) : zero2;
        m[u] = on ? ldw2<true>(Mp + (size_t)(i0 + u) * stride2) : zero2;
      }
    };
    auto consume = [&](double2 (&w)[4], double2 (&m)[4], int i0) {
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int i = i0 + u;
        if (i < pn.cnt) {
          const double d = dsl[i];
          a0 = fma(w[u].x, d, a0);
          a1 = fma(w[u].y, d, a1);
          const double s = lr * d;
          m[u].x = fma(s, h0, m[u].x);
          m[u].y = fma(s, h1, m[u].y);
          w[u].x += m[u].x;
          w[u].y += m[u].y;
          stw2<true>(Wp + (size_t)i * stride2, w[u]);
          m[u].x *= alpha;
          m[u].y *= alpha;
          stw2<true>(Mp + (size_t)i * stride2, m[u]);
        }
      }
    };
    load(wa, ma, 0);
    for (int i0 = 0; i0 < pn.cnt; i0 += 8) {
      load(wb, mb, i0 + 4);
      consume(wa, ma, i0);
      load(wa, ma, i0 + 8);
      consume(wb, mb, i0 + 4);
    }
    stg_cg2(reinterpret_cast<double2 *>(Pout + c), make_double2(a0, a1));
  }
}

// sums of 8 values over the 32 lanes with 9 exchanges instead of 40: every step halves the number
// of values a lane is responsible for.  Returns the complete sum of value ((lane >> 2) & 7).
__device__ __forceinline__ double reduce8(const double (&pr)[8], int lane) {
  const bool b4 = lane & 16, b3 = lane & 8, b2 = lane & 4;
  double s4[4], s2[2];
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    const double keep = b4 ? pr[4 + q] : pr[q], give = b4 ? pr[q] : pr[4 + q];
    s4[q] = keep + __shfl_xor_sync(0xffffffffu, give, 16);
  }
#pragma unroll
  for (int q = 0; q < 2; ++q) {
    const double keep = b3 ? s4[2 + q] : s4[q], give = b3 ? s4[q] : s4[2 + q];
    s2[q] = keep + __shfl_xor_sync(0xffffffffu, give, 8);
  }
  const double keep = b2 ? s2[1] : s2[0], give = b2 ? s2[0] : s2[1];
  double s1 = keep + __shfl_xor_sync(0xffffffffu, give, 4);
  s1 += __shfl_xor_sync(0xffffffffu, s1, 2);
  s1 += __shfl_xor_sync(0xffffffffu, s1, 1);
  return s1;
}

constexpr int kZStride = kWarps;  // partial dot products per row kept in shared memory (>= 8 segments)

// z partials of the own panel, streaming variant for wide layers: like the top-down pass the CTA
// walks its rows one after the other, every thread holding the x values of four column pairs per
// group; the per-row sums are completed inside each warp (reduce8) and parked per warp in
// zpart[i*kZStride + warp] -- no CTA-wide barrier until the whole panel is done.
template <int UPD>
__device__ __noinline__ void panel_z_stream(double *W, int ld, Panel pn, const double *xs, const double *dsl,
                                               double lr, double *zpart) {
  constexpr int G = 4;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const double2 zero2 = make_double2(0.0, 0.0);
  for (int i = lane; i < pn.cnt; i += 32) zpart[i * kZStride + warp] = 0.0;  // own slots only
  __syncwarp();
  // The loads of all (column group, row) pairs form one pipelined sequence (two rows in flight, running
  // on into the next group); every group is padded to an even number of rows so that a row always
  // sits in the same one of the two register buffers.
  const int ng = (ld + G * kPanelCols - 1) / (G * kPanelCols);
  const int cp = pn.cnt + (pn.cnt & 1);
  int lg = 0, li = 0;  // next row to load
  double2 wa[G], wb[G];
  auto load_next = [&](double2 (&w)[G]) {
    const bool on = lg < ng && li < pn.cnt;
    const int cb = lg * G * kPanelCols + 2 * threadIdx.x;
    const double *row = W + (size_t)(pn.r0 + li) * ld;
#pragma unroll
    for (int k = 0; k < G; ++k) {
      const int c = cb + k * kPanelCols;
      w[k] = (on && c < ld) ? ldw2<false>(reinterpret_cast<const double2 *>(row + c)) : zero2;
    }
    if (++li >= cp) {
      li = 0;
      ++lg;
    }
  };
  load_next(wa);
  load_next(wb);
  for (int g = 0; g < ng; ++g) {
    const int cbase = g * G * kPanelCols + 2 * threadIdx.x;
    double2 xr[G];
#pragma unroll
    for (int k = 0; k < G; ++k) {
      const int c = cbase + k * kPanelCols;
      xr[k] = (c < ld) ? *reinterpret_cast<const double2 *>(xs + c) : zero2;
    }
    auto consume = [&](double2 (&w)[G], int i) -> double {
      double a = 0.0, b = 0.0;
      if (UPD && i < pn.cnt) {
        const double s = lr * dsl[i];
        double *row = W + (size_t)(pn.r0 + i) * ld;
#pragma unroll
        for (int k = 0; k < G; ++k) {
          const int c = cbase + k * kPanelCols;
          if (c < ld) {
            w[k].x = fma(s, xr[k].x, w[k].x);
            w[k].y = fma(s, xr[k].y, w[k].y);
            stw2<false>(reinterpret_cast<double2 *>(row + c), w[k]);
          }
        }
      }
#pragma unroll
      for (int k = 0; k < G; ++k) {
        a = fma(w[k].x, xr[k].x, a);
        b = fma(w[k].y, xr[k].y, b);
      }
      return a + b;
    };
    for (int i0 = 0; i0 < cp; i0 += 8) {
      double pr[8];
#pragma unroll
      for (int u = 0; u < 8; u += 2) {
        if (i0 + u < cp) {
          pr[u] = consume(wa, i0 + u);
          load_next(wa);
          pr[u + 1] = consume(wb, i0 + u + 1);
          load_next(wb);
        } else {
          pr[u] = 0.0;
          pr[u + 1] = 0.0;
        }
      }
      const double s = reduce8(pr, lane);
      const int i = i0 + ((lane >> 2) & 7);
      if ((lane & 3) == 0 && i < pn.cnt) zpart[i * kZStride + warp] += s;
    }
  }
}

// z partials of a SHORT panel (<= 8 rows: 8 ranks on the wide shape) when the input vector is still arriving
// from the other ranks: the row sums stay in registers across the whole row, the columns are walked in
// chunks of 1024 starting with the chunk this rank produced itself, the weights of the next chunk are
// already in flight while a warp waits (Barrier::wait_rank) for the ranks that own the columns of the
// current one.  The weight stream never waits for NVLink; only the FMAs do.  x is read from the arena
// (L2), two values per thread and chunk: no staging pass, no CTA-wide barrier until the panel is done.
__device__ __forceinline__ void panel_z_chunks(const double *W, int ld, Panel pn, const double *x, const hpnn_dev_net &n,
                                               int l, const Barrier &bar, double *zpart) {
  constexpr int R = 8;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const double2 zero2 = make_double2(0.0, 0.0);
  const int nch = (ld + kPanelCols - 1) / kPanelCols;
  const int ncols = n.dims[l];  // real columns = rows of layer l-1, split over the ranks
  int own0 = 0;
  for (int q = 0; q < n.rank; ++q) own0 += s_rows_of[l - 1][q];
  const int c_first = min(own0 / kPanelCols, nch - 1);
  double acc[R];
#pragma unroll
  for (int i = 0; i < R; ++i) acc[i] = 0.0;
  double2 wa[R], wb[R];
  auto loadw = [&](double2 (&w)[R], int cc) {
    const int c = (c_first + cc) % nch;
    const int col = c * kPanelCols + 2 * threadIdx.x;
    const bool on = cc < nch && col < ld;
#pragma unroll
    for (int i = 0; i < R; ++i)
      w[i] = (on && i < pn.cnt) ? ldw2<false>(reinterpret_cast<const double2 *>(W + (size_t)(pn.r0 + i) * ld + col)) : zero2;
  };
  auto consume = [&](double2 (&w)[R], int cc) {
    if (cc >= nch) return;
    const int c = (c_first + cc) % nch;
    const int lo = c * kPanelCols, hi = min(ncols, lo + kPanelCols);
    if (lane == 0) {  // the ranks that produced columns [lo, hi)
      int r0 = 0;
      for (int q = 0; q < n.world; ++q) {
        const int r1 = r0 + s_rows_of[l - 1][q];
        if (r0 < hi && r1 > lo) bar.wait_rank(n, q);
        r0 = r1;
      }
    }
    __syncwarp();
    const int col = lo + 2 * threadIdx.x;
    const double2 x2 = col < ld ? ldg_cg2(reinterpret_cast<const double2 *>(x + col)) : zero2;
#pragma unroll
    for (int i = 0; i < R; ++i) {
      acc[i] = fma(w[i].x, x2.x, acc[i]);
      acc[i] = fma(w[i].y, x2.y, acc[i]);
    }
  };
  loadw(wa, 0);
  for (int cc = 0; cc < nch; cc += 2) {
    loadw(wb, cc + 1);
    consume(wa, cc);
    loadw(wa, cc + 2);
    consume(wb, cc + 1);
  }
  const double sres = reduce8(acc, lane);
  const int i = (lane >> 2) & 7;
  if ((lane & 3) == 0 && i < pn.cnt) zpart[i * kZStride + warp] = sres;
}

// Row tasks over the own panel.
//   UPD 0: z_i = S[r,:].x + sum_{j<mcorr} us[i][j] coef[j]   (read-only; us == nullptr: plain)
//   UPD 1: w += lr*dsl[i] * x ; z_i = w.x                     (immediate update; kept for layers
//          that cannot be deferred, not instantiated at present)
// y goes to every rank's activation buffer (dst); while it is at hand the CTA's share of the next layer's v_j . y (j < m) is formed.
template <int UPD>
__device__ __noinline__ void panel_rows(double *W, int ld, Panel pn, const double *xs, const double *dsl, double lr,
                                           int act_kind, const Dst &dst, int segs, double *zpart, const double *us,
                                           const double *coef, int mcorr, int m, const DotOut &dots, double *red,
                                           double *ykeep = nullptr, bool z_ready = false) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int nz;  // partials per row
  if (z_ready) {  // panel_z_chunks has filled zpart[i][0..kWarps)
    nz = kWarps;
  } else if (ld >= 4 * kPanelCols) {
    panel_z_stream<UPD>(W, ld, pn, xs, dsl, lr, zpart);
    nz = kWarps;
  } else {
    // narrow layers: one warp per (row, segment), see phase_rows
    const int n2 = ld >> 1;
    const int seg2 = (n2 + segs - 1) / segs;
    for (int u = warp; u < pn.cnt * segs; u += kWarps) {
      const int i = u / segs, sg = u % segs;
      const int c0 = sg * seg2, c1 = min(n2, c0 + seg2);
      double *wrow = W + (size_t)(pn.r0 + i) * ld + 2 * c0;
      const double s = UPD ? lr * dsl[i] : 0.0;
      const double z = c1 > c0 ? warp_row<UPD>(wrow, nullptr, xs + 2 * c0, 2 * (c1 - c0), s, 0.0, lane) : 0.0;
      if (lane == 0) zpart[i * kZStride + sg] = z;
    }
    nz = segs;
  }
  __syncthreads();
  const bool want_dots = dots.p[0] != nullptr && m > 0 && pn.cnt > 0;
  double acc[kLazyChunk];
#pragma unroll
  for (int j = 0; j < kLazyChunk; ++j) acc[j] = 0.0;
  for (int i = threadIdx.x; i < pn.cnt; i += blockDim.x) {
    double z = 0.0;
    for (int q = 0; q < nz; ++q) z += zpart[i * kZStride + q];
    if (!UPD && us && mcorr > 0) {
      double corr = 0.0;
      for (int j = 0; j < mcorr; ++j) corr = fma(us[(size_t)i * kLazyChunk + j], coef[j], corr);
      z += corr;
    }
    const double y = act_kind == 0 ? act_f(z) : z;
    dst_store(dst, pn.r0 + i, y);
    if (ykeep) ykeep[i] = y;  // own outputs stay in shared memory (folded output layer)
    if (want_dots) chunk_fma(dots.HV + (size_t)(dots.row0 + pn.r0 + i) * kLazyChunk, y, m, acc);
  }
  if (want_dots) {  // CTA-uniform
    __shared__ double mine_dots[kLazyChunk];
    chunk_block_sum(acc, m, mine_dots, red);
    if (threadIdx.x < kLazyChunk) {
      if (!dots.tag && dots.mc) {
        multimem_st_f64(dots.mc + threadIdx.x, mine_dots[threadIdx.x]);
      } else {
        for (int q = 0; q < dots.n; ++q) {
          if (dots.tag) st_pair<true>(dots.p[q], threadIdx.x, mine_dots[threadIdx.x], dots.tag);
          else stg_cg(dots.p[q] + threadIdx.x, mine_dots[threadIdx.x]);
        }
      }
    }
  }
}

// tagged variant of gather_dots: every producer chunk is 8 {value, tag} pairs in this rank's arena,
// polled until the tag of the running iteration shows up (no barrier, no fence)
__device__ __forceinline__ void gather_dots_tagged(const hpnn_dev_net &n, int l, int m, unsigned long long tag,
                                                   double *coef, double *red, int grid) {
  const double *slots = n.arena[n.rank] + n.off_hist_dot[l];
  double acc[kLazyChunk];
#pragma unroll
  for (int j = 0; j < kLazyChunk; ++j) acc[j] = 0.0;
  for (int idx = threadIdx.x; idx < n.world * kLazyDotSlots; idx += kThreads) {
    const int q = idx / kLazyDotSlots, b = idx % kLazyDotSlots;
    const int rows_q = s_rows_of[l - 1][q];
    if (b < grid && b < rows_q) {
      const double *addr[kLazyChunk];
      bool want[kLazyChunk];
      double v[kLazyChunk];
#pragma unroll
      for (int j = 0; j < kLazyChunk; ++j) {
        want[j] = j < m;
        addr[j] = slots + 2 * ((size_t)idx * kLazyChunk + j);
      }
      ld_pairs_wait<kLazyChunk, false>(addr, want, tag, v);
#pragma unroll
      for (int j = 0; j < kLazyChunk; ++j) acc[j] += v[j];
    }
  }
  chunk_block_sum(acc, m, coef, red);
}

// TG: several ranks with the tagged-pair hand-over of K-P (kp_common.cuh): activations, slice sums
// and dot-product chunks cross NVLink as {value, tag} pairs and are polled by their consumers, the
// grid barriers stay local to the GPU and there is no system-scope fence.  TG == false: plain
// buffers + fence / flag barriers (also the single-rank case).
template <bool TG>
__device__ __forceinline__ void train_sample_lazy1_body(const KP &p, const int grid, const int b) {
  extern __shared__ double smem[];
  watch_init(p);
  if (p.fault == 1) return;  // injected fault: this rank never shows up
  rows_table_init(p.net);
  const hpnn_dev_net &n = p.net;
  const int NL = n.n_layers;
  const int pmax = p.max_rows;  // here: widest panel, padded to 8 (see the launcher)
  double *xs = smem;                 // input vector of the running forward phase
  double *dsl = xs + p.max_ld;       // deltas of the own panel of the layer being processed
  double *us = dsl + pmax;           // u factors of the own panels: [layer][row][8]
  double *red = us + (size_t)NL * pmax * kLazyChunk;
  double *rowpart = red + kRedDoubles;
  double *ylast = rowpart + (size_t)pmax * kZStride;  // own outputs of the last hidden layer (folded output layer)
  double *zs = ylast + pmax;                          // output vector / output deltas, kFoldMaxOut each
  double *dlast = zs + kFoldMaxOut;
  __shared__ double coef[kLazyChunk];

  const int n_out = n.dims[NL];
  double *A = n.arena[n.rank];
  const bool snn = p.type == HPNN_TYPE_SNN;
  const int depth = p.lazy_depth;
  const bool multi = n.world > 1;
  // momentum: immediate updates (depth 1, no pending factors), the momentum rows of a panel are
  // only ever touched by the CTA that owns the panel
  const bool bpm = p.train == HPNN_TRAIN_BPM;
  constexpr bool tg = TG;
  // Output layer folded into the layer below it (narrow output layers, plain-buffer protocol).  A
  // 10 x 8192 output layer is 0.04% of the weights but used to cost four phases and four barriers
  // per iteration.  Instead every rank keeps a working copy WL of the whole layer in its arena and
  // uses only the COLUMNS that match its own rows of the last hidden layer:
  //   forward : the CTA that has just produced outputs y of its panel adds WL[:, panel] . y and
  //             publishes <= 32 partial sums per CTA; after the phase barrier of the hidden layer
  //             every CTA adds the partial sums of all CTAs and ranks in a fixed order;
  //   backward: delta_out is known to every CTA (redundant evaluation), so the CTA forms
  //             WL[:, panel]^T delta_out, the hidden deltas of its panel and the BP / BPM update of
  //             those columns locally: no phase, no barrier;
  //   sample end: the column slices are gathered into every rank's copy and the owners of the rows
  //             copy them back to the weight storage.
  const bool fold = !TG && p.fold_out != 0;
  const int NLw = fold ? NL - 1 : NL;  // layers that have phases of their own
  double *WL = A + n.off_wlast, *ML = A + n.off_mlast;
  const int ldl = n.ld[NL - 1];
  const unsigned long long launch = (unsigned long long)ld_relaxed_gpu(n.local_sync + 160) + 1ull;
  auto tag_of = [&](int iteration) { return (launch << 32) | (unsigned long long)(unsigned)iteration; };

  Barrier bar;
  bar.init(n, grid, b);
  int trace_n = 0;
  auto trace = [&](int tag) {
    if (p.trace && b == 0 && threadIdx.x == 0 && trace_n < p.trace_cap) {
      p.trace[2 * trace_n] = tag;
      long long now;
      if (p.trace_ns) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
      else now = clock64();
      p.trace[2 * trace_n + 1] = now;
      ++trace_n;
    }
  };
  int cur = 0;
  auto act_peer = [&](int q, int i, int buf) -> double * {
    return n.arena[q] + n.off_act[i] + ((i == 0) ? 0 : (size_t)buf * pad8(n.dims[i]) * (tg ? 2 : 1));
  };
  auto act_of = [&](int i, int buf) -> double * { return act_peer(n.rank, i, buf); };
  // act[i] (buffer buf) into shared memory; tagged mode: wait until every value carries `tag`
  auto stage_act = [&](double *dst_smem, int i, int buf, unsigned long long tag) {
    if (tg && i > 0) stage_pairs<false>(dst_smem, act_of(i, buf), n.dims[i], pad8(n.dims[i]), tag);
    else stage_vec(dst_smem, act_of(i, buf), pad8(n.dims[i]));
  };
  auto phase_sync = [&](bool remote) {
    if (tg) bar.sync_local();
    else bar.sync_all(n, remote);
  };
  auto panel = [&](int l) { return panel_of(n.rows[l], grid, b); };  // local row indices of this rank's block
  auto us_of = [&](int l) { return us + (size_t)l * pmax * kLazyChunk; };
  // first rows of the own panel of layer l into L2, at most p.prefetch_bytes per rank and call
  auto prefetch_panel = [&](int l) {
    if (threadIdx.x != 0 || l < 0 || l >= NL || p.prefetch_bytes <= 0) return;
    const Panel pn = panel(l);
    const long long row_bytes = 8LL * n.ld[l];
    long long rows = p.prefetch_bytes / grid / row_bytes;
    rows = rows < pn.cnt ? rows : pn.cnt;
    for (int i = 0; i < (int)rows; ++i) prefetch_l2(n.W[l] + (size_t)(pn.r0 + i) * n.ld[l], (unsigned)row_bytes);
  };
  // per-CTA slots: partial column sums (+ 8 partial u_j . delta behind them), two buffers
  auto pslot = [&](int buf, int cta) { return A + n.off_panel + ((size_t)buf * kLazyDotSlots + cta) * n.panel_stride; };

  // Layer 0 sees the same input x in every iteration, so its pending updates collapse into ONE
  // factor pair: W0_true = S0 + ubar (x) x with ubar = sum of lr*delta_0 so far (slot 0 of the
  // layer's u chunks), z = S0.x + ubar*(x.x).  W0 is read-only until the end of the sample.
  __shared__ double coef0[kLazyChunk];
  {
    const double *x0 = act_of(0, 0);
    double s = 0.0;
    for (int c = threadIdx.x; c < n.dims[0]; c += kThreads) {
      const double v = ldg_cg(x0 + c);
      s = fma(v, v, s);
    }
    s = block_sum(s, red);
    if (threadIdx.x < kLazyChunk) coef0[threadIdx.x] = threadIdx.x == 0 ? s : 0.0;
    for (int i = threadIdx.x; i < pmax; i += kThreads) {
      us[(size_t)i * kLazyChunk] = 0.0;      // ubar
      us[(size_t)i * kLazyChunk + 1] = 0.0;  // BPM: mu, the momentum of layer 0 in factored form (M0 = mu (x) x)
    }
    __syncthreads();
  }
  if (bpm) {
    // the momentum is zeroed per sample (src/ann.c:2386); own panel rows only, so no barrier is needed
    // before the owning CTA reads them in its first top-down pass
    for (int l = 1; l < NLw; ++l) {
      const Panel pn = panel_of(n.rows[l], grid, b);
      double2 *m2 = reinterpret_cast<double2 *>(n.dW[l] + (size_t)pn.r0 * n.ld[l]);
      const size_t tot = (size_t)pn.cnt * n.ld[l] / 2;
      for (size_t i = threadIdx.x; i < tot; i += kThreads) stw2<true>(m2 + i, make_double2(0.0, 0.0));
    }
  }
  if (fold) {
    // working copy of the output layer: every rank contributes its rows to every rank's copy (first
    // used after the barrier of the first forward phase); momentum copy starts from zero
    const int lr_ = NL - 1;
    const size_t tot = (size_t)n.rows[lr_] * ldl;
    for (size_t i = (size_t)b * kThreads + threadIdx.x; i < tot; i += (size_t)grid * kThreads) {
      const double v = ldg_cg(n.W[lr_] + i);
      const size_t off = (size_t)n.off_wlast + (size_t)n.row0[lr_] * ldl + i;
      if (n.arena_mc) multimem_st_f64(n.arena_mc + off, v);
      else
        for (int q = 0; q < n.world; ++q) stg_cg(n.arena[q] + off, v);
    }
    if (bpm) {
      const size_t all = (size_t)n_out * ldl;
      for (size_t i = (size_t)b * kThreads + threadIdx.x; i < all; i += (size_t)grid * kThreads) stg_cg(ML + i, 0.0);
    }
  }
  // folded output layer, forward part: this CTA's share of WL . y, one slot per (rank, CTA) on every rank
  auto fold_publish = [&]() {
    const Panel pn = panel_of(n.rows[NL - 2], grid, b);
    const int col0 = n.row0[NL - 2] + pn.r0;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    __syncthreads();  // ylast complete
    for (int r = warp; r < n_out; r += kWarps) {
      double s = 0.0;
      for (int i = lane; i < pn.cnt; i += 32) s = fma(ldg_cg(WL + (size_t)r * ldl + col0 + i), ylast[i], s);
#pragma unroll
      for (int o = 16; o; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
      if (lane == 0) zs[r] = s;
    }
    __syncthreads();
    if ((int)threadIdx.x < n_out && pn.cnt > 0) {
      const size_t off = (size_t)n.off_zpart + ((size_t)(n.rank * kLazyDotSlots + b) * kFoldMaxOut + threadIdx.x);
      if (n.arena_mc) multimem_st_f64(n.arena_mc + off, zs[threadIdx.x]);
      else
        for (int q = 0; q < n.world; ++q) stg_cg(n.arena[q] + off, zs[threadIdx.x]);
    }
  };
  // layer l >= 1 consumes its input slice by slice (panel_z_chunks) instead of behind a full cross-rank barrier
  auto slice_ok = [&](int l) -> bool {
    return !TG && p.slicewise && multi && l >= 1 && l < NLw &&
           panel_max((n.dims[l + 1] + n.world - 1) / n.world, grid) <= 8;
  };
  auto forward = [&](int l, int src_buf, int dst_buf, int m, unsigned long long tag) {
    const bool sw = slice_ok(l);
    if (sw) {
      trace(600 + l);
      panel_z_chunks(n.W[l], n.ld[l], panel(l), act_of(l, src_buf), n, l, bar, rowpart);
      __syncthreads();  // every warp has seen every rank's flag: the whole input (and the dot-product chunks) is there
      if (m > 0) gather_dots(n, l, m, coef, red, grid);
    } else {
      stage_act(xs, l, src_buf, tag);  // produced earlier in the same sweep
      __syncthreads();
      trace(600 + l);  // input vector staged (tagged mode: every producer's value has arrived)
      if (l >= 1 && m > 0) {  // coef[j] = v_j . x
        if (tg) gather_dots_tagged(n, l, m, tag, coef, red, grid);
        else gather_dots(n, l, m, coef, red, grid);
      }
    }
    trace(700 + l);  // partial dot products gathered
    Dst dst;  // all-gather by peer stores
    dst.n = n.world;
    dst.tag = tg ? tag : 0ull;
    dst.mc = (!tg && n.arena_mc) ? n.arena_mc + (act_of(l + 1, dst_buf) - A) + n.row0[l] : nullptr;
    for (int q = 0; q < n.world; ++q) dst.p[q] = act_peer(q, l + 1, dst_buf) + (size_t)n.row0[l] * (tg ? 2 : 1);
    DotOut dots;
    dots.n = n.world;
    dots.tag = tg ? tag : 0ull;
    dots.mc = nullptr;
    dots.p[0] = nullptr;
    if (l + 1 < NLw && m > 0) {
      for (int q = 0; q < n.world; ++q)
        dots.p[q] = n.arena[q] + n.off_hist_dot[l + 1] +
                    (size_t)(n.rank * kLazyDotSlots + b) * kLazyChunk * (tg ? 2 : 1);
      dots.HV = A + n.off_hist_v[l + 1];
      dots.row0 = n.row0[l];
      if (!tg && n.arena_mc) dots.mc = n.arena_mc + (dots.p[n.rank] - A);
    }
    const int kind = (snn && l == NL - 1) ? 1 : 0;
    const int segs = pick_row_segments(n.rows[l], n.ld[l], grid);
    const bool feeds_out = fold && l == NL - 2;
    panel_rows<0>(n.W[l], n.ld[l], panel(l), xs, dsl, p.lr, kind, dst, segs, rowpart, us_of(l),
                  l == 0 ? coef0 : coef, l == 0 ? 1 : m, m, dots, red, feeds_out ? ylast : nullptr, sw);
    if (feeds_out) fold_publish();
  };
  // barrier after the forward pass of layer l: only the arrive half when the next layer waits per producer rank
  auto forward_sync = [&](int l) {
    if (slice_ok(l + 1)) bar.arrive_all(n);
    else phase_sync(panel(l).cnt > 0);
  };

  for (int l = 0; l < NLw; ++l) {
    forward(l, cur, cur, 0, tag_of(0));
    forward_sync(l);
  }
  const double *target = A + n.off_target;
  const int last = NLw - 1;  // top-most layer with phases of its own
  const Panel plast = panel(NL - 1);
  const int out_row0 = n.row0[NL - 1] + plast.r0;
  // folded output layer: the output vector from the per-CTA partial sums of all ranks (rank-major, CTA
  // order: the same order on every CTA and every rank), then the usual statistics; dlast = all output deltas
  auto evaluate_fold = [&](double *o_out) -> OutStats {
    const double *zp = A + n.off_zpart;
    // thread = (output r, part): the valid slots (rank q, CTA bb < nb_q) are dealt to the parts
    // round-robin in rank-major order, eight loads in flight per thread
    const int lanes = n_out, nparts = kThreads / lanes;
    const int r = threadIdx.x % lanes, part = threadIdx.x / lanes;
    double s = 0.0;
    __syncthreads();
    if (part < nparts) {
      // all loads of all ranks' slots in flight at once (one L2 round trip for narrow output layers)
      double v[HPNN_MAX_WORLD * 4];
#pragma unroll
      for (int q = 0; q < HPNN_MAX_WORLD; ++q) {
        const int nbq = q < n.world ? min(grid, s_rows_of[NL - 2][q]) : 0;
        const double *base = zp + (size_t)q * kLazyDotSlots * kFoldMaxOut + r;
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const int bb = part + u * nparts;
          v[q * 4 + u] = bb < nbq ? ldg_cg(base + (size_t)bb * kFoldMaxOut) : 0.0;
        }
      }
#pragma unroll
      for (int j = 0; j < HPNN_MAX_WORLD * 4; ++j) s += v[j];
      for (int q = 0; q < n.world; ++q) {  // wide output layers (few parts): the slots beyond four per part
        const int nbq = min(grid, s_rows_of[NL - 2][q]);
        const double *base = zp + (size_t)q * kLazyDotSlots * kFoldMaxOut + r;
        for (int bb = part + 4 * nparts; bb < nbq; bb += nparts) s += ldg_cg(base + (size_t)bb * kFoldMaxOut);
      }
      red[part * lanes + r] = s;
    }
    __syncthreads();
    if ((int)threadIdx.x < n_out) {
      double z = 0.0;
      for (int q = 0; q < nparts; ++q) z += red[q * lanes + threadIdx.x];
      zs[threadIdx.x] = snn ? z : act_f(z);
    }
    __syncthreads();
    return output_stats<true>(p.type, p.stable, zs, target, n_out, 0, n_out, dlast, o_out, red);
  };
  // output statistics from the gathered output vector (tagged mode: gathered into xs first)
  auto evaluate = [&](unsigned long long tag, int row0, int rows, double *dsp, double *o_out) -> OutStats {
    if (tg) {
      __syncthreads();
      stage_act(xs, NL, cur, tag);
      __syncthreads();
      return output_stats<true>(p.type, p.stable, xs, target, n_out, row0, rows, dsp, o_out, red);
    }
    return output_stats(p.type, p.stable, act_of(NL, cur), target, n_out, row0, rows, dsp, o_out, red);
  };
  OutStats st = fold ? evaluate_fold(nullptr) : evaluate(tag_of(0), out_row0, plast.cnt, dsl, nullptr);
  double E = st.err;
  const double init_err = E;
  int iter = 0, ok = (st.argmax == st.target), first_ok = 0;
  double dE = 0.0;
  int m = 0;  // pending factors of every layer l >= 1 (identical on all CTAs)

  for (;;) {
    ++iter;
    trace(1);
    const unsigned long long abort_seen = watch_peek();
    const unsigned long long tag = tag_of(iter);
    const bool flush = (m + 1 >= depth);
    if (fold) {
      // output layer, backward part: hidden deltas of the own panel from the OLD columns, then the
      // BP / BPM update of exactly those columns (src/ann.c:1644, 1996-1999), all inside this CTA
      const Panel pn = panel(NL - 2);
      const int col0 = n.row0[NL - 2] + pn.r0;
      for (int i = threadIdx.x; i < pn.cnt; i += kThreads) {
        const double h = ylast[i];
        double P = 0.0;
        for (int r = 0; r < n_out; ++r) {
          double *wp = WL + (size_t)r * ldl + col0 + i;
          double w = ldg_cg(wp);
          const double d = dlast[r];
          P = fma(w, d, P);
          if (!bpm) {
            w = fma(p.lr * d, h, w);
          } else {
            double *mp = ML + (size_t)r * ldl + col0 + i;
            const double mm = fma(p.lr * d, h, ldg_cg(mp));
            w += mm;
            stg_cg(mp, mm * p.alpha);
          }
          stg_cg(wp, w);
        }
        dsl[i] = P * dact_f(h);
      }
      __syncthreads();
    }
    // ---- top-down
    for (int l = last; l >= 1; --l) {
      const Panel pn = panel(l);
      double *ul = us_of(l);
      double *HV = A + n.off_hist_v[l];
      const double *h = act_of(l, cur);  // input of layer l in the previous sweep
      if (tg) {  // present for a whole sweep already, but stored as pairs: unpack into shared memory
        stage_act(xs, l, cur, tag - 1);
        __syncthreads();
        h = xs;
      }
      double *Pout = pslot(l & 1, b);
      if (pn.cnt > 0 && threadIdx.x < 32) {
        // this panel's share of u_j . delta (factor j = lane & 7, rows dealt to the four lane groups),
        // parked behind the column sums.  Reads slots < m only: no conflict with the append below.
        const int j = threadIdx.x & 7;
        double s = 0.0;
        if (j < m)
          for (int i = threadIdx.x >> 3; i < pn.cnt; i += 4) s = fma(ul[(size_t)i * kLazyChunk + j], dsl[i], s);
        s += __shfl_xor_sync(0xffffffffu, s, 8);
        s += __shfl_xor_sync(0xffffffffu, s, 16);
        if (threadIdx.x < kLazyChunk) stg_cg(Pout + n.panel_stride - kLazyChunk + j, s);
      }
      if (!flush) {
        for (int i = threadIdx.x; i < pn.cnt; i += kThreads) ul[(size_t)i * kLazyChunk + m] = p.lr * dsl[i];
        const int gt = b * kThreads + threadIdx.x;
        for (int c = gt; c < n.ld[l]; c += grid * kThreads)
          stg_cg(HV + (size_t)c * kLazyChunk + m, tg ? h[c] : ldg_cg(h + c));
        if (pn.cnt > 0) {
          // a layer share that fits L2 is read again soon (top-down of layer 1 -> its forward pass, forward
          // of the top layer -> its top-down pass): keep it cached; otherwise stream with evict-first
          if (p.items_in_flight >= 3) {
            if (p.keep_l2) panel_cols_read<false, 3>(n.W[l], n.ld[l], pn, dsl, Pout);
            else panel_cols_read<true, 3>(n.W[l], n.ld[l], pn, dsl, Pout);
          } else {
            if (p.keep_l2) panel_cols_read<false, 2>(n.W[l], n.ld[l], pn, dsl, Pout);
            else panel_cols_read<true, 2>(n.W[l], n.ld[l], pn, dsl, Pout);
          }
        }
      } else if (pn.cnt > 0) {
        if (bpm) panel_cols_bpm<TG>(n.W[l], n.dW[l], n.ld[l], pn, dsl, h, p.lr, p.alpha, Pout);
        else panel_cols_flush<TG>(n.W[l], n.ld[l], pn, dsl, ul, HV, h, p.lr, m, Pout);
      }
      trace(100 + l);
      prefetch_panel(l - 1);  // streamed next: top-down of layer l-1, or the forward pass of layer 0
      bar.sync_local();
      trace(200 + l);
      // ---- reduce: this rank's partial sums, column slice by column slice.  One rank: the slice
      // of CTA b is its own panel of the layer below, so the deltas are finished here.  Several
      // ranks: the slice sums go to the rank owning that row of the layer below (reduce-scatter by
      // peer stores into the part slots of K-P), and after the cross-rank barrier the owning CTA
      // adds one value per rank.
      const int nbp = grid < n.rows[l] ? grid : n.rows[l];  // CTAs of this rank that own rows of layer l
      if (m > 0) {
        double acc[kLazyChunk];
#pragma unroll
        for (int j = 0; j < kLazyChunk; ++j) acc[j] = 0.0;
        for (int q = threadIdx.x; q < nbp; q += kThreads)
          chunk_fma(pslot(l & 1, q) + n.panel_stride - kLazyChunk, 1.0, m, acc);
        chunk_block_sum(acc, m, coef, red);  // coef[j] = u_j . delta_l over this rank's rows
      }
      const Panel slice = panel_of(n.dims[l], grid, b);  // columns of layer l == rows of layer l-1 (global)
      const double *hb = h;                   // output of layer l-1
      const int cols = n.dims[l];
      for (int ci0 = 0; ci0 < slice.cnt; ci0 += 64) {
        const int ci = ci0 + (threadIdx.x & 63), g = threadIdx.x >> 6;
        double s = 0.0;
        if (ci < slice.cnt) {
          // all (<= 20) partials of this thread in flight at once, added in CTA order
          constexpr int NQ = kLazyDotSlots / (kThreads / 64);
          double v[NQ];
#pragma unroll
          for (int k = 0; k < NQ; ++k) {
            const int q = g + k * (kThreads / 64);
            v[k] = (q < nbp) ? ldg_cg(pslot(l & 1, q) + slice.r0 + ci) : 0.0;
          }
#pragma unroll
          for (int k = 0; k < NQ; ++k) s += v[k];
        }
        __syncthreads();
        red[threadIdx.x] = s;
        __syncthreads();
        if (threadIdx.x < 64 && ci < slice.cnt) {
          double pc = 0.0;
#pragma unroll
          for (int gg = 0; gg < kThreads / 64; ++gg) pc += red[gg * 64 + threadIdx.x];
          const int c = slice.r0 + ci;
          if (m > 0) pc += chunk_apply(HV + (size_t)c * kLazyChunk, coef, m);
          if (!multi) {
            dsl[ci] = pc * dact_f(ldg_cg(hb + c));
          } else {
            const int own = owner_of_row(c, cols, n.world);
            const size_t slot = (size_t)(n.rank * kMaxRsplit) * n.part_stride[l - 1] + (c - row0_of(own, cols, n.world));
            if (tg) st_pair<true>(n.arena[own] + n.off_part[l - 1], (long long)slot, pc, tag);
            else stg_cg(n.arena[own] + n.off_part[l - 1] + slot, pc);
          }
        }
      }
      if (multi) {
        trace(800 + l);  // column-slice sums formed and sent to their owners
        if (!tg) bar.sync_all(n, slice.cnt > 0);
        trace(500 + l);
        const Panel below = panel(l - 1);
        const double *part = A + n.off_part[l - 1];
        for (int i = threadIdx.x; i < below.cnt; i += kThreads) {
          double s = 0.0;
          if (tg) {  // one {value, tag} pair per rank, polled until this iteration's value is there
            const double *addr[HPNN_MAX_WORLD];
            bool want[HPNN_MAX_WORLD];
            double v[HPNN_MAX_WORLD];
#pragma unroll
            for (int g = 0; g < HPNN_MAX_WORLD; ++g) {
              want[g] = g < n.world;
              addr[g] = part + 2 * ((size_t)((want[g] ? g : 0) * kMaxRsplit) * n.part_stride[l - 1] + below.r0 + i);
            }
            ld_pairs_wait<HPNN_MAX_WORLD, false>(addr, want, tag, v);
#pragma unroll
            for (int g = 0; g < HPNN_MAX_WORLD; ++g) s += v[g];
            dsl[i] = s * dact_f(hb[n.row0[l - 1] + below.r0 + i]);  // hb == xs here
          } else {
            for (int g = 0; g < n.world; ++g)
              s += ldg_cg(part + (size_t)(g * kMaxRsplit) * n.part_stride[l - 1] + below.r0 + i);
            dsl[i] = s * dact_f(ldg_cg(hb + n.row0[l - 1] + below.r0 + i));
          }
        }
      }
      __syncthreads();
    }
    m = flush ? 0 : m + 1;
    // ---- bottom-up: layer 0 update fused with its forward, then the read-only layers
    {
      const Panel p0 = panel(0);
      if (!bpm) {
        for (int i = threadIdx.x; i < p0.cnt; i += kThreads) us[(size_t)i * kLazyChunk] += p.lr * dsl[i];  // ubar
      } else {
        // dw += lr*delta (x) x; W += dw; dw *= alpha with dw = mu (x) x: the triple acts on the row scalars
        for (int i = threadIdx.x; i < p0.cnt; i += kThreads) {
          const double mu = us[(size_t)i * kLazyChunk + 1] + p.lr * dsl[i];
          us[(size_t)i * kLazyChunk] += mu;
          us[(size_t)i * kLazyChunk + 1] = mu * p.alpha;
        }
      }
    }
    forward(0, 0, cur ^ 1, m, tag);  // (stages x and synchronises before ubar is read)
    trace(300);
    prefetch_panel(1);
    forward_sync(0);
    trace(400);
    for (int l = 1; l < NLw; ++l) {
      forward(l, cur ^ 1, cur ^ 1, m, tag);
      trace(300 + l);
      prefetch_panel(l + 1 < NLw ? l + 1 : last - 1);  // after the last layer: the first wide top-down pass
      forward_sync(l);
      trace(400 + l);
    }
    cur ^= 1;
    watch_publish(abort_seen);
    st = fold ? evaluate_fold(nullptr) : evaluate(tag, out_row0, plast.cnt, dsl, nullptr);
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
  {  // layer 0: S0 += ubar (x) x, one warp per own row
    const Panel p0 = panel(0);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    stage_vec(xs, act_of(0, 0), pad8(n.dims[0]));
    __syncthreads();
    const double2 *x2 = reinterpret_cast<const double2 *>(xs);
    for (int i = warp; i < p0.cnt; i += kWarps) {
      const double ub = us[(size_t)i * kLazyChunk];
      double2 *w2 = reinterpret_cast<double2 *>(n.W[0] + (size_t)(p0.r0 + i) * n.ld[0]);
      for (int c = lane; c < (n.ld[0] >> 1); c += 32) {
        double2 w = ldg_cg2(w2 + c);
        w.x = fma(ub, x2[c].x, w.x);
        w.y = fma(ub, x2[c].y, w.y);
        stg_cg2(w2 + c, w);
      }
      if (bpm) {  // leave the momentum matrix of layer 0 as the immediate updates would have: mu (x) x
        const double mu = us[(size_t)i * kLazyChunk + 1];
        double2 *m2 = reinterpret_cast<double2 *>(n.dW[0] + (size_t)(p0.r0 + i) * n.ld[0]);
        for (int c = lane; c < (n.ld[0] >> 1); c += 32) stg_cg2(m2 + c, make_double2(mu * x2[c].x, mu * x2[c].y));
      }
    }
  }
  if (m > 0) {
    __syncthreads();
    for (int i = threadIdx.x; i < pmax; i += kThreads) dsl[i] = 0.0;  // lr == 0 and no sums are wanted
    __syncthreads();
    for (int l = 1; l < NLw; ++l) {
      const Panel pn = panel(l);
      if (pn.cnt > 0)
        panel_cols_flush<false>(n.W[l], n.ld[l], pn, dsl, us_of(l), A + n.off_hist_v[l], act_of(l, cur), 0.0, m, nullptr);
    }
  }
  if (fold) {
    if (b == 0) evaluate_fold(A + n.off_out);  // publish the final (normalised) output
    // the updated column slice of the output layer into every rank's working copy ...
    const Panel pn = panel(NL - 2);
    const int col0 = n.row0[NL - 2] + pn.r0;
    if (n.world > 1) {
      for (int idx = threadIdx.x; idx < n_out * pn.cnt; idx += kThreads) {
        const size_t off = (size_t)n.off_wlast + (size_t)(idx / pn.cnt) * ldl + col0 + idx % pn.cnt;
        const double v = ldg_cg(A + off);
        if (n.arena_mc) multimem_st_f64(n.arena_mc + off, v);
        else
          for (int q = 0; q < n.world; ++q)
            if (q != n.rank) stg_cg(n.arena[q] + off, v);
      }
    }
    bar.sync_all(n, true);
    // ... and the rows this rank owns back to the weight storage
    const int lr_ = NL - 1;
    const size_t tot = (size_t)n.rows[lr_] * ldl;
    const double *src = WL + (size_t)n.row0[lr_] * ldl;
    for (size_t i = (size_t)b * kThreads + threadIdx.x; i < tot; i += (size_t)grid * kThreads)
      stg_cg(n.W[lr_] + i, ldg_cg(src + i));
  } else if (b == 0) {
    evaluate(tag_of(iter), 0, 0, nullptr, A + n.off_out);  // publish the final (normalised) output
  }
  bar.finish();
  if (b == 0 && threadIdx.x == 0) n.local_sync[160] = (unsigned)launch;
  if (b == 0 && threadIdx.x == 0 && p.result) {
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

template <bool TG>
__global__ void __launch_bounds__(kThreads, 1) k_train_sample_lazy1(const __grid_constant__ KP p) {
  train_sample_lazy1_body<TG>(p, (int)gridDim.x, (int)blockIdx.x);
}

// several ranks of a row-split job inside ONE cooperative launch (see k_train_sample_vranks)
template <bool TG>
__global__ void __launch_bounds__(kThreads, 1) k_train_sample_lazy1_vranks(const KP *pv, int ctas_per_rank) {
  __shared__ KP sp;
  const int vr = (int)blockIdx.x / ctas_per_rank, bid = (int)blockIdx.x % ctas_per_rank;
  {
    const unsigned long long *src = reinterpret_cast<const unsigned long long *>(pv + vr);
    unsigned long long *dst = reinterpret_cast<unsigned long long *>(&sp);
    static_assert(sizeof(KP) % 8 == 0, "KP is copied in 8-byte words");
    for (int i = threadIdx.x; i < (int)(sizeof(KP) / 8); i += kThreads) dst[i] = src[i];
  }
  __syncthreads();
  train_sample_lazy1_body<TG>(sp, ctas_per_rank, bid);
}

int g_lazy_panel = -1;  // -1 = HPNN_LAZY_PANEL (default on), 0 = off, 1 = on
int g_slicewise = -1;   // -1 = HPNN_SLICEWISE (default off), 0 = off, 1 = on
// Eligibility + launch parameters of K-PL1 for one rank: 0 = p_out / smem_out filled, 1 = not eligible.
int prepare_lazy1(const KP &p_in, int grid, KP &p_out, size_t &smem_out) {
  static const int disabled_env = [] {
    const char *e = getenv("HPNN_LAZY_PANEL");
    return (e && !strcmp(e, "0")) ? 1 : 0;
  }();
  const int disabled = g_lazy_panel >= 0 ? !g_lazy_panel : disabled_env;
  const int depth = lazy_depth_setting();
  const hpnn_dev_net &n = p_in.net;
  const bool bpm = p_in.train == HPNN_TRAIN_BPM;
  if (disabled || depth < 2 || grid > kLazyDotSlots || p_in.mode != 0 || (p_in.train != HPNN_TRAIN_BP && !bpm) ||
      n.hist_depth < 2 || n.panel_stride <= 0 || n.n_layers < 2)
    return 1;
  KP p = p_in;
  // momentum cannot be deferred cheaply (the stored momentum matrix would need a second factor series,
  // docs/ROADMAP.md): BPM runs the same row-panel organisation with immediate updates (depth 1)
  p.lazy_depth = bpm ? 1 : (depth < n.hist_depth ? depth : n.hist_depth);
  {
    static const long long prefetch_mb = [] {
      const char *e = getenv("HPNN_PREFETCH_MB");
      return e ? atoll(e) : 0LL;  // measured at 4 GPUs: 48 MB costs 3%, so off unless asked for
    }();
    // one rank: the barrier is short and a layer is many times L2, nothing to hide behind
    p.prefetch_bytes = n.world > 1 ? prefetch_mb << 20 : 0;
  }
  int pmax = 1, rowpart = 8;
  for (int l = 0; l < n.n_layers; ++l) {
    const int pm = panel_max(n.rows[l], grid);
    pmax = pm > pmax ? pm : pmax;
  }
  pmax = (pmax + 7) & ~7;
  rowpart = pmax * kZStride;
  p.max_rows = pmax;  // the kernel's carve-up uses the panel height, not the layer height
  const size_t smem = sizeof(double) * ((size_t)p.max_ld + pmax + (size_t)n.n_layers * pmax * kLazyChunk + kRedDoubles + rowpart +
                                       pmax + 2 * kFoldMaxOut);
  if (smem > kMaxDynSmem) return 1;
  {
    static const int fold_env = [] {
      const char *e = getenv("HPNN_FOLD_OUT");
      return (e && !strcmp(e, "0")) ? 0 : 1;
    }();
    static const int depth_env = [] {
      const char *e = getenv("HPNN_ITEMS_IN_FLIGHT");
      return e ? atoi(e) : 3;  // 7-row panels: 162.7 vs 166.9 us per iteration with 2; wide panels: no difference
    }();
    p.items_in_flight = depth_env;
    static const int keep_env = [] {
      const char *e = getenv("HPNN_L2_KEEP");
      return e ? atoi(e) : -1;
    }();
    long long widest = 0;
    for (int l = 1; l < n.n_layers; ++l) {
      const long long bytes = 8LL * n.rows[l] * n.ld[l];
      widest = bytes > widest ? bytes : widest;
    }
    p.keep_l2 = keep_env >= 0 ? (keep_env ? 1 : 0) : (widest <= (80LL << 20) ? 1 : 0);
    // Measured at 8 GPUs on the wide shape (profiles/r2_probe8e_*.jsonl): 234 us/iteration against 190 with the
    // full barrier.  Every rank's flag arrives at about the same time, so only the two chunks of weights
    // that are in flight during the wait overlap it, and the chunked pass itself streams at 2.1 TB/s
    // (eight rows x 16 bytes per thread and chunk) against 4.9 TB/s.  Off unless HPNN_SLICEWISE=1.
    static const int slice_env = [] {
      const char *e = getenv("HPNN_SLICEWISE");
      return (e && !strcmp(e, "1")) ? 1 : 0;
    }();
    const int slice_on = g_slicewise >= 0 ? g_slicewise : slice_env;
    p.slicewise = (slice_on && !p.tagged && n.world > 1) ? 1 : 0;
    p.fold_out = (fold_env && !p.tagged && n.off_zpart > 0 && n.dims[n.n_layers] <= kFoldMaxOut && n.n_layers >= 3) ? 1 : 0;
  }
  p_out = p;
  smem_out = smem;
  return 0;
}

// 0 = launched, < 0 = error, 1 = not eligible.  HPNN_LAZY_PANEL=0 falls back to the column-task
// variant of train_lazy.cu.
int launch_lazy1_if_eligible(const KP &p_in, int grid, void *stream) {
  KP p;
  size_t smem = 0;
  if (prepare_lazy1(p_in, grid, p, smem) != 0) return 1;
  const void *kfn = p.tagged ? (const void *)k_train_sample_lazy1<true> : (const void *)k_train_sample_lazy1<false>;
  cudaError_t e = cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) {
    set_error("cudaFuncSetAttribute(k_train_sample_lazy1)", e);
    return -1;
  }
  void *args[] = {(void *)&p};
  e = cudaLaunchCooperativeKernel(kfn, dim3(grid), dim3(kThreads), args, smem,
                                  (cudaStream_t)stream);
  if (e != cudaSuccess) {
    set_error("cudaLaunchCooperativeKernel(k_train_sample_lazy1)", e);
    return -1;
  }
  return 0;
}

// virtual-rank launch of K-PL1 (every rank's KP already prepared, same smem for all)
int launch_lazy1_vranks(const KP *pv_dev, int world, int ctas_per_rank, bool tagged, size_t smem, void *stream) {
  const void *kfn = tagged ? (const void *)k_train_sample_lazy1_vranks<true> : (const void *)k_train_sample_lazy1_vranks<false>;
  cudaError_t e = cudaFuncSetAttribute(kfn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) {
    set_error("cudaFuncSetAttribute(k_train_sample_lazy1_vranks)", e);
    return -1;
  }
  void *args[] = {(void *)&pv_dev, (void *)&ctas_per_rank};
  e = cudaLaunchCooperativeKernel(kfn, dim3(world * ctas_per_rank), dim3(kThreads), args, smem, (cudaStream_t)stream);
  if (e != cudaSuccess) {
    set_error("cudaLaunchCooperativeKernel(k_train_sample_lazy1_vranks)", e);
    return -1;
  }
  return 0;
}

}  // namespace hpnn

extern "C" void hpnn_dev_set_lazy_panel(int on) { hpnn::g_lazy_panel = on < 0 ? -1 : (on ? 1 : 0); }
extern "C" void hpnn_dev_set_slicewise(int on) { hpnn::g_slicewise = on < 0 ? -1 : (on ? 1 : 0); }
