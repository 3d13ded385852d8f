// Device-side building blocks shared by the persistent trainer (K-P) and the
// stand-alone phase kernels (K-A..K-E).  sm_100a only.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "hpnn_core.h"

namespace hpnn {

constexpr int kThreads = 512;            // CTA size of every engine kernel
constexpr int kWarps = kThreads / 32;
constexpr int kRedDoubles = 2 * kThreads;  // shared reduction scratch (8 KB)

// ---- activation: src/ann.c:883-888, cuda_ann.cu:41-61 ----------------------
__device__ __forceinline__ double act_f(double x) { return 2.0 / (1.0 + exp(-1.0 * x)) - 1.0; }
__device__ __forceinline__ double dact_f(double y) { return -0.5 * (y * y - 1.0); }

// ---- memory-model helpers --------------------------------------------------
__device__ __forceinline__ unsigned ld_acquire_gpu(const unsigned *p) {
  unsigned v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_gpu(unsigned *p, unsigned v) {
  asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned atom_add_acqrel_gpu(unsigned *p, unsigned v) {
  unsigned old;
  asm volatile("atom.acq_rel.gpu.global.add.u32 %0, [%1], %2;" : "=r"(old) : "l"(p), "r"(v) : "memory");
  return old;
}
__device__ __forceinline__ void red_add_release_gpu(unsigned *p, unsigned v) {
  asm volatile("red.release.gpu.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long *p) {
  unsigned long long v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_release_sys(unsigned long long *p, unsigned long long v) {
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned ld_relaxed_gpu(const unsigned *p) {
  unsigned v;
  asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned long long ld_relaxed_sys(const unsigned long long *p) {
  unsigned long long v;
  asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_relaxed_sys(unsigned long long *p, unsigned long long v) {
  asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
// {value, tag} pairs are moved with the SCALAR 128-bit type (PTX ISA 8.3, ld/st .b128): a vector
// access (.v2.b64) is, in the memory model, two 64-bit accesses that merely share an instruction,
// whereas a naturally aligned scalar access is single-copy atomic -- the property the tagged
// hand-over rests on (a reader never sees a new tag with an old value).  SASS: STG.E.128 / LDG.E.128.
__device__ __forceinline__ void st_b128_relaxed_sys(void *p, unsigned long long lo, unsigned long long hi) {
  asm volatile("{\n\t.reg .b128 q;\n\tmov.b128 q, {%1, %2};\n\tst.relaxed.sys.global.b128 [%0], q;\n\t}" ::"l"(p), "l"(lo), "l"(hi) : "memory");
}
__device__ __forceinline__ void st_b128_relaxed_gpu(void *p, unsigned long long lo, unsigned long long hi) {
  asm volatile("{\n\t.reg .b128 q;\n\tmov.b128 q, {%1, %2};\n\tst.relaxed.gpu.global.b128 [%0], q;\n\t}" ::"l"(p), "l"(lo), "l"(hi) : "memory");
}
__device__ __forceinline__ void ld_b128_relaxed_sys(const void *p, unsigned long long &lo, unsigned long long &hi) {
  asm volatile("{\n\t.reg .b128 q;\n\tld.relaxed.sys.global.b128 q, [%2];\n\tmov.b128 {%0, %1}, q;\n\t}" : "=l"(lo), "=l"(hi) : "l"(p) : "memory");
}
__device__ __forceinline__ void ld_b128_relaxed_gpu(const void *p, unsigned long long &lo, unsigned long long &hi) {
  asm volatile("{\n\t.reg .b128 q;\n\tld.relaxed.gpu.global.b128 q, [%2];\n\tmov.b128 {%0, %1}, q;\n\t}" : "=l"(lo), "=l"(hi) : "l"(p) : "memory");
}
// NVLS: one store to the multicast address of a symmetric allocation (hpnn_symm_mc_ptr) is replicated
// by the NVSwitch into every rank's copy; a reduction is applied to every copy; a load-reduce returns
// the sum over the copies.  The address decides the routing (SASS shows plain STG / REDG / LDGMC).
__device__ __forceinline__ void multimem_st_f64(double *mc, double v) {
  asm volatile("multimem.st.relaxed.sys.global.f64 [%0], %1;" ::"l"(mc), "d"(v) : "memory");
}
__device__ __forceinline__ void multimem_st_u64(unsigned long long *mc, unsigned long long v) {
  asm volatile("multimem.st.relaxed.sys.global.b64 [%0], %1;" ::"l"(mc), "l"(v) : "memory");
}
__device__ __forceinline__ void multimem_red_add_u64(unsigned long long *mc, unsigned long long v) {
  asm volatile("multimem.red.relaxed.sys.global.add.u64 [%0], %1;" ::"l"(mc), "l"(v) : "memory");
}
__device__ __forceinline__ void red_add_relaxed_sys_u64(unsigned long long *p, unsigned long long v) {
  asm volatile("red.relaxed.sys.global.add.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ void multimem_red_add_f64(double *mc, double v) {
  asm volatile("multimem.red.relaxed.sys.global.add.f64 [%0], %1;" ::"l"(mc), "d"(v) : "memory");
}
__device__ __forceinline__ double multimem_ld_reduce_add_f64(const double *mc) {
  double v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.f64 %0, [%1];" : "=d"(v) : "l"(mc) : "memory");
  return v;
}
__device__ __forceinline__ void fence_sys() { asm volatile("fence.acq_rel.sys;" ::: "memory"); }
__device__ __forceinline__ void fence_gpu() { asm volatile("fence.acq_rel.gpu;" ::: "memory"); }

// L2-coherent (L1-bypassing) accesses for everything another CTA / GPU rewrites
// while a persistent kernel is running.
__device__ __forceinline__ double ldg_cg(const double *p) { return __ldcg(p); }
__device__ __forceinline__ double2 ldg_cg2(const double2 *p) { return __ldcg(p); }
__device__ __forceinline__ void stg_cg(double *p, double v) { __stcg(p, v); }
__device__ __forceinline__ void stg_cg2(double2 *p, double2 v) { __stcg(p, v); }

// ---- block-wide reductions with a fixed summation order ---------------------
// Every thread of the CTA gets the same value; `red` needs kWarps doubles.
__device__ __forceinline__ double block_sum(double v, double *red) {
#pragma unroll
  for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  const int w = threadIdx.x >> 5;
  __syncthreads();  // protect `red` from the previous use
  if ((threadIdx.x & 31) == 0) red[w] = v;
  __syncthreads();
  double s = 0.0;
#pragma unroll
  for (int i = 0; i < kWarps; ++i) s += red[i];
  return s;
}
__device__ __forceinline__ double block_max(double v, double *red) {
#pragma unroll
  for (int o = 16; o; o >>= 1) v = fmax(v, __shfl_xor_sync(0xffffffffu, v, o));
  const int w = threadIdx.x >> 5;
  __syncthreads();
  if ((threadIdx.x & 31) == 0) red[w] = v;
  __syncthreads();
  double s = red[0];
#pragma unroll
  for (int i = 1; i < kWarps; ++i) s = fmax(s, red[i]);
  return s;
}
// first index holding the largest value (ties -> lowest index); NaN never wins
__device__ __forceinline__ int block_argmax(double v, int idx, double *red) {
  if (!(v == v)) v = -INFINITY;
#pragma unroll
  for (int o = 16; o; o >>= 1) {
    double ov = __shfl_xor_sync(0xffffffffu, v, o);
    int oi = __shfl_xor_sync(0xffffffffu, idx, o);
    if (ov > v || (ov == v && oi < idx)) {
      v = ov;
      idx = oi;
    }
  }
  const int w = threadIdx.x >> 5;
  int *redi = reinterpret_cast<int *>(red + kWarps);
  __syncthreads();
  if ((threadIdx.x & 31) == 0) {
    red[w] = v;
    redi[w] = idx;
  }
  __syncthreads();
  double bv = red[0];
  int bi = redi[0];
#pragma unroll
  for (int i = 1; i < kWarps; ++i) {
    double ov = red[i];
    int oi = redi[i];
    if (ov > bv || (ov == bv && oi < bi)) {
      bv = ov;
      bi = oi;
    }
  }
  // reference scan starts from probe=-1.0 with a strict '<' (src/ann.c:2346-2351)
  return (bv > -1.0) ? bi : 0;
}
__device__ __forceinline__ int block_max_int(int v, double *red) {
#pragma unroll
  for (int o = 16; o; o >>= 1) v = max(v, __shfl_xor_sync(0xffffffffu, v, o));
  int *redi = reinterpret_cast<int *>(red);
  const int w = threadIdx.x >> 5;
  __syncthreads();
  if ((threadIdx.x & 31) == 0) redi[w] = v;
  __syncthreads();
  int s = redi[0];
#pragma unroll
  for (int i = 1; i < kWarps; ++i) s = max(s, redi[i]);
  return s;
}

// copy n doubles (n % 2 == 0, 16-byte aligned) global -> shared, bypassing L1
// Eight loads per thread are issued before the first one is consumed: an 8192-wide vector costs one
// L2 round trip instead of eight (measured on the wide shape: 5.8 k -> cycles per staging, and every
// phase of every iteration starts with one).
__device__ __forceinline__ void stage_vec(double *dst, const double *src, int n) {
  const double2 *s2 = reinterpret_cast<const double2 *>(src);
  double2 *d2 = reinterpret_cast<double2 *>(dst);
  const int n2 = n >> 1;
  constexpr int U = 8;
  for (int base = threadIdx.x; base < n2; base += U * (int)blockDim.x) {
    double2 v[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int i = base + u * (int)blockDim.x;
      v[u] = i < n2 ? ldg_cg2(s2 + i) : make_double2(0.0, 0.0);
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int i = base + u * (int)blockDim.x;
      if (i < n2) d2[i] = v[u];
    }
  }
}

// where a forward result goes: the same offset in up to HPNN_MAX_WORLD buffers
// (all-gather by peer stores when n > 1)
struct Dst {
  double *p[HPNN_MAX_WORLD];
  int n;
  unsigned long long tag;  // != 0: publish {value, tag} pairs (tagged hand-over, see kp_common.cuh)
  double *mc;              // != nullptr (plain buffers only): the same offset at the multicast address
};
__device__ __forceinline__ void dst_store(const Dst &dst, int r, double y) {
  if (!dst.tag && dst.mc) {  // all-gather in ONE store: the switch writes every rank's copy
    multimem_st_f64(dst.mc + r, y);
    return;
  }
  for (int q = 0; q < dst.n; ++q) {
    if (dst.tag)
      st_b128_relaxed_sys(dst.p[q] + 2 * (size_t)r, (unsigned long long)__double_as_longlong(y), dst.tag);
    else
      stg_cg(dst.p[q] + r, y);
  }
}

// ---------------------------------------------------------------------------
// Row tasks (K-A / K-D / K-E): one warp per owned row.
//   UPD 0: y = f(w.x)
//   UPD 1: w += (lr*delta_r) * x ; y = f(w.x)                       (BP)
//   UPD 2: m += (lr*delta_r) * x ; w += m ; m *= alpha ; y = f(w.x)  (BPM)
// xs: input vector in shared memory (ld doubles, pad = 0); ds: local deltas.
// ---------------------------------------------------------------------------
template <int UPD>
__device__ __forceinline__ void row_apply(double2 &w, double2 &m, const double2 x, double s, double alpha,
                                          double2 *wp, double2 *mp, double &a) {
  if (UPD == 1) {
    w.x = fma(s, x.x, w.x);
    w.y = fma(s, x.y, w.y);
    stg_cg2(wp, w);
  } else if (UPD == 2) {
    m.x = fma(s, x.x, m.x);
    m.y = fma(s, x.y, m.y);
    w.x += m.x;
    w.y += m.y;
    stg_cg2(wp, w);
    m.x *= alpha;
    m.y *= alpha;
    stg_cg2(mp, m);
  }
  a = fma(w.x, x.x, a);
  a = fma(w.y, x.y, a);
}

// One warp, one row.  The row is walked in chunks of 4 x 32 double2; the loads of chunk
// k+1 are issued before chunk k is consumed, so a 1000-wide row costs ~one L2 round trip
// instead of one per chunk (the small nets are latency bound) and a wide
// row keeps 8 x 512 B per warp in flight.
template <int UPD>
__device__ __forceinline__ double warp_row(double *wrow, double *mrow, const double *xs, int ld, double s,
                                           double alpha, int lane) {
  double2 *w2 = reinterpret_cast<double2 *>(wrow);
  double2 *m2 = reinterpret_cast<double2 *>(mrow);
  const double2 *x2 = reinterpret_cast<const double2 *>(xs);
  const int n2 = ld >> 1;
  const double2 zero2 = make_double2(0.0, 0.0);
  double a0 = 0.0, a1 = 0.0, a2 = 0.0, a3 = 0.0;
  double2 wa[4], ma[4], wb[4], mb[4];
  auto load = [&](double2 (&w)[4], double2 (&m)[4], int base) {
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int i = base + lane + 32 * u;
      w[u] = (i < n2) ? ldg_cg2(w2 + i) : zero2;
      if (UPD == 2) m[u] = (i < n2) ? ldg_cg2(m2 + i) : zero2;
    }
  };
  auto consume = [&](double2 (&w)[4], double2 (&m)[4], int base) {
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int i = base + lane + 32 * u;
      if (i < n2) {
        double &a = (u == 0) ? a0 : (u == 1) ? a1 : (u == 2) ? a2 : a3;
        row_apply<UPD>(w[u], m[u], x2[i], s, alpha, w2 + i, m2 + i, a);
      }
    }
  };
  load(wa, ma, 0);
  for (int base = 0; base < n2; base += 256) {
    load(wb, mb, base + 128);
    consume(wa, ma, base);
    load(wa, ma, base + 256);
    consume(wb, mb, base + 128);
  }
  double a = (a0 + a1) + (a2 + a3);
#pragma unroll
  for (int o = 16; o; o >>= 1) a += __shfl_xor_sync(0xffffffffu, a, o);
  return a;
}

// Segments per row chosen so that (rows of this CTA) x segments divides evenly over the 16
// warps: 8192 rows on 148 CTAs are 55.4 rows per CTA = 3.5 rows per warp, i.e. every fourth
// round is half empty; two half rows per row make it 6.9 -> 7 even rounds.
__host__ __device__ inline int pick_row_segments(int rows, int ld, int grid) {
  const int rpc = (rows + grid - 1) / grid;
  int best = 1;
  long long best_cost = -1;
  for (int sg = 1; sg <= 8; sg *= 2) {
    if (sg > 1 && ld / sg < 256) break;  // keep segments >= 2 KB
    const long long cost = (long long)((rpc * sg + kWarps - 1) / kWarps) * ((ld + sg - 1) / sg);
    if (best_cost < 0 || cost < best_cost) {
      best_cost = cost;
      best = sg;
    }
  }
  return best;
}
__host__ __device__ inline int row_part_doubles(int rows, int grid) { return ((rows + grid - 1) / grid) * 8; }

// all owned rows of one layer, spread over the whole grid: row r belongs to CTA r % grid, its
// `segs` segments are dealt round-robin to the warps; partial dot products meet in `rowpart`
// (shared, row_part_doubles() doubles) and are added in segment order.  Result r goes to
// dst.p[q][r] for every q (caller pre-offsets the pointers to the owned slice).
template <int UPD>
__device__ __forceinline__ void phase_rows(double *W, double *dW, int ld, int rows, const double *xs,
                                           const double *ds, double lr, double alpha, int act_kind,
                                           const Dst &dst, int segs, double *rowpart, int grid, int b) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int mine = b < rows ? (rows - b + grid - 1) / grid : 0;  // rows b, b+grid, ...
  const int n2 = ld >> 1;
  const int seg2 = (n2 + segs - 1) / segs;  // double2 per segment
  for (int u = warp; u < mine * segs; u += kWarps) {
    const int i = u / segs, sg = u % segs;
    const int r = b + i * grid;
    const int c0 = sg * seg2, c1 = min(n2, c0 + seg2);
    double *wrow = W + (size_t)r * ld + 2 * c0;
    double *mrow = (UPD == 2) ? dW + (size_t)r * ld + 2 * c0 : nullptr;
    const double s = UPD ? lr * ds[r] : 0.0;
    const double z = c1 > c0 ? warp_row<UPD>(wrow, mrow, xs + 2 * c0, 2 * (c1 - c0), s, alpha, lane) : 0.0;
    if (lane == 0) rowpart[i * 8 + sg] = z;
  }
  __syncthreads();
  for (int i = threadIdx.x; i < mine; i += blockDim.x) {
    double z = 0.0;
    for (int sg = 0; sg < segs; ++sg) z += rowpart[i * 8 + sg];
    const double y = act_kind == 0 ? act_f(z) : z;
    const int r = b + i * grid;
    dst_store(dst, r, y);
  }
}

// ---------------------------------------------------------------------------
// Column tasks (K-C fused with K-D/K-E): one CTA owns CB columns of the local
// row block, reads every row segment once, accumulates W^T.delta from the OLD
// weights and applies the BP/BPM update in the same pass.
// The CB partial sums are handed to `sink(col_in_block, value)` by threads < CB.
// ---------------------------------------------------------------------------
template <int CB, int UPD, typename Sink>
__device__ __forceinline__ void col_task(double *W, double *dW, int ld, int row_begin, int rows, int c0, const double *ds,
                                         const double *h, bool h_shared, double lr, double alpha, double *red, Sink sink) {
  constexpr int TPR = CB / 2;          // threads per row segment (double2 each)
  constexpr int RL = kThreads / TPR;   // row lanes
  const int cp = threadIdx.x % TPR, rl = threadIdx.x / TPR;
  const int c = c0 + 2 * cp;
  double h0 = 0.0, h1 = 0.0;
  if (UPD) {
    h0 = h_shared ? h[c] : ldg_cg(h + c);
    h1 = h_shared ? h[c + 1] : ldg_cg(h + c + 1);
  }
  double a0 = 0.0, a1 = 0.0;
  const size_t stride2 = (size_t)(ld >> 1);
  double2 *Wp = reinterpret_cast<double2 *>(W + c);
  double2 *Mp = reinterpret_cast<double2 *>((UPD == 2 ? dW : W) + c);
  const double2 zero2 = make_double2(0.0, 0.0);
  double2 wa[4], ma[4], wb[4], mb[4];
  auto load = [&](double2 (&w)[4], double2 (&m)[4], int r0) {
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int r = r0 + u * RL;
      w[u] = (r < rows) ? ldg_cg2(Wp + (size_t)r * stride2) : zero2;
      if (UPD == 2) m[u] = (r < rows) ? ldg_cg2(Mp + (size_t)r * stride2) : zero2;
    }
  };
  auto consume = [&](double2 (&w)[4], double2 (&m)[4], int r0) {
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int r = r0 + u * RL;
      if (r < rows) {
        const double d = ds[r];
        a0 = fma(w[u].x, d, a0);
        a1 = fma(w[u].y, d, a1);
        if (UPD == 1) {
          const double s = lr * d;
          w[u].x = fma(s, h0, w[u].x);
          w[u].y = fma(s, h1, w[u].y);
          stg_cg2(Wp + (size_t)r * stride2, w[u]);
        } else if (UPD == 2) {
          const double s = lr * d;
          m[u].x = fma(s, h0, m[u].x);
          m[u].y = fma(s, h1, m[u].y);
          w[u].x += m[u].x;
          w[u].y += m[u].y;
          stg_cg2(Wp + (size_t)r * stride2, w[u]);
          m[u].x *= alpha;
          m[u].y *= alpha;
          stg_cg2(Mp + (size_t)r * stride2, m[u]);
        }
      }
    }
  };
  // two batches of 4 rows in flight per thread (software pipeline, see warp_row)
  load(wa, ma, row_begin + rl);
  for (int r0 = row_begin + rl; r0 < rows; r0 += 8 * RL) {
    load(wb, mb, r0 + 4 * RL);
    consume(wa, ma, r0);
    load(wa, ma, r0 + 8 * RL);
    consume(wb, mb, r0 + 4 * RL);
  }
  // row lanes inside a warp share a column pair when lane % TPR matches
#pragma unroll
  for (int o = 16; o >= TPR; o >>= 1) {
    a0 += __shfl_xor_sync(0xffffffffu, a0, o);
    a1 += __shfl_xor_sync(0xffffffffu, a1, o);
  }
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  __syncthreads();  // `red` may still be read by the previous task
  if (lane < TPR && lane < 32) {
    red[warp * CB + 2 * lane] = a0;
    red[warp * CB + 2 * lane + 1] = a1;
  }
  __syncthreads();
  if (threadIdx.x < CB) {
    double s = 0.0;
    // with TPR == 32 (CB == 64) every warp holds one row lane; otherwise kWarps partials
#pragma unroll
    for (int w = 0; w < kWarps; ++w) s += red[w * CB + threadIdx.x];
    sink(threadIdx.x, s);
  }
}

// One task = (column block, row part): the owned rows are cut into `rsplit` contiguous parts so
// that the task count divides evenly over the grid (512 column blocks on 148 CTAs are 3.46
// rounds; with two row parts 6.9 -> 7 even rounds).  make_sink(c0, row_part) says where the CB
// partial sums of that task go; the consumer adds the row parts in order.
template <int UPD, typename SinkFactory>
__device__ __forceinline__ void phase_cols(double *W, double *dW, int ld, int rows, int cb, int rsplit, const double *ds,
                                           const double *h, bool h_shared, double lr, double alpha, double *red,
                                           SinkFactory make_sink, int grid, int b) {
  const int ncb = ld / cb;
  const int ntasks = ncb * rsplit;
  const int part_rows = (rows + rsplit - 1) / rsplit;
  for (int t = b; t < ntasks; t += grid) {
    const int c0 = (t % ncb) * cb, rp = t / ncb;
    const int rb = rp * part_rows, re = min(rows, rb + part_rows);
    auto sink = make_sink(c0, rp);
    if (cb == 8) col_task<8, UPD>(W, dW, ld, rb, re, c0, ds, h, h_shared, lr, alpha, red, sink);
    else if (cb == 16) col_task<16, UPD>(W, dW, ld, rb, re, c0, ds, h, h_shared, lr, alpha, red, sink);
    else col_task<32, UPD>(W, dW, ld, rb, re, c0, ds, h, h_shared, lr, alpha, red, sink);
  }
}

struct ColPlan {
  int cb, rsplit;
};
// cheapest (column block, row split) for a layer: rounds x work per task, narrow blocks
// penalised (64-byte row segments use DRAM pages poorly), row split capped by `max_rsplit`
__host__ __device__ inline ColPlan pick_col_plan(int ld, int rows, int grid, int max_rsplit) {
  ColPlan best = {8, 1};
  double best_cost = -1.0;
  const int cbs[3] = {32, 16, 8};
  const double pen[3] = {0.97, 1.0, 1.12};
  for (int ci = 0; ci < 3; ++ci) {
    const int cb = cbs[ci];
    if (ld % cb) continue;
    for (int rs = 1; rs <= max_rsplit; rs *= 2) {
      if (rs > 1 && rows / rs < 256) break;
      const long long tasks = (long long)(ld / cb) * rs;
      const double cost = (double)((tasks + grid - 1) / grid) * cb * ((rows + rs - 1) / rs) * pen[ci];
      if (best_cost < 0 || cost < best_cost * 0.999) {
        best_cost = cost;
        best.cb = cb;
        best.rsplit = rs;
      }
    }
  }
  return best;
}
__host__ __device__ inline int pick_col_block(int ld, int grid) { return pick_col_plan(ld, 1 << 30, grid, 1).cb; }

// ---------------------------------------------------------------------------
// Output statistics (K-B), computed redundantly by every CTA from the gathered
// output vector: softmax (SNN), error, argmax, target index, output deltas.
// ---------------------------------------------------------------------------
struct OutStats {
  double err;
  int argmax;
  int target;
};

// z: gathered output (ANN: activated, SNN: raw).  For SNN the normalised output
// is exp(z-1)/(TINY+sum) (src/snn.c:282-334) unless stable.  ds receives the
// deltas of rows [row0,row0+rows) when non-null; o_out (global) the outputs.
// Small output layers (n <= 32, the MNIST shape has 10): every warp evaluates the whole
// vector redundantly with shuffles only, so the per-iteration evaluation costs one L2 round
// trip and a single __syncthreads instead of ~10 block-wide reductions.
// ZS: z lives in shared memory (already gathered by the caller) instead of global memory
template <bool ZS>
__device__ __forceinline__ double ld_z(const double *p) {
  return ZS ? *p : ldg_cg(p);
}

template <bool ZS>
__device__ __forceinline__ OutStats output_stats_small(int type, int stable, const double *z, const double *t,
                                                       int n, int row0, int rows, double *ds, double *o_out) {
  const int lane = threadIdx.x & 31;
  const bool in = lane < n;
  double o = in ? ld_z<ZS>(z + lane) : 0.0;
  const double tj = in ? ldg_cg(t + lane) : 0.0;
  if (type == HPNN_TYPE_SNN) {
    double shift = 1.0;
    if (stable) {
      double m = in ? o : -INFINITY;
#pragma unroll
      for (int k = 16; k; k >>= 1) m = fmax(m, __shfl_xor_sync(0xffffffffu, m, k));
      shift = m;
    }
    const double e = in ? exp(o - shift) : 0.0;
    double dv = e;  // butterfly sum: every lane ends with the same value
#pragma unroll
    for (int k = 16; k; k >>= 1) dv += __shfl_xor_sync(0xffffffffu, dv, k);
    dv += HPNN_TINY;
    o = e / dv;
  }
  double term, d;
  if (type == HPNN_TYPE_SNN) {
    term = (in && o > 0.) ? tj * log(o + HPNN_TINY) : 0.0;
    d = tj - o;
  } else {
    term = in ? (tj - o) * (tj - o) : 0.0;
    d = (tj - o) * dact_f(o);
  }
  double e = term;
#pragma unroll
  for (int k = 16; k; k >>= 1) e += __shfl_xor_sync(0xffffffffu, e, k);
  double bv = (in && o == o) ? o : -INFINITY;
  int bi = in ? lane : 0x7fffffff;
  int trg = (in && tj == 1.0) ? lane : -1;
#pragma unroll
  for (int k = 16; k; k >>= 1) {
    const double ov = __shfl_xor_sync(0xffffffffu, bv, k);
    const int oi = __shfl_xor_sync(0xffffffffu, bi, k);
    if (ov > bv || (ov == bv && oi < bi)) {
      bv = ov;
      bi = oi;
    }
    trg = max(trg, __shfl_xor_sync(0xffffffffu, trg, k));
  }
  if (threadIdx.x < 32 && in) {
    if (ds && lane >= row0 && lane < row0 + rows) ds[lane - row0] = d;
    if (o_out) stg_cg(o_out + lane, o);
  }
  OutStats st;
  st.err = (type == HPNN_TYPE_SNN) ? e * (-1.0 / n) : 0.5 * e;
  st.argmax = (bv > -1.0) ? bi : 0;
  st.target = trg < 0 ? 0 : trg;
  __syncthreads();
  return st;
}

template <bool ZS = false>
__device__ __forceinline__ OutStats output_stats(int type, int stable, const double *z, const double *t,
                                                 int n, int row0, int rows, double *ds, double *o_out,
                                                 double *red) {
  if (n <= 32) return output_stats_small<ZS>(type, stable, z, t, n, row0, rows, ds, o_out);
  OutStats st;
  double shift = 1.0, dv = 1.0;
  if (type == HPNN_TYPE_SNN) {
    if (stable) {
      double m = -INFINITY;
      for (int j = threadIdx.x; j < n; j += blockDim.x) m = fmax(m, ld_z<ZS>(z + j));
      shift = block_max(m, red);
    }
    double s = 0.0;
    for (int j = threadIdx.x; j < n; j += blockDim.x) s += exp(ld_z<ZS>(z + j) - shift);
    dv = block_sum(s, red) + HPNN_TINY;
  }
  double e = 0.0, best = -INFINITY;
  int besti = 0x7fffffff, trg = -1;
  for (int j = threadIdx.x; j < n; j += blockDim.x) {
    double o = ld_z<ZS>(z + j);
    if (type == HPNN_TYPE_SNN) o = exp(o - shift) / dv;
    const double tj = ldg_cg(t + j);
    double d;
    if (type == HPNN_TYPE_SNN) {
      if (o > 0.) e += tj * log(o + HPNN_TINY);
      d = tj - o;
    } else {
      e += (tj - o) * (tj - o);
      d = (tj - o) * dact_f(o);
    }
    if (ds && j >= row0 && j < row0 + rows) ds[j - row0] = d;
    if (o_out) stg_cg(o_out + j, o);
    if (o > best) {  // ascending j per thread: keeps the first maximum
      best = o;
      besti = j;
    }
    if (tj == 1.0) trg = j;
  }
  e = block_sum(e, red);
  st.err = (type == HPNN_TYPE_SNN) ? e * (-1.0 / n) : 0.5 * e;
  st.argmax = block_argmax(best, besti, red);
  trg = block_max_int(trg, red);
  st.target = trg < 0 ? 0 : trg;
  __syncthreads();
  return st;
}

}  // namespace hpnn
