// CPU / OpenMP engine: the fallback back-end and the oracle every GPU kernel is
// tested against.  It restates the numerical spec of the reference:
//   activation   src/ann.c:883-888        forward     src/ann.c:892-1242, src/snn.c:79-443
//   errors       src/ann.c:1246-1275, src/snn.c:447-477
//   deltas       src/ann.c:1279-1592, src/snn.c:481-794
//   BP / BPM     src/ann.c:1596-1872, 1943-2277
//   sample loops src/ann.c:2281-2467, src/snn.c:1417-1595
// written as one generic L-layer loop instead of the reference's per-BLAS-flavour copies.
#include <math.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#include <algorithm>
#endif

#include <vector>

#include "hpnn_core.h"
#include "kernel_host.hpp"

namespace {

int g_threads = 0;

inline double act(double x) { return 2.0 / (1.0 + exp(-1.0 * x)) - 1.0; }
inline double dact(double y) { return -0.5 * (y * y - 1.0); }

void ensure_workspace(hpnn_kernel *k, bool momentum) {
  const int L = k->n_layers();
  if ((int)k->act.size() != L + 1) {
    k->act.resize(L + 1);
    for (int i = 0; i <= L; ++i) k->act[i].assign(k->dims[i], 0.0);
    k->delta.resize(L);
    for (int l = 0; l < L; ++l) k->delta[l].assign(k->dims[l + 1], 0.0);
    k->hold.resize(L);
    for (int l = 0; l < L; ++l) k->hold[l].assign(k->dims[l], 0.0);
  }
  if (momentum && (int)k->dW.size() != L) {
    k->dW.resize(L);
    for (int l = 0; l < L; ++l) k->dW[l].assign(k->W[l].size(), 0.0);
  }
}

// Dot products of `rows` consecutive rows with x, every row summed strictly left to right as the
// reference does (src/ann.c:892-1242 without BLAS), so the results are bit-identical to it.  The
// speed comes from walking 8 rows at once: 8 independent add chains instead of one that waits
// 4 cycles for every element.
template <int R>
inline void dot_rows(const double *w, int M, const double *x, double *s) {
  double a[R];
  for (int q = 0; q < R; ++q) a[q] = 0.0;
  for (int c = 0; c < M; ++c) {
    const double xc = x[c];
    for (int q = 0; q < R; ++q) a[q] += w[(size_t)q * M + c] * xc;
  }
  for (int q = 0; q < R; ++q) s[q] = a[q];
}

// The phase functions below use ORPHANED worksharing (`omp for`): called from inside the
// training loop's parallel region they split the work over the team and end with a barrier;
// called from serial code (hpnn_cpu_forward, the single-step entry points) they run on one thread.
void forward_layer(hpnn_kernel *k, int type, int l) {
  const int L = k->n_layers();
  const int N = k->dims[l + 1], M = k->dims[l];
  const double *W = k->W[l].data();
  const double *x = k->act[l].data();
  double *y = k->act[l + 1].data();
  const bool raw = (type == HPNN_TYPE_SNN && l == L - 1);
  const int nblk = (N + 7) / 8;
#pragma omp for schedule(static)
  for (int b = 0; b < nblk; ++b) {
    const int j0 = b * 8, n = N - j0 < 8 ? N - j0 : 8;
    double s[8];
    if (n == 8) dot_rows<8>(W + (size_t)j0 * M, M, x, s);
    else
      for (int q = 0; q < n; ++q) dot_rows<1>(W + (size_t)(j0 + q) * M, M, x, s + q);
    for (int q = 0; q < n; ++q) y[j0 + q] = raw ? s[q] : act(s[q]);
  }
}

// softmax of the raw output layer; every thread of a team may call it on its own copy of the
// loop state, so it writes only through `single`
void softmax_output(hpnn_kernel *k, int stable) {
  const int L = k->n_layers();
#pragma omp single
  {
    std::vector<double> &o = k->act[L];
    const int N = k->dims[L];
    double shift = 1.0;
    if (stable) {
      shift = o[0];
      for (int j = 1; j < N; ++j) shift = o[j] > shift ? o[j] : shift;
    }
    double dv = HPNN_TINY;
    for (int j = 0; j < N; ++j) {
      o[j] = exp(o[j] - shift);
      dv += o[j];
    }
    for (int j = 0; j < N; ++j) o[j] /= dv;
  }  // implicit barrier
}

void forward(hpnn_kernel *k, int type, int stable) {
  const int L = k->n_layers();
  for (int l = 0; l < L; ++l) forward_layer(k, type, l);
  if (type == HPNN_TYPE_SNN) softmax_output(k, stable);
}

double error(const hpnn_kernel *k, int type, const double *t) {
  const int L = k->n_layers(), N = k->dims[L];
  const double *o = k->act[L].data();
  double E = 0.0;
  if (type == HPNN_TYPE_SNN) {
    for (int j = 0; j < N; ++j)
      if (o[j] > 0.) E += t[j] * log(o[j] + HPNN_TINY);
    return E * (-1.0 / N);
  }
  for (int j = 0; j < N; ++j) E += (t[j] - o[j]) * (t[j] - o[j]);
  return 0.5 * E;
}

// all deltas from the OLD weights, top-down.  Column c is summed over the rows in ascending
// order exactly as the reference's transposed GEMV does (src/ann.c:1279-1592); walking the rows in
// the outer loop makes the inner loop contiguous (and vectorisable) without changing that order.
void deltas(hpnn_kernel *k, int type, const double *t) {
  const int L = k->n_layers();
#pragma omp single
  {
    const int N = k->dims[L];
    const double *o = k->act[L].data();
    double *d = k->delta[L - 1].data();
    for (int j = 0; j < N; ++j) d[j] = (type == HPNN_TYPE_SNN) ? (t[j] - o[j]) : (t[j] - o[j]) * dact(o[j]);
  }  // implicit barrier
  for (int l = L - 1; l >= 1; --l) {
    const int N = k->dims[l + 1], M = k->dims[l];
    const double *W = k->W[l].data();
    const double *dn = k->delta[l].data();
    const double *h = k->act[l].data();
    double *d = k->delta[l - 1].data();
    const int CB = 256;  // columns per task: 2 KB of accumulators stay in L1
    const int nblk = (M + CB - 1) / CB;
#pragma omp for schedule(static)
    for (int b = 0; b < nblk; ++b) {
      const int c0 = b * CB, n = M - c0 < CB ? M - c0 : CB;
      double s[CB];
      for (int c = 0; c < n; ++c) s[c] = 0.0;
      for (int r = 0; r < N; ++r) {
        const double dr = dn[r];
        const double *w = W + (size_t)r * M + c0;
        for (int c = 0; c < n; ++c) s[c] += w[c] * dr;
      }
      for (int c = 0; c < n; ++c) d[c0 + c] = s[c] * dact(h[c0 + c]);
    }  // implicit barrier
  }
}


// BP / BPM update of R consecutive rows fused with their forward dot products: one pass over the
// weights instead of two.  Element by element this is the reference's update (src/ann.c:1596-1872,
// 1943-2277) followed by its left-to-right dot product, so every value is bit-identical; h is the
// layer's input of the PREVIOUS sweep (what the update needs), x the new one.
template <int R, bool MOM>
inline void update_dot_rows(double *w, double *m, int M, const double *sc, const double *h, const double *x,
                            double alpha, double *out) {
  double a[R];
  for (int q = 0; q < R; ++q) a[q] = 0.0;
  constexpr int CB = 256;  // R rows x 2 KB stay in L1 between the two sweeps of a block
  for (int c0 = 0; c0 < M; c0 += CB) {
    const int c1 = c0 + CB < M ? c0 + CB : M;
    for (int q = 0; q < R; ++q) {  // elementwise: vectorised by the compiler
      double *wq = w + (size_t)q * M;
      const double s = sc[q];
      if (MOM) {
        double *mq = m + (size_t)q * M;
        for (int c = c0; c < c1; ++c) {
          const double mv = mq[c] + s * h[c];
          wq[c] += mv;
          mq[c] = mv * alpha;
        }
      } else {
        for (int c = c0; c < c1; ++c) wq[c] += s * h[c];
      }
    }
    for (int c = c0; c < c1; ++c) {  // R independent left-to-right chains
      const double xc = x[c];
      for (int q = 0; q < R; ++q) a[q] += w[(size_t)q * M + c] * xc;
    }
  }
  for (int q = 0; q < R; ++q) out[q] = a[q];
}

void update_forward(hpnn_kernel *k, int type, int stable, bool momentum, double lr, double alpha) {
  const int L = k->n_layers();
#pragma omp single
  for (int l = 1; l < L; ++l) k->hold[l] = k->act[l];  // same sizes: plain copies, no allocation
  // (implicit barrier)
  for (int l = 0; l < L; ++l) {
    const int N = k->dims[l + 1], M = k->dims[l];
    double *W = k->W[l].data();
    double *dW = momentum ? k->dW[l].data() : nullptr;
    const double *d = k->delta[l].data();
    const double *h = l == 0 ? k->act[0].data() : k->hold[l].data();
    const double *x = k->act[l].data();
    double *y = k->act[l + 1].data();
    const bool raw = (type == HPNN_TYPE_SNN && l == L - 1);
    const int nblk = (N + 7) / 8;
#pragma omp for schedule(static)
    for (int b = 0; b < nblk; ++b) {
      const int j0 = b * 8, n = N - j0 < 8 ? N - j0 : 8;
      double sc[8], s[8];
      for (int q = 0; q < n; ++q) sc[q] = lr * d[j0 + q];
      double *w = W + (size_t)j0 * M;
      double *m = momentum ? dW + (size_t)j0 * M : nullptr;
      if (n == 8) {
        if (momentum) update_dot_rows<8, true>(w, m, M, sc, h, x, alpha, s);
        else update_dot_rows<8, false>(w, m, M, sc, h, x, alpha, s);
      } else {
        for (int q = 0; q < n; ++q) {
          if (momentum) update_dot_rows<1, true>(w + (size_t)q * M, m + (size_t)q * M, M, sc + q, h, x, alpha, s + q);
          else update_dot_rows<1, false>(w + (size_t)q * M, nullptr, M, sc + q, h, x, alpha, s + q);
        }
      }
      for (int q = 0; q < n; ++q) y[j0 + q] = raw ? s[q] : act(s[q]);
    }  // implicit barrier: the next layer reads y
  }
  if (type == HPNN_TYPE_SNN) softmax_output(k, stable);
}

// ---------------------------------------------------------------------------------------
// Packed engine for the per-sample convergence loop.
//
// The loop spends its time in two operations per layer: the elementwise update and one
// left-to-right dot product per row.  The dot product is a dependent add chain per row, so the
// only way to go faster without changing a single rounding is to run many rows side by side.
// For the duration of a sample the weights are therefore kept with 8 rows interleaved per
// column, P[(b*M + c)*8 + q] = W[(8b+q)*M + c]: one 32-byte vector holds column c of four
// neighbouring rows, `acc += w * x[c]` advances four row sums at once, and 4 blocks (32 rows,
// 8 vectors) in flight hide the add latency.  Update and dot product share one pass over the
// data.  Every row still sees exactly the reference's operations in the reference's order
// (src/ann.c:1596-1872: w += (lr*d)*h, then s += w*x for c = 0..M-1), mul and add separately
// rounded: the targets below enable AVX2 but not FMA.
// ---------------------------------------------------------------------------------------
namespace packed {

constexpr int RB = 8;
// Naturally aligned vector types (attributes such as aligned(8) are dropped from template arguments, so the
// packed weights -- 16-byte aligned at best -- are always reached through loadu / storeu).
typedef double v4 __attribute__((vector_size(32)));
template <typename V>
__attribute__((always_inline)) inline V loadu(const double *p) {
  V v;
  __builtin_memcpy(&v, p, sizeof(V));
  return v;
}
template <typename V>
__attribute__((always_inline)) inline void storeu(double *p, V v) {
  __builtin_memcpy(p, &v, sizeof(V));
}

inline int nblk(int N) { return (N + RB - 1) / RB; }

void pack(const double *W, int N, int M, std::vector<double> &P) {
  P.assign((size_t)nblk(N) * M * RB, 0.0);
  for (int r = 0; r < N; ++r) {
    double *dst = P.data() + (size_t)(r / RB) * M * RB + (r % RB);
    const double *src = W + (size_t)r * M;
    for (int c = 0; c < M; ++c) dst[(size_t)c * RB] = src[c];
  }
}
void unpack(const std::vector<double> &P, int N, int M, double *W) {
  for (int r = 0; r < N; ++r) {
    const double *src = P.data() + (size_t)(r / RB) * M * RB + (r % RB);
    double *dst = W + (size_t)r * M;
    for (int c = 0; c < M; ++c) dst[c] = src[(size_t)c * RB];
  }
}

// NB blocks (8*NB rows) of one layer: optional update (UPD: 0 none, 1 BP, 2 BPM), then the dot
// products with x; sums[8*NB].  sc[8*NB] = lr * delta of the rows (0 for padding rows).
// V is a vector of VW doubles (4: AVX2 / generic, 8: AVX-512, one vector = one block of 8 rows).
template <typename V, int VW, int NB, int UPD>
__attribute__((always_inline)) inline void rows_pass(double *P, double *Mo, int M, const double *sc, const double *h,
                                                     const double *x, double alpha, double *sums) {
  constexpr int VPB = RB / VW;  // vectors per block
  constexpr int NV = VPB * NB;
  V acc[NV], s[NV];
  for (int j = 0; j < NV; ++j) {
    V z = {}, t = {};
    for (int q = 0; q < VW; ++q) t[q] = sc[VW * j + q];
    acc[j] = z;
    s[j] = t;
  }
  const size_t bstride = (size_t)M * RB;
  for (int c = 0; c < M; ++c) {
    const double hc = UPD ? h[c] : 0.0, xc = x[c];
    for (int j = 0; j < NV; ++j) {
      double *p = P + (size_t)(j / VPB) * bstride + (size_t)c * RB + (j % VPB) * VW;
      V w = loadu<V>(p);
      if (UPD == 1) {
        const V t = s[j] * hc;  // (scalar operands are broadcast)
        w = w + t;
        storeu<V>(p, w);
      } else if (UPD == 2) {
        double *pm = Mo + (p - P);
        V m = loadu<V>(pm);
        const V t = s[j] * hc;
        m = m + t;
        w = w + m;
        storeu<V>(p, w);
        m = m * alpha;
        storeu<V>(pm, m);
      }
      const V t2 = w * xc;
      acc[j] = acc[j] + t2;
    }
  }
  for (int j = 0; j < NV; ++j)
    for (int q = 0; q < VW; ++q) sums[VW * j + q] = acc[j][q];
}

// blocks [b0, b1) of one layer, in groups of GMAX / ... / 2 / 1 blocks: enough independent row sums in
// flight to hide the add latency (8 vectors: 4 blocks of two v4, 8 blocks of one v8)
template <typename V, int VW, int GMAX, int UPD>
__attribute__((always_inline)) inline void blocks_pass(double *P, double *Mo, int M, int b0, int b1, const double *sc,
                                                       const double *h, const double *x, double alpha, double *sums) {
  const size_t bstride = (size_t)M * RB;
  int b = b0;
  if (GMAX >= 8)
    for (; b + 8 <= b1; b += 8)
      rows_pass<V, VW, 8, UPD>(P + b * bstride, Mo ? Mo + b * bstride : nullptr, M, sc + b * RB, h, x, alpha, sums + b * RB);
  for (; b + 4 <= b1; b += 4)
    rows_pass<V, VW, 4, UPD>(P + b * bstride, Mo ? Mo + b * bstride : nullptr, M, sc + b * RB, h, x, alpha, sums + b * RB);
  for (; b + 2 <= b1; b += 2)
    rows_pass<V, VW, 2, UPD>(P + b * bstride, Mo ? Mo + b * bstride : nullptr, M, sc + b * RB, h, x, alpha, sums + b * RB);
  for (; b < b1; ++b)
    rows_pass<V, VW, 1, UPD>(P + b * bstride, Mo ? Mo + b * bstride : nullptr, M, sc + b * RB, h, x, alpha, sums + b * RB);
}

template <typename V, int VW, int GMAX>
__attribute__((always_inline)) inline void blocks_any(int upd, double *P, double *Mo, int M, int b0, int b1, const double *sc,
                                                      const double *h, const double *x, double alpha, double *sums) {
  if (upd == 0) blocks_pass<V, VW, GMAX, 0>(P, Mo, M, b0, b1, sc, h, x, alpha, sums);
  else if (upd == 1) blocks_pass<V, VW, GMAX, 1>(P, Mo, M, b0, b1, sc, h, x, alpha, sums);
  else blocks_pass<V, VW, GMAX, 2>(P, Mo, M, b0, b1, sc, h, x, alpha, sums);
}

void blocks_default(int upd, double *P, double *Mo, int M, int b0, int b1, const double *sc, const double *h,
                    const double *x, double alpha, double *sums) {
  blocks_any<v4, 4, 4>(upd, P, Mo, M, b0, b1, sc, h, x, alpha, sums);
}
#if defined(__x86_64__)
__attribute__((target("avx2"))) void blocks_avx2(int upd, double *P, double *Mo, int M, int b0, int b1, const double *sc,
                                                 const double *h, const double *x, double alpha, double *sums) {
  blocks_any<v4, 4, 4>(upd, P, Mo, M, b0, b1, sc, h, x, alpha, sums);
}
// AVX-512: one 64-byte vector per block of 8 rows.  AVX-512F contains fused multiply-add, so this file is
// compiled with -ffp-contract=off: every product and every sum is rounded on its own, as in the reference.
typedef double v8 __attribute__((vector_size(64)));
__attribute__((target("avx512f"))) void blocks_avx512(int upd, double *P, double *Mo, int M, int b0, int b1, const double *sc,
                                                      const double *h, const double *x, double alpha, double *sums) {
  blocks_any<v8, 8, 8>(upd, P, Mo, M, b0, b1, sc, h, x, alpha, sums);
}
#endif
typedef void (*blocks_fn)(int, double *, double *, int, int, int, const double *, const double *, const double *, double,
                          double *);
blocks_fn pick_blocks() {
#if defined(__x86_64__)
  // HPNN_CPU_ISA = avx512 | avx2 | generic.  Default avx2: on the 784-300-10 tutorial net the 64-byte path is no
  // faster (34.4 k vs 35.3 k iterations/s with 4 threads: the pass is bound by L2 traffic, not by the adds).
  const char *isa = getenv("HPNN_CPU_ISA");
  const bool want512 = isa && !strcmp(isa, "avx512");
  const bool want2 = !isa || want512 || !strcmp(isa, "avx2");
  if (want512 && __builtin_cpu_supports("avx512f")) return blocks_avx512;
  if (want2 && __builtin_cpu_supports("avx2")) return blocks_avx2;
#endif
  return blocks_default;
}

// this thread's share of `n` items (contiguous, balanced); the whole range outside a team
inline void my_range(int n, int &lo, int &hi) {
  int t = 0, nt = 1;
#ifdef _OPENMP
  t = omp_get_thread_num();
  nt = omp_get_num_threads();
#endif
  lo = (int)((long)n * t / nt);
  hi = (int)((long)n * (t + 1) / nt);
}

// one layer: (update +) forward, y = f(W.x); called by every thread of the team, ends with a barrier
void layer(hpnn_kernel *k, int type, int l, int upd, double lr, double alpha, const double *h, blocks_fn fn,
           double *scratch /* 2 * 8*nblk doubles, shared */) {
  const int L = k->n_layers();
  const int N = k->dims[l + 1], M = k->dims[l];
  const int nb = nblk(N);
  double *sc = scratch, *sums = scratch + (size_t)nb * RB;
  const double *x = k->act[l].data();
  double *y = k->act[l + 1].data();
  const bool raw = (type == HPNN_TYPE_SNN && l == L - 1);
  int b0, b1;
  my_range(nb, b0, b1);
  if (upd) {
    const double *d = k->delta[l].data();
    for (int r = b0 * RB; r < b1 * RB; ++r) sc[r] = r < N ? lr * d[r] : 0.0;
  }
  fn(upd, k->packW[l].data(), upd == 2 ? k->packM[l].data() : nullptr, M, b0, b1, sc, h, x, alpha, sums);
  for (int r = b0 * RB; r < b1 * RB && r < N; ++r) y[r] = raw ? sums[r] : act(sums[r]);
#pragma omp barrier
}

// deltas of the layers below the output layer from the OLD packed weights; column c is summed
// over the rows in ascending order (src/ann.c:1279-1592)
void deltas_hidden(hpnn_kernel *k) {
  const int L = k->n_layers();
  for (int l = L - 1; l >= 1; --l) {
    const int N = k->dims[l + 1], M = k->dims[l];
    const double *P = k->packW[l].data();
    const double *dn = k->delta[l].data();
    const double *hh = k->act[l].data();
    double *d = k->delta[l - 1].data();
    int c0, c1;
    my_range(M, c0, c1);
    constexpr int CI = 8;  // independent column chains
    for (int cb = c0; cb < c1; cb += CI) {
      const int nc = c1 - cb < CI ? c1 - cb : CI;
      double s[CI];
      for (int i = 0; i < CI; ++i) s[i] = 0.0;
      for (int r = 0; r < N; ++r) {
        const double dr = dn[r];
        const double *p = P + ((size_t)(r / RB) * M + cb) * RB + (r % RB);
        for (int i = 0; i < nc; ++i) s[i] += p[(size_t)i * RB] * dr;
      }
      for (int i = 0; i < nc; ++i) d[cb + i] = s[i] * dact(hh[cb + i]);
    }
#pragma omp barrier
  }
}

}  // namespace packed

// a team only pays off when a phase is longer than a barrier
bool worth_a_team(const hpnn_kernel *k) {
  long w = 0;
  for (int l = 0; l < k->n_layers(); ++l) w += (long)k->dims[l] * k->dims[l + 1];
  return w > 16384;
}

}  // namespace

extern "C" {

void hpnn_set_omp_threads(int n) {
  g_threads = n;
#ifdef _OPENMP
  if (n > 0) omp_set_num_threads(n);
#endif
}

void hpnn_train_opts_default(hpnn_train_opts *o, int type, int train, int cuda_defaults) {
  if (!o) return;
  memset(o, 0, sizeof(*o));
  o->type = type;
  o->train = train;
  // CUDA build uses 0.01 everywhere; CPU ANN uses the header constants
  if (cuda_defaults || type == HPNN_TYPE_SNN) o->lr = 0.01;
  else o->lr = (train == HPNN_TRAIN_BPM) ? 0.0005 : 0.001;
  o->alpha = 0.2;
  o->delta = 1e-6;
  o->min_iter = (train == HPNN_TRAIN_BPM) ? 15 : 31;
  o->max_iter = 102399;
  o->stable_softmax = 0;
}

int hpnn_cpu_forward(const hpnn_kernel *kc, int type, int stable, const double *in, double *out) {
  hpnn_kernel *k = const_cast<hpnn_kernel *>(kc);
  if (!k || !in) return -1;
  ensure_workspace(k, false);
  memcpy(k->act[0].data(), in, sizeof(double) * k->dims[0]);
  const bool team = worth_a_team(k);
#pragma omp parallel if (team)
  forward(k, type, stable);
  const int L = k->n_layers();
  if (out) memcpy(out, k->act[L].data(), sizeof(double) * k->dims[L]);
  return 0;
}

/* ---- single-step entry points (ann_kernel_run / ann_kernel_train* of the reference) ---- */
double *hpnn_cpu_act(hpnn_kernel *k, int i) {
  if (!k) return nullptr;
  ensure_workspace(k, false);
  return (i >= 0 && i <= k->n_layers()) ? k->act[i].data() : nullptr;
}
double *hpnn_cpu_momentum(hpnn_kernel *k, int l) {
  if (!k || (int)k->dW.size() != k->n_layers() || l < 0 || l >= k->n_layers()) return nullptr;
  return k->dW[l].data();
}
void hpnn_cpu_momentum_init(hpnn_kernel *k) {
  if (k) ensure_workspace(k, true);
}
void hpnn_cpu_momentum_zero(hpnn_kernel *k) {
  if (!k) return;
  for (auto &m : k->dW) std::fill(m.begin(), m.end(), 0.0);
}
void hpnn_cpu_momentum_free(hpnn_kernel *k) {
  if (k) k->dW.clear();
}
// forward from act[0] (already filled by the caller)
int hpnn_cpu_forward_inplace(hpnn_kernel *k, int type, int stable) {
  if (!k) return -1;
  ensure_workspace(k, false);
  const bool team = worth_a_team(k);
#pragma omp parallel if (team)
  forward(k, type, stable);
  return 0;
}
double hpnn_cpu_error(hpnn_kernel *k, int type, const double *target) {
  if (!k || !target) return 0.0;
  ensure_workspace(k, false);
  return error(k, type, target);
}
// one BP / BPM iteration on an already forwarded net; returns E_before - E_after.
// Momentum persists across calls (the caller zeroes it per sample).
double hpnn_cpu_train_iteration(hpnn_kernel *k, const hpnn_train_opts *opt, const double *target) {
  if (!k || !opt || !target) return 0.0;
  hpnn_train_opts d;
  hpnn_train_opts_default(&d, opt->type, opt->train, 0);
  const bool momentum = opt->train == HPNN_TRAIN_BPM;
  ensure_workspace(k, momentum);
  const double lr = opt->lr > 0 ? opt->lr : d.lr;
  const double Ep = error(k, opt->type, target);
  const bool team = worth_a_team(k);
#pragma omp parallel if (team)
  {
    deltas(k, opt->type, target);
    update_forward(k, opt->type, opt->stable_softmax, momentum, lr, opt->alpha);
  }
  return Ep - error(k, opt->type, target);
}

int hpnn_cpu_train_sample(hpnn_kernel *k, const hpnn_train_opts *opt, const double *in, const double *t,
                          hpnn_train_result *res, double *out) {
  if (!k || !opt || !in || !t) return -1;
  hpnn_train_opts d;
  hpnn_train_opts_default(&d, opt->type, opt->train, 0);
  const bool momentum = opt->train == HPNN_TRAIN_BPM;
  const double lr = opt->lr > 0 ? opt->lr : d.lr;
  const double alpha = opt->alpha;
  const double thr = opt->delta > 0 ? opt->delta : d.delta;
  const int min_iter = opt->min_iter >= 0 ? opt->min_iter : d.min_iter;
  const int max_iter = opt->max_iter >= 0 ? opt->max_iter : d.max_iter;
  ensure_workspace(k, momentum);
  const int L = k->n_layers(), N = k->dims[L];
  memcpy(k->act[0].data(), in, sizeof(double) * k->dims[0]);
  hpnn_train_result r;
  memset(&r, 0, sizeof(r));
  const bool team = worth_a_team(k);
  // ONE parallel region for the whole convergence loop (src/ann.c:2281-2372 forks a team per
  // BLAS-less loop nest, ~6 times per iteration): the phases are orphaned `omp for`s separated
  // by their barriers, and every thread evaluates the tiny output vector and the exit test itself,
  // from the same shared data and therefore with the same outcome.
  const int upd = momentum ? 2 : 1;
  k->packW.resize(L);
  if (momentum) k->packM.resize(L);
  int max_nb = 1;
  for (int l = 0; l < L; ++l) {
    packed::pack(k->W[l].data(), k->dims[l + 1], k->dims[l], k->packW[l]);
    if (momentum) k->packM[l].assign(k->packW[l].size(), 0.0);  // zeroed per sample
    max_nb = std::max(max_nb, packed::nblk(k->dims[l + 1]));
  }
  std::vector<double> scratch((size_t)2 * packed::RB * max_nb, 0.0);
  const packed::blocks_fn fn = packed::pick_blocks();
#pragma omp parallel if (team)
  {
    for (int l = 0; l < L; ++l) packed::layer(k, opt->type, l, 0, 0.0, 0.0, nullptr, fn, scratch.data());
    if (opt->type == HPNN_TYPE_SNN) softmax_output(k, opt->stable_softmax);
    double E = error(k, opt->type, t);
    const double init_err = E;
    int iter = 0, ok = 0, first_ok = 0;
    double dE = 0.0;
    do {
      ++iter;
#pragma omp single
      {
        const double *o = k->act[L].data();
        double *d = k->delta[L - 1].data();
        for (int j = 0; j < N; ++j)
          d[j] = (opt->type == HPNN_TYPE_SNN) ? (t[j] - o[j]) : (t[j] - o[j]) * dact(o[j]);
        for (int l = 1; l < L; ++l) k->hold[l] = k->act[l];  // inputs of the update (equal sizes: no allocation)
      }  // implicit barrier
      packed::deltas_hidden(k);
      for (int l = 0; l < L; ++l)
        packed::layer(k, opt->type, l, upd, lr, alpha, l == 0 ? k->act[0].data() : k->hold[l].data(), fn, scratch.data());
      if (opt->type == HPNN_TYPE_SNN) softmax_output(k, opt->stable_softmax);
      const double En = error(k, opt->type, t);
      dE = E - En;
      E = En;
      const double *o = k->act[L].data();
      double probe = -1.0;
      int max_p = 0, p_trg = 0;
      for (int j = 0; j < N; ++j) {
        if (probe < o[j]) {
          probe = o[j];
          max_p = j;
        }
        if (t[j] == 1.0) p_trg = j;
      }
      ok = (max_p == p_trg);
      if (iter == 1) first_ok = ok;
      if (iter > max_iter) break;
      ok = ok && (iter > min_iter);
    } while (dE > thr || !ok);
#pragma omp master
    {
      r.init_err = init_err;
      r.first_ok = first_ok;
      r.n_iter = iter;
      r.final_de = dE;
      r.last_err = E;
      r.success = ok;
    }
  }
  for (int l = 0; l < L; ++l) {
    packed::unpack(k->packW[l], k->dims[l + 1], k->dims[l], k->W[l].data());
    if (momentum) packed::unpack(k->packM[l], k->dims[l + 1], k->dims[l], k->dW[l].data());
  }
  if (res) *res = r;
  if (out) memcpy(out, k->act[L].data(), sizeof(double) * N);
  return 0;
}

}  // extern "C"
