// CPU / OpenMP engine: the fallback back-end and the oracle every GPU kernel is
// tested against.  It restates the numerical spec of the reference:
//   activation   src/ann.c:883-888        forward     src/ann.c:892-1242, src/snn.c:79-443
//   errors       src/ann.c:1246-1275, src/snn.c:447-477
//   deltas       src/ann.c:1279-1592, src/snn.c:481-794
//   BP / BPM     src/ann.c:1596-1872, 1943-2277
//   sample loops src/ann.c:2281-2467, src/snn.c:1417-1595
// written as one generic L-layer loop instead of the reference's per-BLAS-flavour copies.
#include <math.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
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
  }
  if (momentum && (int)k->dW.size() != L) {
    k->dW.resize(L);
    for (int l = 0; l < L; ++l) k->dW[l].assign(k->W[l].size(), 0.0);
  }
}

void forward(hpnn_kernel *k, int type, int stable) {
  const int L = k->n_layers();
  for (int l = 0; l < L; ++l) {
    const int N = k->dims[l + 1], M = k->dims[l];
    const double *W = k->W[l].data();
    const double *x = k->act[l].data();
    double *y = k->act[l + 1].data();
    const bool raw = (type == HPNN_TYPE_SNN && l == L - 1);
#pragma omp parallel for schedule(static) if (N * (long)M > 16384)
    for (int j = 0; j < N; ++j) {
      const double *w = W + (size_t)j * M;
      double s = 0.0;
      for (int c = 0; c < M; ++c) s += w[c] * x[c];
      y[j] = raw ? s : act(s);
    }
  }
  if (type == HPNN_TYPE_SNN) {
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
  }
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

// all deltas from the OLD weights, top-down
void deltas(hpnn_kernel *k, int type, const double *t) {
  const int L = k->n_layers();
  {
    const int N = k->dims[L];
    const double *o = k->act[L].data();
    double *d = k->delta[L - 1].data();
    for (int j = 0; j < N; ++j) d[j] = (type == HPNN_TYPE_SNN) ? (t[j] - o[j]) : (t[j] - o[j]) * dact(o[j]);
  }
  for (int l = L - 1; l >= 1; --l) {
    const int N = k->dims[l + 1], M = k->dims[l];
    const double *W = k->W[l].data();
    const double *dn = k->delta[l].data();
    const double *h = k->act[l].data();
    double *d = k->delta[l - 1].data();
#pragma omp parallel for schedule(static) if (N * (long)M > 16384)
    for (int c = 0; c < M; ++c) {
      double s = 0.0;
      for (int r = 0; r < N; ++r) s += W[(size_t)r * M + c] * dn[r];
      d[c] = s * dact(h[c]);
    }
  }
}

void update(hpnn_kernel *k, bool momentum, double lr, double alpha) {
  const int L = k->n_layers();
  for (int l = L - 1; l >= 0; --l) {
    const int N = k->dims[l + 1], M = k->dims[l];
    double *W = k->W[l].data();
    double *dW = momentum ? k->dW[l].data() : nullptr;
    const double *d = k->delta[l].data();
    const double *h = k->act[l].data();
#pragma omp parallel for schedule(static) if (N * (long)M > 16384)
    for (int r = 0; r < N; ++r) {
      const double s = lr * d[r];
      double *w = W + (size_t)r * M;
      if (!momentum) {
        for (int c = 0; c < M; ++c) w[c] += s * h[c];
      } else {
        double *m = dW + (size_t)r * M;
        for (int c = 0; c < M; ++c) {
          m[c] += s * h[c];
          w[c] += m[c];
          m[c] *= alpha;
        }
      }
    }
  }
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
  deltas(k, opt->type, target);
  update(k, momentum, lr, opt->alpha);
  forward(k, opt->type, opt->stable_softmax);
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
  if (momentum)
    for (auto &m : k->dW) std::fill(m.begin(), m.end(), 0.0);  // zeroed per sample
  memcpy(k->act[0].data(), in, sizeof(double) * k->dims[0]);
  forward(k, opt->type, opt->stable_softmax);
  double E = error(k, opt->type, t);
  hpnn_train_result r;
  memset(&r, 0, sizeof(r));
  r.init_err = E;
  int iter = 0, ok = 0;
  double dE = 0.0;
  do {
    ++iter;
    deltas(k, opt->type, t);
    update(k, momentum, lr, alpha);
    forward(k, opt->type, opt->stable_softmax);
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
    if (iter == 1) r.first_ok = ok;
    if (iter > max_iter) break;
    ok = ok && (iter > min_iter);
  } while (dE > thr || !ok);
  r.n_iter = iter;
  r.final_de = dE;
  r.last_err = E;
  r.success = ok;
  if (res) *res = r;
  if (out) memcpy(out, k->act[L].data(), sizeof(double) * N);
  return 0;
}

}  // extern "C"
