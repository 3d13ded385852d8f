// hpnn_net: a device-resident network driven from ONE host process, for C/C++ callers
// (the libhpnn.h shim first of all).  With n_gpus > 1 it is the same row-split as the
// PyTorch HpnnNet -- weights partitioned, one symmetric arena per GPU reachable by every
// peer (cudaDeviceEnablePeerAccess; no IPC needed inside one process), one persistent
// kernel per GPU per sample -- which is also how the reference drives its GPUs
// (single process, src/libhpnn.c:201-305), minus its "everything lives on GPU 0" layout.
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <vector>

#include "hpnn_core.h"
#include "kernel_host.hpp"

namespace hpnn {
void set_error(const char *what, cudaError_t e);
void set_error_msg(const char *what);
}  // namespace hpnn

struct hpnn_net {
  int type = 0, world = 1;
  bool momentum = false;
  std::vector<int> dims;
  struct Rank {
    int dev = 0;
    hpnn_dev_net net;
    std::vector<double *> W, dW;
    double *arena = nullptr;
    unsigned *sync = nullptr;
    hpnn_train_result *res = nullptr;
    cudaStream_t stream = nullptr;
  };
  std::vector<Rank> ranks;
  hpnn_symm *symm = nullptr;  // several GPUs: the arenas are one symmetric set (multicast when the fabric has NVLS)
  double *pin = nullptr;  // pinned staging: input | target | output
  hpnn_train_result *pin_res = nullptr;
  size_t bytes_dev = 0;
  // batched inference work space (grown on demand): two ping-pong activation matrices
  double *bx = nullptr, *by = nullptr;
  size_t bcap = 0;  // doubles per matrix
};

#define NET_CHECK(call, what)          \
  do {                                 \
    cudaError_t e__ = (call);          \
    if (e__ != cudaSuccess) {          \
      hpnn::set_error(what, e__);      \
      return -1;                       \
    }                                  \
  } while (0)

extern "C" {

void hpnn_net_destroy(hpnn_net *n) {
  if (!n) return;
  for (auto &r : n->ranks) {
    cudaSetDevice(r.dev);
    if (r.stream) cudaStreamSynchronize(r.stream);
    for (double *p : r.W) cudaFree(p);
    for (double *p : r.dW) cudaFree(p);
    if (!n->symm) cudaFree(r.arena);
    cudaFree(r.sync);
    cudaFree(r.res);
    if (r.stream) cudaStreamDestroy(r.stream);
  }
  if (n->symm) hpnn_symm_destroy(n->symm);
  if (n->pin) cudaFreeHost(n->pin);
  if (n->pin_res) cudaFreeHost(n->pin_res);
  if (n->bx) cudaFree(n->bx);
  if (n->by) cudaFree(n->by);
  if (!n->ranks.empty()) cudaSetDevice(n->ranks[0].dev);
  delete n;
}

int hpnn_net_upload(hpnn_net *n, const hpnn_kernel *k) {
  if (!n || !k || k->dims != n->dims) return -1;
  const int NL = (int)n->dims.size() - 1;
  for (auto &r : n->ranks) {
    cudaSetDevice(r.dev);
    for (int l = 0; l < NL; ++l) {
      if (r.net.rows[l] == 0) continue;
      const int M = n->dims[l];
      const double *src = k->W[l].data() + (size_t)r.net.row0[l] * M;
      NET_CHECK(cudaMemcpy2DAsync(r.W[l], sizeof(double) * r.net.ld[l], src, sizeof(double) * M, sizeof(double) * M,
                                  r.net.rows[l], cudaMemcpyHostToDevice, r.stream),
                "hpnn_net_upload");
    }
    NET_CHECK(cudaStreamSynchronize(r.stream), "hpnn_net_upload(sync)");
  }
  return 0;
}

int hpnn_net_download(hpnn_net *n, hpnn_kernel *k) {
  if (!n || !k || k->dims != n->dims) return -1;
  const int NL = (int)n->dims.size() - 1;
  for (auto &r : n->ranks) {
    cudaSetDevice(r.dev);
    for (int l = 0; l < NL; ++l) {
      if (r.net.rows[l] == 0) continue;
      const int M = n->dims[l];
      double *dst = k->W[l].data() + (size_t)r.net.row0[l] * M;
      NET_CHECK(cudaMemcpy2DAsync(dst, sizeof(double) * M, r.W[l], sizeof(double) * r.net.ld[l], sizeof(double) * M,
                                  r.net.rows[l], cudaMemcpyDeviceToHost, r.stream),
                "hpnn_net_download");
    }
    NET_CHECK(cudaStreamSynchronize(r.stream), "hpnn_net_download(sync)");
  }
  return 0;
}

hpnn_net *hpnn_net_create(const hpnn_kernel *k, int type, int momentum, int n_gpus) {
  return hpnn_net_create_on(k, type, momentum, n_gpus, 0);
}

hpnn_net *hpnn_net_create_on(const hpnn_kernel *k, int type, int momentum, int n_gpus, int first_dev) {
  if (!k) return nullptr;
  int avail = 0;
  if (cudaGetDeviceCount(&avail) != cudaSuccess || avail < 1) {
    cudaGetLastError();
    hpnn::set_error_msg("hpnn_net_create: no CUDA device");
    return nullptr;
  }
  if ((int)k->dims.size() - 1 > HPNN_MAX_LAYERS) {
    // the host kernel, the file formats and the CPU engine take any depth (like the reference);
    // only the fixed-size device descriptor is capped, and callers fall back to the CPU engine
    hpnn::set_error_msg("hpnn_net_create: more than HPNN_MAX_LAYERS weight layers (CPU engine only)");
    return nullptr;
  }
  if (first_dev < 0 || first_dev >= avail) first_dev = 0;
  avail -= first_dev;
  if (n_gpus <= 0 || n_gpus > avail) n_gpus = avail;
  if (n_gpus > HPNN_MAX_WORLD) n_gpus = HPNN_MAX_WORLD;
  if (n_gpus > 1) {
    // Row-splitting pays only when streaming the weights dominates an iteration.  Measured on 8 B200:
    // 784-8192x4-10 (1.6 GB) runs 3.3x faster on 8 GPUs, but SNN 784-1000-1000-10 (14 MB) and the
    // 851-230-230 tutorial net (1.6 MB) take ~38 / ~23 us per iteration at 2..8 ranks against 20 / 12 us
    // on one GPU: their iterations are a chain of ~10 dependent hand-overs, and one that crosses NVLink
    // (~2 us) costs several times one that stays in one GPU's L2 (~0.5 us), whatever the protocol
    // (profiles/r2_probe8_small.jsonl).  Below HPNN_TP_MIN_MB (default 128 MB of weights) the net
    // therefore stays on ONE GPU even when more are visible.
    long long bytes = 0;
    for (size_t l = 0; l + 1 < k->dims.size(); ++l) bytes += 8LL * k->dims[l] * k->dims[l + 1];
    const char *e = getenv("HPNN_TP_MIN_MB");
    const long long min_mb = e ? atoll(e) : 128;
    if (bytes < (min_mb << 20)) n_gpus = 1;
  }
  hpnn_net *n = new hpnn_net();
  n->type = type;
  n->world = n_gpus;
  n->momentum = momentum != 0;
  n->dims = k->dims;
  const int NL = (int)k->dims.size() - 1;
  n->ranks.resize(n_gpus);
  bool ok = true;
  // peer access all-to-all (the reference only enables it towards GPU 0, src/libhpnn.c:282-288)
  for (int a = 0; a < n_gpus && ok; ++a)
    for (int b = 0; b < n_gpus && ok; ++b)
      if (a != b && hpnn_enable_peer_access(first_dev + a, first_dev + b) != 0) {
        hpnn::set_error_msg("hpnn_net_create: GPUs cannot reach each other (peer access)");
        ok = false;
      }
  for (int g = 0; g < n_gpus && ok; ++g) {
    auto &r = n->ranks[g];
    r.dev = first_dev + g;
    cudaSetDevice(g);
    memset(&r.net, 0, sizeof(r.net));
    r.net.n_layers = NL;
    for (int i = 0; i <= NL; ++i) r.net.dims[i] = k->dims[i];
    r.net.rank = g;
    r.net.world = n_gpus;
    const long long doubles = hpnn_dev_arena_layout(&r.net);
    if (doubles <= 0) {
      ok = false;
      break;
    }
    r.W.assign(NL, nullptr);
    r.dW.assign(NL, nullptr);
    for (int l = 0; l < NL && ok; ++l) {
      const size_t bytes = sizeof(double) * (size_t)r.net.rows[l] * r.net.ld[l];
      r.W[l] = (double *)hpnn_dev_alloc(bytes);
      if (n->momentum) r.dW[l] = (double *)hpnn_dev_alloc(bytes);
      ok = r.W[l] && (!n->momentum || r.dW[l]);
      r.net.W[l] = r.W[l];
      r.net.dW[l] = r.dW[l];
      n->bytes_dev += bytes * (n->momentum ? 2 : 1);
    }
    r.sync = (unsigned *)hpnn_dev_alloc(256 * sizeof(unsigned));
    r.res = (hpnn_train_result *)hpnn_dev_alloc(sizeof(hpnn_train_result));
    n->bytes_dev += sizeof(double) * (size_t)doubles;
    ok = ok && r.sync && r.res && cudaStreamCreateWithFlags(&r.stream, cudaStreamNonBlocking) == cudaSuccess;
    r.net.local_sync = r.sync;
  }
  if (ok) {
    // arenas: the same layout on every rank (hpnn_dev_arena_layout does not depend on the rank's
    // row counts).  Several GPUs: one symmetric set, so that the kernels can also reach all copies
    // through the multicast address; plain allocations if the VMM path is unavailable.
    const size_t arena_bytes = sizeof(double) * (size_t)n->ranks[0].net.arena_doubles;
    const char *no_symm = getenv("HPNN_SYMM");
    if (n_gpus > 1 && !(no_symm && !strcmp(no_symm, "ipc"))) {
      int devs[HPNN_MAX_WORLD];
      for (int g = 0; g < n_gpus; ++g) devs[g] = n->ranks[g].dev;
      const char *no_mc = getenv("HPNN_MULTICAST");
      n->symm = hpnn_symm_create_local(devs, n_gpus, arena_bytes, !(no_mc && !strcmp(no_mc, "0")));
    }
    for (int g = 0; g < n_gpus && ok; ++g) {
      auto &r = n->ranks[g];
      cudaSetDevice(r.dev);
      r.arena = n->symm ? (double *)hpnn_symm_ptr(n->symm, g) : (double *)hpnn_dev_alloc(arena_bytes);
      ok = r.arena != nullptr;
    }
  }
  if (ok) {
    for (auto &r : n->ranks) {
      for (int q = 0; q < n_gpus; ++q) r.net.arena[q] = n->ranks[q].arena;
      r.net.arena_mc = n->symm ? (double *)hpnn_symm_mc_ptr(n->symm) : nullptr;
    }
    const size_t pin_doubles = (size_t)k->dims[0] + 2 * (size_t)k->dims[NL];
    ok = cudaMallocHost((void **)&n->pin, sizeof(double) * pin_doubles) == cudaSuccess &&
         cudaMallocHost((void **)&n->pin_res, sizeof(hpnn_train_result)) == cudaSuccess;
  }
  if (ok) ok = hpnn_net_upload(n, k) == 0;
  for (auto &r : n->ranks) {
    cudaSetDevice(r.dev);
    cudaDeviceSynchronize();
  }
  cudaSetDevice(first_dev);
  if (!ok) {
    hpnn_net_destroy(n);
    return nullptr;
  }
  return n;
}

int hpnn_net_world(const hpnn_net *n) { return n ? n->world : 0; }
void hpnn_net_set_type(hpnn_net *n, int type) {
  if (n) n->type = type;
}
long long hpnn_net_device_bytes(const hpnn_net *n) { return n ? (long long)n->bytes_dev : 0; }

// After an abandoned sample (hpnn_train_result.status != 0) or a failed launch: barrier counters,
// flag words, abort words and launch ids are in an undefined state on every rank.  Bring them back to
// the state of a fresh net so that the next call starts clean.
static void net_reset_sync(hpnn_net *n) {
  for (auto &r : n->ranks) {
    cudaSetDevice(r.dev);
    cudaStreamSynchronize(r.stream);
    // the whole arena, not only the flag words: the launch id restarts from zero, so no {value, tag}
    // pair of an earlier launch may survive (its tag could match a future one)
    cudaMemsetAsync(r.sync, 0, 256 * sizeof(unsigned), r.stream);
    cudaMemsetAsync(r.arena, 0, sizeof(double) * (size_t)r.net.arena_doubles, r.stream);
    cudaStreamSynchronize(r.stream);
  }
  cudaSetDevice(n->ranks[0].dev);
}

// Raise the abort word of ranks [0, upto) from the host while their persistent kernels are running
// (a side stream: the kernels occupy the ranks' own streams), then wait for them to leave.
static void net_abort_running(hpnn_net *n, int upto) {
  static const unsigned long long one = 1ull;
  for (int g = 0; g < upto; ++g) {
    auto &r = n->ranks[g];
    cudaSetDevice(r.dev);
    cudaStream_t side = nullptr;
    if (cudaStreamCreateWithFlags(&side, cudaStreamNonBlocking) != cudaSuccess) continue;
    cudaMemcpyAsync(reinterpret_cast<unsigned long long *>(r.arena + r.net.off_flags) + HPNN_MAX_WORLD, &one, sizeof(one),
                    cudaMemcpyHostToDevice, side);
    cudaStreamSynchronize(side);
    cudaStreamDestroy(side);
  }
  for (int g = 0; g < upto; ++g) {
    cudaSetDevice(n->ranks[g].dev);
    cudaStreamSynchronize(n->ranks[g].stream);
  }
}

static int net_launch(hpnn_net *n, const hpnn_train_opts *o, int mode, const double *in, const double *target) {
  const int NL = (int)n->dims.size() - 1;
  const int n_in = n->dims[0], n_out = n->dims[NL];
  memcpy(n->pin, in, sizeof(double) * n_in);
  if (target) memcpy(n->pin + n_in, target, sizeof(double) * n_out);
  else memset(n->pin + n_in, 0, sizeof(double) * n_out);
  for (auto &r : n->ranks) {
    cudaSetDevice(r.dev);
    NET_CHECK(cudaMemcpyAsync(r.arena + r.net.off_act[0], n->pin, sizeof(double) * n_in, cudaMemcpyHostToDevice, r.stream),
              "sample H2D");
    NET_CHECK(cudaMemcpyAsync(r.arena + r.net.off_target, n->pin + n_in, sizeof(double) * n_out, cudaMemcpyHostToDevice,
                              r.stream),
              "target H2D");
  }
  // one persistent kernel per GPU; they meet in the cross-GPU barriers
  for (int g = 0; g < (int)n->ranks.size(); ++g) {
    auto &r = n->ranks[g];
    cudaSetDevice(r.dev);
    if (hpnn_dev_train_sample(&r.net, o, mode, r.res, r.stream) != 0) {
      // the ranks launched so far are waiting for this one: tell them to leave instead of letting
      // them spin until their time-out (the message of the failed launch is kept)
      net_abort_running(n, g);
      net_reset_sync(n);
      return -1;
    }
  }
  auto &r0 = n->ranks[0];
  cudaSetDevice(r0.dev);
  NET_CHECK(cudaMemcpyAsync(n->pin_res, r0.res, sizeof(hpnn_train_result), cudaMemcpyDeviceToHost, r0.stream), "result D2H");
  NET_CHECK(cudaMemcpyAsync(n->pin + n_in + n_out, r0.arena + r0.net.off_out, sizeof(double) * n_out,
                            cudaMemcpyDeviceToHost, r0.stream),
            "output D2H");
  for (auto &r : n->ranks) {
    cudaSetDevice(r.dev);
    NET_CHECK(cudaStreamSynchronize(r.stream), "hpnn_net sync");
  }
  cudaSetDevice(r0.dev);
  if (n->pin_res->status != 0) {
    net_reset_sync(n);
    hpnn::set_error_msg("hpnn_net: the sample was abandoned: a wait for another CTA or GPU exceeded the spin time-out "
                        "(hpnn_dev_set_spin_timeout_ms / HPNN_SPIN_TIMEOUT_MS)");
    return -1;
  }
  return 0;
}

int hpnn_net_train_sample(hpnn_net *n, const hpnn_train_opts *o, const double *in, const double *target,
                          hpnn_train_result *res, double *out) {
  if (!n || !o || !in || !target) return -1;
  if (o->train == HPNN_TRAIN_BPM && !n->momentum) {
    hpnn::set_error_msg("hpnn_net_train_sample: net was created without momentum buffers");
    return -1;
  }
  if (net_launch(n, o, 0, in, target) != 0) return -1;
  if (res) *res = *n->pin_res;
  const int NL = (int)n->dims.size() - 1;
  if (out) memcpy(out, n->pin + n->dims[0] + n->dims[NL], sizeof(double) * n->dims[NL]);
  return 0;
}

// Batched forward of `batch` samples (row-major X [batch x n_in] on the host) through the FP64
// DMMA GEMM path; Y [batch x n_out] on the host.  Single-GPU nets only (a row-split net is
// served sample by sample through hpnn_net_run).  Extension: the reference is batch 1.
int hpnn_net_forward_batch(hpnn_net *n, const double *X, int batch, double *Y, int stable_softmax) {
  if (!n || !X || !Y || batch < 1) return -1;
  if (n->world != 1) {
    hpnn::set_error_msg("hpnn_net_forward_batch: single-GPU nets only");
    return -2;
  }
  const int NL = (int)n->dims.size() - 1;
  auto &r = n->ranks[0];
  cudaSetDevice(r.dev);
  int wmax = 0;
  for (int d : n->dims) wmax = d > wmax ? d : wmax;
  wmax = (wmax + 7) & ~7;
  const size_t need = (size_t)batch * wmax;
  if (need > n->bcap) {
    if (n->bx) cudaFree(n->bx);
    if (n->by) cudaFree(n->by);
    n->bx = (double *)hpnn_dev_alloc(sizeof(double) * need);
    n->by = (double *)hpnn_dev_alloc(sizeof(double) * need);
    n->bcap = (n->bx && n->by) ? need : 0;
    if (!n->bcap) return -1;
  }
  const int ld0 = (n->dims[0] + 7) & ~7;
  NET_CHECK(cudaMemsetAsync(n->bx, 0, sizeof(double) * (size_t)batch * ld0, r.stream), "batch memset");
  NET_CHECK(cudaMemcpy2DAsync(n->bx, sizeof(double) * ld0, X, sizeof(double) * n->dims[0], sizeof(double) * n->dims[0],
                              batch, cudaMemcpyHostToDevice, r.stream),
            "batch H2D");
  double *cur = n->bx, *nxt = n->by;
  int ldc = ld0;
  for (int l = 0; l < NL; ++l) {
    const int rows = n->dims[l + 1], ldn = (rows + 7) & ~7;
    const int kind = (l == NL - 1 && n->type == HPNN_TYPE_SNN) ? 1 : 0;
    if (hpnn_dev_forward_batch_f64(r.W[l], r.net.ld[l], rows, n->dims[l], cur, ldc, batch, nxt, ldn, kind, r.stream) != 0)
      return -1;
    double *t = cur;
    cur = nxt;
    nxt = t;
    ldc = ldn;
  }
  if (n->type == HPNN_TYPE_SNN && hpnn_dev_softmax_rows(cur, ldc, batch, n->dims[NL], stable_softmax, r.stream) != 0) return -1;
  NET_CHECK(cudaMemcpy2DAsync(Y, sizeof(double) * n->dims[NL], cur, sizeof(double) * ldc, sizeof(double) * n->dims[NL], batch,
                              cudaMemcpyDeviceToHost, r.stream),
            "batch D2H");
  NET_CHECK(cudaStreamSynchronize(r.stream), "batch sync");
  return 0;
}

int hpnn_net_run(hpnn_net *n, const double *in, double *out, int stable_softmax) {
  if (!n || !in || !out) return -1;
  hpnn_train_opts o;
  hpnn_train_opts_default(&o, n->type, HPNN_TRAIN_BP, 1);
  o.stable_softmax = stable_softmax;
  if (net_launch(n, &o, 1, in, nullptr) != 0) return -1;
  const int NL = (int)n->dims.size() - 1;
  memcpy(out, n->pin + n->dims[0] + n->dims[NL], sizeof(double) * n->dims[NL]);
  return 0;
}

}  // extern "C"
