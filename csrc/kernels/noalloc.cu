// hpnn_dev_forward_noalloc: a forward pass that never touches the CUDA allocator.
//
// Who needs this.  libhpnn-style networks are meant to be trained and queried from inside
// another running program.  Two kinds of host cannot let a library allocate device memory:
//   * hosts that hand the whole GPU to their own memory pool at start-up (simulation codes,
//     inference servers): a later cudaMalloc either fails or escapes their accounting;
//   * hosts with several streams in flight in their steady-state loop: cudaMalloc / cudaFree
//     synchronise the device, so a single hidden allocation stalls work that has nothing to
//     do with the network.
// Every other device entry point of this library either allocates (hpnn_net_create, the
// PyTorch layer) or expects the caller to allocate the work space (hpnn_dev_train_sample
// wants an arena and sync words inside hpnn_dev_net).  Here the work space of the persistent
// kernel -- activation buffers, partial sums, barrier words -- is a fixed array that is part
// of the loaded library image instead.
//
// Cost and limits, stated plainly:
//   * the array is HPNN_NOALLOC_SCRATCH_BYTES (1 MiB + 1 KiB) of global memory on EVERY GPU
//     on which this entry point has been called, until the process exits.  CUDA loads modules
//     lazily by default, so the driver reserves the array (and the kernel image) during the
//     FIRST call on a device, not when the library is mapped: that first call does allocate
//     inside the driver.  A host that is about to hand the rest of the GPU to its own pool
//     makes one call per device before doing so (or sets CUDA_MODULE_LOADING=EAGER);
//   * a network fits if its arena layout (hpnn_dev_arena_layout, forward-pass regions only)
//     needs no more than HPNN_NOALLOC_MAX_DOUBLES doubles.  The layout reserves, per layer
//     output of width d, two double buffers of {value, tag} pairs (4 pad8(d)) and the partial-sum
//     slots of the trainers (8 pad8(d)):
//         need = pad8(n_in) + 12 * sum of pad8(d) over hidden and output widths
//                + 2 * pad8(n_out) + 16,
//     so the hidden and output widths may add up to about 10 900 (784-5000-5000-10 fits,
//     784-20000-10 does not).  Larger networks are refused with an error, nothing is launched;
//   * there is one array per GPU and every call uses it.  It still holds whatever the last
//     call left there, so the part this call will use is cleared first (the kernel relies on
//     zero padding behind every vector and on barrier words that start from a known state).
//     Two calls whose GPU work overlaps -- different streams, or different host threads on
//     the same device -- race on it; serialise them on one stream.
//   * the first call on each device looks up the array's device address (see below).  That
//     first call should be made during start-up, before any stream capture is begun.  The
//     address is cached per device for the life of the process; it does not survive a
//     cudaDeviceReset by the host.
#include <cuda_runtime.h>
#include <stdio.h>
#include <string.h>

#include "hpnn_core.h"

namespace hpnn {
void set_error(const char *what, cudaError_t e);
void set_error_msg(const char *what);
}  // namespace hpnn

namespace {

constexpr long long kSyncWords = 256;  // barrier words of the persistent kernel (u32)
constexpr long long kSyncDoubles = kSyncWords * sizeof(unsigned) / sizeof(double);

// The scratch itself.  In HOST code this name is only a stand-in the runtime uses to find the
// symbol: it is not a device address, and a kernel handed `g_noalloc_scratch` from the host
// side would fault or quietly use unrelated memory.  The real address is per device and is
// obtained with cudaGetSymbolAddress below.
static __device__ __align__(256) double g_noalloc_scratch[HPNN_NOALLOC_MAX_DOUBLES + kSyncDoubles];

constexpr int kMaxDevices = 64;
double *g_scratch_addr[kMaxDevices] = {};  // device address of the array, filled once per device

// Set-up, once per device, with that device current.
double *scratch_for_current_device() {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess || dev < 0 || dev >= kMaxDevices) {
    hpnn::set_error("hpnn_dev_forward_noalloc: cudaGetDevice", e);
    return nullptr;
  }
  if (!g_scratch_addr[dev]) {
    double *ptr = nullptr;
    e = cudaGetSymbolAddress((void **)&ptr, g_noalloc_scratch);
    if (e != cudaSuccess || !ptr) {
      hpnn::set_error("hpnn_dev_forward_noalloc: cudaGetSymbolAddress(g_noalloc_scratch)", e);
      return nullptr;
    }
    g_scratch_addr[dev] = ptr;
  }
  return g_scratch_addr[dev];
}

}  // namespace

extern "C" int hpnn_dev_forward_noalloc(int type, int stable_softmax, int n_layers, const int *dims,
                                        const double *const *W, const double *x, double *out, void *stream) {
  if (!dims || !W || !x || !out || n_layers < 2 || n_layers > HPNN_MAX_LAYERS) {
    hpnn::set_error_msg("hpnn_dev_forward_noalloc: bad arguments (need 1..15 hidden layers, non-null pointers)");
    return -1;
  }
  hpnn_dev_net net;
  memset(&net, 0, sizeof(net));
  net.n_layers = n_layers;
  for (int i = 0; i <= n_layers; ++i) {
    if (dims[i] < 1) {
      hpnn::set_error_msg("hpnn_dev_forward_noalloc: every layer width must be >= 1");
      return -1;
    }
    net.dims[i] = dims[i];
  }
  net.rank = 0;
  net.world = 1;
  long long need = hpnn_dev_arena_layout(&net);
  if (need > 0 && net.cta_slots > 0) {
    // the optional per-CTA slots of the shared-memory-resident trainer are the last region of
    // the layout; a forward pass does not need them, so they do not count against the limit
    need = net.off_cta_part[1];
    net.cta_slots = 0;
    net.arena_doubles = need;
  }
  if (need > 0 && net.hist_depth > 0) {  // training-only region behind everything a forward pass uses
    if (net.off_hist_u[1] < need) need = net.off_hist_u[1];
    net.hist_depth = 0;
    net.arena_doubles = need;
  }
  if (need <= 0 || need > (long long)HPNN_NOALLOC_MAX_DOUBLES) {
    char msg[256];
    snprintf(msg, sizeof(msg),
             "hpnn_dev_forward_noalloc: this network needs %lld doubles of work space, the built-in buffer "
             "holds %lld (HPNN_NOALLOC_MAX_DOUBLES); use hpnn_dev_train_sample with your own arena instead",
             need, (long long)HPNN_NOALLOC_MAX_DOUBLES);
    hpnn::set_error_msg(msg);
    return -2;
  }
  for (int l = 0; l < n_layers; ++l) {
    if (!W[l]) {
      hpnn::set_error_msg("hpnn_dev_forward_noalloc: null weight pointer");
      return -1;
    }
    net.W[l] = const_cast<double *>(W[l]);  // forward only: never written
    net.dW[l] = nullptr;
  }
  double *scratch = scratch_for_current_device();
  if (!scratch) return -1;
  net.arena[0] = scratch;
  net.local_sync = reinterpret_cast<unsigned *>(scratch + HPNN_NOALLOC_MAX_DOUBLES);

  cudaStream_t s = (cudaStream_t)stream;
  // the array keeps the previous call's contents: clear what this call depends on being zero
  cudaError_t e = cudaMemsetAsync(scratch, 0, sizeof(double) * (size_t)need, s);
  if (e == cudaSuccess) e = cudaMemsetAsync(net.local_sync, 0, sizeof(unsigned) * kSyncWords, s);
  if (e == cudaSuccess)
    e = cudaMemcpyAsync(scratch + net.off_act[0], x, sizeof(double) * dims[0], cudaMemcpyDeviceToDevice, s);
  if (e != cudaSuccess) {
    hpnn::set_error("hpnn_dev_forward_noalloc: staging", e);
    return -1;
  }
  hpnn_train_opts o;
  hpnn_train_opts_default(&o, type, HPNN_TRAIN_BP, 1);
  o.stable_softmax = stable_softmax;
  if (hpnn_dev_train_sample(&net, &o, /*mode: forward only*/ 1, nullptr, stream) != 0) return -1;
  e = cudaMemcpyAsync(out, scratch + net.off_out, sizeof(double) * dims[n_layers], cudaMemcpyDeviceToDevice, s);
  if (e != cudaSuccess) {
    hpnn::set_error("hpnn_dev_forward_noalloc: output copy", e);
    return -1;
  }
  return 0;
}
