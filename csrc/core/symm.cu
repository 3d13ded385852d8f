// Symmetric device memory for one NVLink / NVSwitch domain, built on the CUDA virtual memory
// management API: the same-sized allocation on every rank, every rank's copy mapped into every
// participant (plain peer loads / stores), and -- where the fabric supports it -- ONE extra
// "multicast" mapping of the whole set: a store to the multicast address is replicated by the
// NVSwitch into all ranks' copies (multimem.st), a reduction is applied to all of them
// (multimem.red) and a load returns the element-wise sum over the ranks (multimem.ld_reduce), all
// without the issuing SM touching more than one address.  This is the NVLS path that SURVEY 5
// names as the B200 counterpart of the reference's "collectives":
//   * gather-to-GPU-0 + broadcast by cudaMemcpyAsync, src/cuda_ann.cu:610-666 (EXP model),
//   * everything on GPU 0 behind peer access, src/libhpnn.c:282-288 (P2P model),
//   * MPI_Allgather of activations / deltas / weight rows, src/ann.c:925,1300,1638.
//
// Two ways to build a set:
//   * one process, several devices (the C API's hpnn_net): hpnn_symm_create_local();
//   * one process per device (torchrun): every rank calls hpnn_symm_create_shared(), exports its
//     allocation as a POSIX file descriptor, the Python layer passes the descriptors between the
//     processes (SCM_RIGHTS over a unix socket, hpnn_b200/parallel/symm.py) and every rank
//     imports its peers' with hpnn_symm_import_peer(); the multicast object is created by rank 0
//     and shared the same way.
// Everything degrades gracefully: no multicast support -> mc pointer NULL (callers use per-peer
// stores); no VMM support at all -> creation fails and callers fall back to cudaMalloc + CUDA IPC.
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <unistd.h>

#include <vector>

#include "hpnn_core.h"

namespace hpnn {
void set_error_msg(const char *what);
}

struct hpnn_symm {
  int world = 0, rank = -1;      // rank < 0: single-process set (all ranks are local devices)
  std::vector<int> devs;         // device ordinal per rank (local set) / own device at [rank]
  size_t bytes = 0, padded = 0;  // requested / allocated size of every copy
  std::vector<CUmemGenericAllocationHandle> mem;  // per rank (0 = not imported)
  std::vector<CUdeviceptr> va;                    // per rank unicast mapping
  CUmemGenericAllocationHandle mc = 0;
  CUdeviceptr mc_va = 0;
  bool mc_bound = false;
  int mc_devices_added = 0;
};

namespace {

// The driver API is reached through cudaGetDriverEntryPoint instead of linking libcuda: the library
// must stay loadable on hosts without a GPU driver (CPU engine, file formats, the CPU test-suite).
#define HPNN_DRV_FUNCS(X)                                                                               \
  X(cuInit) X(cuDeviceGet) X(cuDeviceGetAttribute) X(cuGetErrorName) X(cuMemGetAllocationGranularity)    \
  X(cuMulticastGetGranularity) X(cuMemSetAccess) X(cuMemAddressReserve) X(cuMemMap) X(cuMemUnmap)        \
  X(cuMemAddressFree) X(cuMemCreate) X(cuMemRelease) X(cuMemExportToShareableHandle)                     \
  X(cuMemImportFromShareableHandle) X(cuMulticastCreate) X(cuMulticastAddDevice) X(cuMulticastBindMem)
#define HPNN_DRV_DECL(name) decltype(&name) p_##name = nullptr;
HPNN_DRV_FUNCS(HPNN_DRV_DECL)
#undef HPNN_DRV_DECL

bool load_driver() {
  static int state = 0;  // 0 = not tried, 1 = ok, -1 = unavailable
  if (state != 0) return state > 0;
  bool ok = true;
#define HPNN_DRV_LOAD(name)                                                                              \
  {                                                                                                      \
    void *fn = nullptr;                                                                                  \
    cudaDriverEntryPointQueryResult q;                                                                   \
    if (cudaGetDriverEntryPoint(#name, &fn, cudaEnableDefault, &q) != cudaSuccess || !fn) ok = false;     \
    p_##name = reinterpret_cast<decltype(&name)>(fn);                                                     \
  }
  HPNN_DRV_FUNCS(HPNN_DRV_LOAD)
#undef HPNN_DRV_LOAD
  cudaGetLastError();
  if (ok) ok = p_cuInit(0) == CUDA_SUCCESS;
  state = ok ? 1 : -1;
  if (!ok) hpnn::set_error_msg("hpnn_symm: CUDA driver API (virtual memory management) is not available");
  return ok;
}
// from here on every cuXxx call goes through the table
#define cuDeviceGet p_cuDeviceGet
#define cuDeviceGetAttribute p_cuDeviceGetAttribute
#define cuGetErrorName p_cuGetErrorName
#define cuMemGetAllocationGranularity p_cuMemGetAllocationGranularity
#define cuMulticastGetGranularity p_cuMulticastGetGranularity
#define cuMemSetAccess p_cuMemSetAccess
#define cuMemAddressReserve p_cuMemAddressReserve
#define cuMemMap p_cuMemMap
#define cuMemUnmap p_cuMemUnmap
#define cuMemAddressFree p_cuMemAddressFree
#define cuMemCreate p_cuMemCreate
#define cuMemRelease p_cuMemRelease
#define cuMemExportToShareableHandle p_cuMemExportToShareableHandle
#define cuMemImportFromShareableHandle p_cuMemImportFromShareableHandle
#define cuMulticastCreate p_cuMulticastCreate
#define cuMulticastAddDevice p_cuMulticastAddDevice
#define cuMulticastBindMem p_cuMulticastBindMem

bool drv_fail(CUresult r, const char *what) {
  if (r == CUDA_SUCCESS) return false;
  const char *name = nullptr;
  cuGetErrorName(r, &name);
  char msg[256];
  snprintf(msg, sizeof(msg), "hpnn_symm: %s failed: %s", what, name ? name : "?");
  hpnn::set_error_msg(msg);
  return true;
}

CUmemAllocationProp alloc_prop(int dev) {
  CUmemAllocationProp p;
  memset(&p, 0, sizeof(p));
  p.type = CU_MEM_ALLOCATION_TYPE_PINNED;
  p.location.type = CU_MEM_LOCATION_TYPE_DEVICE;
  p.location.id = dev;
  p.requestedHandleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  return p;
}

bool multicast_supported(int dev) {
  int v = 0;
  CUdevice d;
  if (cuDeviceGet(&d, dev) != CUDA_SUCCESS) return false;
  if (cuDeviceGetAttribute(&v, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, d) != CUDA_SUCCESS) return false;
  return v != 0;
}

// size every copy must have: a multiple of the allocation granularity and of the multicast one
size_t padded_size(size_t bytes, int dev, int world, bool want_mc) {
  CUmemAllocationProp p = alloc_prop(dev);
  size_t g = 0;
  if (cuMemGetAllocationGranularity(&g, &p, CU_MEM_ALLOC_GRANULARITY_RECOMMENDED) != CUDA_SUCCESS || g == 0) g = 2u << 20;
  if (want_mc) {
    CUmulticastObjectProp mp;
    memset(&mp, 0, sizeof(mp));
    mp.numDevices = (unsigned)world;
    mp.size = bytes;
    mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    size_t mg = 0;
    if (cuMulticastGetGranularity(&mg, &mp, CU_MULTICAST_GRANULARITY_RECOMMENDED) == CUDA_SUCCESS && mg > g) g = mg;
  }
  return (bytes + g - 1) / g * g;
}

bool map_for(CUdeviceptr va, size_t size, const std::vector<int> &devs) {
  std::vector<CUmemAccessDesc> acc(devs.size());
  for (size_t i = 0; i < devs.size(); ++i) {
    acc[i].location.type = CU_MEM_LOCATION_TYPE_DEVICE;
    acc[i].location.id = devs[i];
    acc[i].flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE;
  }
  return !drv_fail(cuMemSetAccess(va, size, acc.data(), acc.size()), "cuMemSetAccess");
}

bool reserve_and_map(CUmemGenericAllocationHandle h, size_t size, const std::vector<int> &devs, CUdeviceptr *out) {
  CUdeviceptr va = 0;
  if (drv_fail(cuMemAddressReserve(&va, size, 0, 0, 0), "cuMemAddressReserve")) return false;
  if (drv_fail(cuMemMap(va, size, 0, h, 0), "cuMemMap")) {
    cuMemAddressFree(va, size);
    return false;
  }
  if (!map_for(va, size, devs)) {
    cuMemUnmap(va, size);
    cuMemAddressFree(va, size);
    return false;
  }
  *out = va;
  return true;
}

void ensure_context(int dev) {
  cudaSetDevice(dev);
  cudaFree(nullptr);  // makes the primary context current for the driver API calls below
}

}  // namespace

extern "C" {

int hpnn_symm_multicast_supported(int dev) {
  if (!load_driver()) return 0;
  return multicast_supported(dev) ? 1 : 0;
}

void hpnn_symm_destroy(hpnn_symm *s) {
  if (!s) return;
  if (s->mc_va) {
    cuMemUnmap(s->mc_va, s->padded);
    cuMemAddressFree(s->mc_va, s->padded);
  }
  if (s->mc) cuMemRelease(s->mc);
  for (int r = 0; r < s->world; ++r) {
    if (s->va[r]) {
      cuMemUnmap(s->va[r], s->padded);
      cuMemAddressFree(s->va[r], s->padded);
    }
    if (s->mem[r]) cuMemRelease(s->mem[r]);
  }
  delete s;
}

// ---- one process, several devices --------------------------------------------------------------
hpnn_symm *hpnn_symm_create_local(const int *devs, int n, size_t bytes, int want_multicast) {
  if (!devs || n < 1 || n > HPNN_MAX_WORLD || bytes == 0) return nullptr;
  if (!load_driver()) return nullptr;
  int cur = 0;
  cudaGetDevice(&cur);
  hpnn_symm *s = new hpnn_symm();
  s->world = n;
  s->rank = -1;
  s->devs.assign(devs, devs + n);
  s->bytes = bytes;
  s->mem.assign(n, 0);
  s->va.assign(n, 0);
  bool mc_ok = want_multicast && n > 1;
  for (int r = 0; r < n && mc_ok; ++r) mc_ok = multicast_supported(devs[r]);
  s->padded = padded_size(bytes, devs[0], n, mc_ok);
  bool ok = true;
  for (int r = 0; r < n && ok; ++r) {
    ensure_context(devs[r]);
    CUmemAllocationProp p = alloc_prop(devs[r]);
    ok = !drv_fail(cuMemCreate(&s->mem[r], s->padded, &p, 0), "cuMemCreate") &&
         reserve_and_map(s->mem[r], s->padded, s->devs, &s->va[r]);
    if (ok) ok = cudaMemset((void *)s->va[r], 0, s->padded) == cudaSuccess;
  }
  if (ok && mc_ok) {
    CUmulticastObjectProp mp;
    memset(&mp, 0, sizeof(mp));
    mp.numDevices = (unsigned)n;
    mp.size = s->padded;
    mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
    bool m = cuMulticastCreate(&s->mc, &mp) == CUDA_SUCCESS;
    for (int r = 0; r < n && m; ++r) {
      CUdevice d;
      m = cuDeviceGet(&d, devs[r]) == CUDA_SUCCESS && cuMulticastAddDevice(s->mc, d) == CUDA_SUCCESS;
    }
    for (int r = 0; r < n && m; ++r) {
      ensure_context(devs[r]);
      m = cuMulticastBindMem(s->mc, 0, s->mem[r], 0, s->padded, 0) == CUDA_SUCCESS;
    }
    if (m) m = reserve_and_map(s->mc, s->padded, s->devs, &s->mc_va);
    if (!m) {  // multicast is an optimisation: keep the unicast mappings
      if (s->mc) cuMemRelease(s->mc);
      s->mc = 0;
      s->mc_va = 0;
    } else {
      s->mc_bound = true;
    }
  }
  cudaSetDevice(cur);
  if (!ok) {
    hpnn_symm_destroy(s);
    return nullptr;
  }
  return s;
}

// ---- one process per device --------------------------------------------------------------------
hpnn_symm *hpnn_symm_create_shared(int dev, int rank, int world, size_t bytes, int want_multicast) {
  if (rank < 0 || rank >= world || world < 1 || world > HPNN_MAX_WORLD || bytes == 0) return nullptr;
  if (!load_driver()) return nullptr;
  ensure_context(dev);
  hpnn_symm *s = new hpnn_symm();
  s->world = world;
  s->rank = rank;
  s->devs.assign(1, dev);
  s->bytes = bytes;
  s->mem.assign(world, 0);
  s->va.assign(world, 0);
  const bool mc_ok = want_multicast && world > 1 && multicast_supported(dev);
  s->padded = padded_size(bytes, dev, world, mc_ok);
  CUmemAllocationProp p = alloc_prop(dev);
  bool ok = !drv_fail(cuMemCreate(&s->mem[rank], s->padded, &p, 0), "cuMemCreate") &&
            reserve_and_map(s->mem[rank], s->padded, s->devs, &s->va[rank]);
  if (ok) ok = cudaMemset((void *)s->va[rank], 0, s->padded) == cudaSuccess && cudaDeviceSynchronize() == cudaSuccess;
  if (!ok) {
    hpnn_symm_destroy(s);
    return nullptr;
  }
  return s;
}

// POSIX file descriptor of this rank's allocation (caller closes it after sending it to the peers)
int hpnn_symm_export_fd(hpnn_symm *s) {
  if (!s || s->rank < 0) return -1;
  int fd = -1;
  if (drv_fail(cuMemExportToShareableHandle(&fd, s->mem[s->rank], CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0),
               "cuMemExportToShareableHandle"))
    return -1;
  return fd;
}

// map a peer's allocation from the descriptor received from it (the descriptor stays the caller's)
int hpnn_symm_import_peer(hpnn_symm *s, int peer, int fd) {
  if (!s || s->rank < 0 || peer < 0 || peer >= s->world || peer == s->rank || fd < 0) return -1;
  if (drv_fail(cuMemImportFromShareableHandle(&s->mem[peer], (void *)(uintptr_t)fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR),
               "cuMemImportFromShareableHandle"))
    return -1;
  return reserve_and_map(s->mem[peer], s->padded, s->devs, &s->va[peer]) ? 0 : -1;
}

// Multicast, step 1 (rank 0): create the object, return its descriptor for the peers (-1: unsupported)
int hpnn_symm_mc_create(hpnn_symm *s) {
  if (!s || s->rank != 0 || s->world < 2 || !multicast_supported(s->devs[0])) return -1;
  CUmulticastObjectProp mp;
  memset(&mp, 0, sizeof(mp));
  mp.numDevices = (unsigned)s->world;
  mp.size = s->padded;
  mp.handleTypes = CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR;
  if (cuMulticastCreate(&s->mc, &mp) != CUDA_SUCCESS) {
    s->mc = 0;
    return -1;
  }
  int fd = -1;
  if (cuMemExportToShareableHandle(&fd, s->mc, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR, 0) != CUDA_SUCCESS) {
    cuMemRelease(s->mc);
    s->mc = 0;
    return -1;
  }
  return fd;
}
// step 1 (other ranks): import rank 0's object
int hpnn_symm_mc_import(hpnn_symm *s, int fd) {
  if (!s || s->rank <= 0 || fd < 0) return -1;
  if (cuMemImportFromShareableHandle(&s->mc, (void *)(uintptr_t)fd, CU_MEM_HANDLE_TYPE_POSIX_FILE_DESCRIPTOR) != CUDA_SUCCESS) {
    s->mc = 0;
    return -1;
  }
  return 0;
}
// step 2 (every rank): add the own device.  ALL ranks must have done this before anyone binds.
int hpnn_symm_mc_add_device(hpnn_symm *s) {
  if (!s || !s->mc) return -1;
  CUdevice d;
  if (cuDeviceGet(&d, s->devs[0]) != CUDA_SUCCESS || cuMulticastAddDevice(s->mc, d) != CUDA_SUCCESS) return -1;
  return 0;
}
// step 3 (every rank, after a barrier): bind the own copy and map the multicast address
int hpnn_symm_mc_bind(hpnn_symm *s) {
  if (!s || !s->mc || s->rank < 0) return -1;
  if (cuMulticastBindMem(s->mc, 0, s->mem[s->rank], 0, s->padded, 0) != CUDA_SUCCESS) return -1;
  if (!reserve_and_map(s->mc, s->padded, s->devs, &s->mc_va)) {
    s->mc_va = 0;
    return -1;
  }
  s->mc_bound = true;
  return 0;
}
// give up on multicast (some rank failed a step): peer mappings stay
void hpnn_symm_mc_drop(hpnn_symm *s) {
  if (!s) return;
  if (s->mc_va) {
    cuMemUnmap(s->mc_va, s->padded);
    cuMemAddressFree(s->mc_va, s->padded);
    s->mc_va = 0;
  }
  if (s->mc) cuMemRelease(s->mc);
  s->mc = 0;
  s->mc_bound = false;
}

void *hpnn_symm_ptr(const hpnn_symm *s, int rank) {
  return (s && rank >= 0 && rank < s->world) ? (void *)s->va[rank] : nullptr;
}
void *hpnn_symm_mc_ptr(const hpnn_symm *s) { return (s && s->mc_bound) ? (void *)s->mc_va : nullptr; }
size_t hpnn_symm_size(const hpnn_symm *s) { return s ? s->padded : 0; }

}  // extern "C"
