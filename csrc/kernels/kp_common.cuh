// Pieces shared by the two persistent trainers (streaming K-P in train_persistent.cu and the
// shared-memory-resident variant in train_resident.cu): kernel parameters, the grid / cross-rank
// barrier, the row-partition helpers.
#pragma once
#include "dev_common.cuh"

namespace hpnn {

constexpr int kMaxRsplit = 4;  // row parts of a column task (see phase_cols)
constexpr int kLazyChunk = 8;  // update factors kept per row / column by the deferred-update trainer
constexpr int kLazyDotSlots = 160;  // per-CTA partial dot product slots per rank (>= grid size)
constexpr int kFoldMaxOut = 32;  // widest output layer the row-panel trainer folds into the layer below it
constexpr size_t kMaxDynSmem = 226 * 1024;  // dynamic shared memory per CTA (227 KB minus the static variables)
constexpr long long kLazyMinBytes = 64LL << 20;  // local weights of layers >= 1 below this: not worth it

struct KP {
  hpnn_dev_net net;
  int type, train, mode, stable;
  double lr, alpha, thr;
  int min_iter, max_iter;
  int max_ld, max_rows;  // shared-memory carve-up (doubles)
  int tagged;            // several ranks: tagged-pair hand-over instead of fence + flag barriers
  int lazy_depth;        // deferred-update trainer: weights are written back every lazy_depth iterations
  long long prefetch_bytes;  // K-PL1: L2 prefetch volume per rank and barrier (0 = off)
  hpnn_train_result *result;
  long long *trace;  // optional phase trace: (tag, clock64) pairs written by CTA 0, thread 0
  int trace_cap;
  int items_in_flight;  // row items (4 x 16 bytes per thread) the read-only top-down pass keeps in flight: 2 or 3
  int trace_ns;
  int jitter_ns;  // > 0: schedule perturbation (tests): random delays of up to this many ns before every arrival / hand-over  // phase stamps in nanoseconds (%globaltimer) instead of SM cycles: HPNN_TRACE_NS=1
  int keep_l2;   // K-PL1: a rank's share of a layer fits L2, so the passes use caching loads instead of evict-first ones
  int slicewise;  // K-PL1, several ranks, short panels: forward passes consume the peers' activations slice by slice
  int fold_out;  // K-PL1: the output layer is folded into the phases of the layer below (see train_lazy1.cu)
  int fault;  // fault injection (tests): 1 = this rank leaves at kernel entry, as if its launch had failed
  unsigned long long spin_budget_ns;  // longest time any wait may take before the job is abandoned (0 = unbounded)
};

// ---- bounded waiting --------------------------------------------------------------------------
// Every wait in the persistent kernels (grid barrier, cross-rank flags, tagged pairs) is a spin on
// memory somebody else writes.  If that somebody never comes -- a rank whose launch failed, a peer
// process that died -- the reference at least exits (CUDA_* macros call exit(-1),
// include/libhpnn/common.h:337-353); an unbounded spin would hang every GPU of the job.  So:
//   * each rank has an ABORT word in its arena (behind the barrier flags, so peers can write it);
//   * a spinning thread looks at the clock every 1024 polls; when one wait has lasted longer than
//     spin_budget_ns it raises the abort word of EVERY rank and leaves the loop;
//   * the same slow path looks at the local abort word, so all other waits end within 1024 polls;
//   * the iteration loop tests the word once per iteration (loaded at the start of the iteration,
//     consumed at its end: off the critical path) and leaves; hpnn_train_result.status = 1.
// After an abort the weights, the barrier words and the tags of that sample are meaningless; the
// host reports the error and must reset the sync state (hpnn_net does, HpnnNet.reset_sync()).
constexpr int kAbortSlot = HPNN_MAX_WORLD;        // u64 index of the abort word in the arena's flag area
constexpr int kCounterSlot = HPNN_MAX_WORLD + 1;  // ... and of the cross-rank barrier counter
struct WatchState {
  unsigned long long *abort_word[HPNN_MAX_WORLD];  // [rank] is the local one
  int rank, world;
  unsigned long long budget_ns;
  unsigned tripped;   // a thread of this CTA gave up or saw the abort word
  unsigned exit_now;  // CTA-uniform copy, written by thread 0 once per iteration
  unsigned jitter_ns; // schedule perturbation, see jitter()
};
static __shared__ WatchState s_watch;
static __shared__ unsigned s_kx;  // cross-rank barrier index of the last arrive (for Barrier::wait_rank)

__device__ __forceinline__ unsigned long long global_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
__device__ __forceinline__ void watch_init(const KP &p) {
  if (threadIdx.x == 0) {
    const hpnn_dev_net &n = p.net;
    for (int q = 0; q < HPNN_MAX_WORLD; ++q)
      s_watch.abort_word[q] =
          q < n.world ? reinterpret_cast<unsigned long long *>(n.arena[q] + n.off_flags) + kAbortSlot : nullptr;
    s_watch.rank = n.rank;
    s_watch.world = n.world;
    s_watch.budget_ns = p.spin_budget_ns;
    s_watch.tripped = 0;
    s_watch.exit_now = 0;
    s_watch.jitter_ns = p.jitter_ns > 0 ? (unsigned)p.jitter_ns : 0u;
  }
  __syncthreads();
}
// Schedule perturbation (hpnn_dev_set_jitter_ns; the stand-in for a race detector on a pool where
// compute-sanitizer is closed): a pseudo-random delay before every barrier arrival, cross-rank signal and
// tagged store.  All summation orders are fixed, so a run with delays must reproduce the undisturbed run bit
// for bit; a hand-over that only works because "the other side is always later" shows up as a difference.
__device__ __forceinline__ void jitter(unsigned salt) {
  const unsigned j = s_watch.jitter_ns;
  if (j != 0u) {
    unsigned h = ((unsigned)blockIdx.x * 2654435761u) ^ (salt * 40503u) ^ ((unsigned)threadIdx.x * 97u) ^ (unsigned)clock64();
    h ^= h >> 13;
    h *= 0x5bd1e995u;
    h ^= h >> 15;
    __nanosleep(h % j);
  }
}
// slow path of a spin: true = stop waiting
static __device__ __noinline__ bool spin_check(unsigned long long &t0) {
  if (s_watch.tripped) return true;
  if (ld_relaxed_sys(s_watch.abort_word[s_watch.rank]) != 0ull) {
    s_watch.tripped = 1;
    return true;
  }
  if (s_watch.budget_ns == 0ull) return false;
  const unsigned long long now = global_ns();
  if (t0 == 0ull) {
    t0 = now;
    return false;
  }
  if (now - t0 <= s_watch.budget_ns) return false;
  for (int q = 0; q < s_watch.world; ++q) st_relaxed_sys(s_watch.abort_word[q], 1ull);
  s_watch.tripped = 1;
  return true;
}
struct Spin {
  unsigned n = 0;
  unsigned long long t0 = 0;
  // call after every poll that did not succeed
  __device__ __forceinline__ bool give_up() {
    if ((++n & 1023u) != 0u) return false;
    return spin_check(t0);
  }
};
// once per iteration: thread 0 passes the abort word it loaded at the START of the iteration
// (`seen`); call before a phase that ends in __syncthreads, read s_watch.exit_now after it
__device__ __forceinline__ void watch_publish(unsigned long long seen) {
  if (threadIdx.x == 0) s_watch.exit_now = (seen != 0ull || s_watch.tripped) ? 1u : 0u;
}
// for hpnn_train_result.status (thread 0 of CTA 0, at the very end)
__device__ __forceinline__ int watch_status() {
  return (s_watch.tripped || s_watch.exit_now || ld_relaxed_sys(s_watch.abort_word[s_watch.rank]) != 0ull) ? 1 : 0;
}
__device__ __forceinline__ unsigned long long watch_peek() {
  return threadIdx.x == 0 ? ld_relaxed_sys(s_watch.abort_word[s_watch.rank]) : 0ull;
}

// Grid barrier, extended to all ranks when world > 1.
//
// Arrival is a fire-and-forget `red.release.gpu` on one of three rotating counters (no
// round trip for a return value); everybody polls that counter with relaxed loads until it
// reads the number of CTAs of the rank and then issues one acquire fence.  CTA 0 zeroes the counter used by the
// PREVIOUS barrier once it has passed the current one: all CTAs have left it by then, and it
// is not needed again before barrier k+2, which cannot complete without CTA 0's own arrival
// at k+1 (ordered after the reset).  The barrier index survives kernel launches in
// local_sync, so the rotation simply continues with the next sample.
//
// With several ranks CTA 0 alone watches the counter, then adds 1 to a barrier counter in every
// rank's arena (fence.sys + one multimem.red, or a red per rank); every CTA polls the counter of
// its own arena and ends with one acquire load.
// local_sync words: [0],[32],[64] counters (own 128 B lines), [128] saved index, [129] cross-rank barrier count, [160] launch id.
struct Barrier {
  unsigned *sync;
  unsigned k;     // barrier index, meaningful in thread 0
  unsigned kx;    // number of cross-rank barriers so far (the arena's barrier counter reads world * kx)
  unsigned grid;  // CTAs of THIS rank (the whole grid, or the rank's share of a virtual-rank launch)
  int bid;        // index of this CTA within the rank
  __device__ __forceinline__ void init(const hpnn_dev_net &n, int grid_, int bid_) {
    sync = n.local_sync;
    grid = (unsigned)grid_;
    bid = bid_;
    k = (threadIdx.x == 0) ? ld_relaxed_gpu(sync + 128) : 0u;
    kx = (threadIdx.x == 0) ? ld_relaxed_gpu(sync + 129) : 0u;
  }
  // grid barrier of this GPU only (no cross-rank part, no system-scope fences)
  __device__ __forceinline__ void sync_local() {
    __syncthreads();
    if (threadIdx.x == 0) {
      unsigned *c = sync + 32 * (k % 3u);
      unsigned *prev = sync + 32 * ((k + 2u) % 3u);
      ++k;
      jitter(k);
      red_add_release_gpu(c, 1u);
      Spin sp;
      while (ld_relaxed_gpu(c) != grid) {
        if (sp.give_up()) break;
      }
      if (bid == 0) *prev = 0;
      fence_gpu();
    }
    __syncthreads();
  }
  // CTA 0, after its fence.sys: this rank's flag word (= kx) and +1 on the barrier counter, on every rank
  // (with_flag: the per-rank flag is only read by wait_rank, i.e. after an arrive_all)
  __device__ __forceinline__ void signal(const hpnn_dev_net &n, bool with_flag = false) {
    if (n.arena_mc) {
      unsigned long long *f = reinterpret_cast<unsigned long long *>(n.arena_mc + n.off_flags);
      if (with_flag) multimem_st_u64(f + n.rank, (unsigned long long)kx);
      multimem_red_add_u64(f + kCounterSlot, 1ull);
    } else {
      for (int q = 0; q < n.world; ++q) {
        unsigned long long *f = reinterpret_cast<unsigned long long *>(n.arena[q] + n.off_flags);
        if (with_flag) st_relaxed_sys(f + n.rank, (unsigned long long)kx);
        red_add_relaxed_sys_u64(f + kCounterSlot, 1ull);
      }
    }
  }
  // Split barrier, arrive half (several ranks only): everything of sync_all except the wait for the other
  // ranks.  The phase that follows consumes the peers' data slice by slice and waits per producer rank
  // (wait_rank) right before it needs that rank's slice, so the NVLink flight and the other ranks' skew
  // hide behind the weight stream.  Every CTA must wait for ALL ranks (its own included) before it
  // arrives at the next barrier: that keeps the three rotating local counters safe.
  __device__ __forceinline__ void arrive_all(const hpnn_dev_net &n) {
    __syncthreads();
    if (threadIdx.x == 0) {
      unsigned *c = sync + 32 * (k % 3u);
      unsigned *prev = sync + 32 * ((k + 2u) % 3u);
      ++k;
      ++kx;
      s_kx = kx;
      jitter(k);
      red_add_release_gpu(c, 1u);
      if (bid == 0) {
        Spin sp;
        while (ld_relaxed_gpu(c) != grid) {
          if (sp.give_up()) break;
        }
        *prev = 0;
        fence_sys();
        jitter(k + 7u);
        signal(n, true);
      }
    }
    __syncthreads();
  }
  // wait half: rank q's data of the last arrive_all is complete and visible (call from ONE lane per warp,
  // __syncwarp afterwards)
  __device__ __forceinline__ void wait_rank(const hpnn_dev_net &n, int q) const {
    const unsigned long long want = (unsigned long long)s_kx;
    const unsigned long long *f = reinterpret_cast<const unsigned long long *>(n.arena[n.rank] + n.off_flags) + q;
    Spin sp;
    while (ld_relaxed_sys(f) < want) {
      if (sp.give_up()) return;
    }
    if (ld_acquire_sys(f) < want) s_watch.tripped = 1;
  }
  __device__ __forceinline__ void finish() {  // call once, after the last barrier
    if (bid == 0 && threadIdx.x == 0) {
      sync[128] = k;
      sync[129] = kx;
    }
  }
  // `remote`: this CTA stored into a peer's arena since the last barrier (only then does
  // it pay for a system-scope fence before arriving).
  __device__ __forceinline__ void sync_all(const hpnn_dev_net &n, bool remote = true) {
    __syncthreads();
    if (threadIdx.x == 0) {
      unsigned *c = sync + 32 * (k % 3u);
      unsigned *prev = sync + 32 * ((k + 2u) % 3u);
      ++k;
      // No system-scope fence here: the arrive is a gpu-scope release (ordered after this CTA's stores,
      // peer stores included, through the __syncthreads above), CTA 0 acquires it by polling the
      // counter and then issues ONE system-scope fence before it publishes the flag -- fences are
      // cumulative, so that fence orders every local CTA's peer stores before the flag.  A fence.sys
      // costs ~2.5 us on this part (measured: barrier with three of them 10.8 us, grid barrier 2 us).
      (void)remote;
      jitter(k);
      red_add_release_gpu(c, 1u);
      Spin sp;
      if (n.world == 1) {
        while (ld_relaxed_gpu(c) != grid) {
          if (sp.give_up()) break;
        }
        if (bid == 0) *prev = 0;
        fence_gpu();
      } else {
        // Several ranks.  CTA 0 alone watches the local counter, issues the ONE system-scope fence of the
        // barrier and then adds 1 to the barrier counter of EVERY rank: a single multimem.red when the
        // arenas have an NVLS multicast mapping (the switch applies it to all copies), one red per rank
        // otherwise.  Every CTA -- not only CTA 0 -- polls the counter of its own arena (one word) until
        // it reads world * barrier-index and finishes with one acquire load: no relay through CTA 0, no
        // second fence.  (fence + red heads a release sequence that the other ranks' reds extend.)
        ++kx;
        const unsigned long long want = (unsigned long long)kx * (unsigned long long)n.world;
        if (bid == 0) {
          while (ld_relaxed_gpu(c) != grid) {
            if (sp.give_up()) break;
          }
          *prev = 0;
          fence_sys();  // release: everything this rank wrote is visible before the counter moves
          jitter(k + 7u);
          signal(n);
        }
        const unsigned long long *ctr = reinterpret_cast<const unsigned long long *>(n.arena[n.rank] + n.off_flags) + kCounterSlot;
        while (ld_relaxed_sys(ctr) < want) {
          if (sp.give_up()) break;
        }
        if (ld_acquire_sys(ctr) < want) s_watch.tripped = 1;  // only after a time-out
      }
    }
    __syncthreads();
  }
};

// ---- tagged hand-over: every value another CTA (or GPU) waits for travels as a 16-byte
// {value, tag} pair written with ONE scalar 128-bit store (st.b128, see dev_common.cuh); the consumer spins on the pair itself until
// the tag is the one it expects.  Data and "ready" flag arrive in the same access, so a phase
// boundary costs one write + one read round trip instead of a barrier plus a staging pass, and
// across GPUs no system-scope fence is needed at all (measured: a fence-based cross-GPU barrier
// costs ~9.6 us on 4 B200s, an NVLink store is visible after ~1.5 us).  The tag is
// (launch id << 32 | iteration): unique per buffer write, never equal to stale contents.
// SYS selects system scope (peer memory) instead of GPU scope.
template <bool SYS>
__device__ __forceinline__ void st_pair(double *base, long long idx, double v, unsigned long long tag) {
  jitter((unsigned)idx);
  if (SYS) st_b128_relaxed_sys(base + 2 * idx, (unsigned long long)__double_as_longlong(v), tag);
  else st_b128_relaxed_gpu(base + 2 * idx, (unsigned long long)__double_as_longlong(v), tag);
}
#ifndef HPNN_POLL_BACKOFF_NS
#define HPNN_POLL_BACKOFF_NS 0  /* measured: any back-off only adds latency (20.1 -> 24.2 us/iteration at 64 ns) */
#endif
constexpr unsigned kPollBackoffNs = HPNN_POLL_BACKOFF_NS;

// U tagged loads in flight per thread: all pending loads are issued back to back, then checked;
// only the ones whose tag is still old are re-issued (one round trip instead of U).
template <int U, bool SYS>
__device__ __forceinline__ void ld_pairs_wait(const double *const (&addr)[U], const bool (&want)[U],
                                              unsigned long long tag, double (&v)[U]) {
  bool done[U];
#pragma unroll
  for (int u = 0; u < U; ++u) {
    done[u] = !want[u];
    v[u] = 0.0;
  }
  bool all;
  Spin sp;
  do {
    unsigned long long a[U];
    unsigned long long t[U];
#pragma unroll
    for (int u = 0; u < U; ++u)
      if (!done[u]) {
        if (SYS) ld_b128_relaxed_sys(addr[u], a[u], t[u]);
        else ld_b128_relaxed_gpu(addr[u], a[u], t[u]);
      }
    all = true;
#pragma unroll
    for (int u = 0; u < U; ++u)
      if (!done[u]) {
        if (t[u] == tag) {
          v[u] = __longlong_as_double((long long)a[u]);
          done[u] = true;
        } else {
          all = false;
        }
      }
    if (kPollBackoffNs && !all) __nanosleep(kPollBackoffNs);
    if (!all && sp.give_up()) break;  // abandoned job: the values are meaningless from here on
  } while (!all);
}
template <bool SYS>
__device__ __forceinline__ double ld_pair_wait(const double *base, long long idx, unsigned long long tag) {
  const double *addr[1] = {base + 2 * idx};
  const bool want[1] = {true};
  double v[1];
  ld_pairs_wait<1, SYS>(addr, want, tag, v);
  return v[0];
}
// gather a whole tagged vector (n values, padded to npad with zeros) into shared memory
template <bool SYS>
__device__ __forceinline__ void stage_pairs(double *dst, const double *pairs, int n, int npad, unsigned long long tag) {
  constexpr int U = 4;
  for (int base = threadIdx.x; base < npad; base += U * blockDim.x) {
    const double *addr[U];
    bool want[U];
    double v[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int i = base + u * blockDim.x;
      want[u] = i < n;
      addr[u] = pairs + 2 * (size_t)(want[u] ? i : 0);
    }
    ld_pairs_wait<U, SYS>(addr, want, tag, v);
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int i = base + u * blockDim.x;
      if (i < npad) dst[i] = v[u];
    }
  }
}

__device__ __forceinline__ int pad8(int v) { return (v + 7) & ~7; }
// inverse of row0(g) = floor(g*n/world)
__device__ __forceinline__ int owner_of_row(int row, int n, int world) {
  return (int)(((long long)(row + 1) * world - 1) / n);
}
__device__ __forceinline__ int row0_of(int g, int n, int world) { return (int)((long long)g * n / world); }

// delta_l[r] = dact(h_l[r]) * sum_g part_l[g][r] for the owned rows, into shared memory
// column-task plan of layer l; computed from the LARGEST row slice so that every rank agrees
__device__ __forceinline__ ColPlan layer_plan(const hpnn_dev_net &n, int l, int grid) {
  return pick_col_plan(n.ld[l], (n.dims[l + 1] + n.world - 1) / n.world, grid, kMaxRsplit);
}

// (rank g, row part rp of the producing layer) lives in slot g*kMaxRsplit + rp
// tagged: partials and activations are {value, tag} pairs; the partials are polled until `tag`
__device__ __forceinline__ void stage_delta(const hpnn_dev_net &n, int l, int rsplit, const double *act_next, double *ds,
                                            bool tagged, unsigned long long tag) {
  const double *part = n.arena[n.rank] + n.off_part[l];
  const int stride = n.part_stride[l];
  const int combos = n.world * rsplit;  // (rank, row part) slots, summed rank-major
  if (!tagged) {
    // four rows per thread and round, every load of a round issued before the first use: the
    // staging pass costs ~rows/2048 L2 round trips instead of one per row and slot
    constexpr int U = 4;
    const int rows = n.rows[l];
    for (int r0 = threadIdx.x; r0 < rows; r0 += U * blockDim.x) {
      double s[U], hn[U];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int r = r0 + u * blockDim.x;
        s[u] = 0.0;
        hn[u] = (r < rows) ? ldg_cg(act_next + n.row0[l] + r) : 0.0;
      }
      for (int g = 0; g < n.world; ++g)
        for (int rp = 0; rp < rsplit; ++rp) {
          const double *src = part + (size_t)(g * kMaxRsplit + rp) * stride;
#pragma unroll
          for (int u = 0; u < U; ++u) {
            const int r = r0 + u * blockDim.x;
            if (r < rows) s[u] += ldg_cg(src + r);
          }
        }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int r = r0 + u * blockDim.x;
        if (r < rows) ds[r] = s[u] * dact_f(hn[u]);
      }
    }
    return;
  }
  for (int r = threadIdx.x; r < n.rows[l]; r += blockDim.x) {
    double s = 0.0;
    if (tagged) {
      for (int c0 = 0; c0 < combos; c0 += 8) {  // 8 polls in flight
        const double *addr[8];
        bool want[8];
        double v[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) {
          const int c = c0 + u;
          want[u] = c < combos;
          const int g = want[u] ? c / rsplit : 0, rp = want[u] ? c % rsplit : 0;
          addr[u] = part + 2 * ((size_t)(g * kMaxRsplit + rp) * stride + r);
        }
        ld_pairs_wait<8, false>(addr, want, tag, v);  // local arena: GPU-scope loads see the peers' stores
#pragma unroll
        for (int u = 0; u < 8; ++u) s += v[u];
      }
    } else {
      for (int g = 0; g < n.world; ++g)
        for (int rp = 0; rp < rsplit; ++rp) s += ldg_cg(part + (size_t)(g * kMaxRsplit + rp) * stride + r);
    }
    const size_t hidx = (size_t)(n.row0[l] + r) * (tagged ? 2 : 1);  // own output of the previous sweep
    ds[r] = s * dact_f(ldg_cg(act_next + hidx));
  }
}


}  // namespace hpnn
