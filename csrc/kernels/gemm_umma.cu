// K-T: batched forward on the 5th-generation tensor cores (opt-in reduced precision).
//
//   Y[b, n] = f( sum_k X[b, k] * W[n, k] )     X: [B x K] bf16, W: [N x K] bf16 (both K-major)
//
// The reference has no batched path and no reduced precision (batch = 1, FP64), so
// this kernel is never used for parity: it serves `run_nn --batch` / HpnnNet.forward_batch with
// precision="bf16".  Three tile mappings (launch_umma picks; HPNN_KT_CLUSTER / hpnn_dev_set_kt_cluster):
//   k_gemm_bf16_umma_pair     (default) CTA pairs on 256x256 tiles, tcgen05.mma.cta_group::2 -- see its comment;
//   k_gemm_bf16_umma<.., 2>   CTA pairs, each CTA its own 128x256 tile, W shared through TMA multicast;
//   k_gemm_bf16_umma<.., 1>   one CTA per 128x256 tile (single-tile batches, operands outside this GPU).
// Common structure (192 threads, persistent over tiles):
//   warp 0      TMA producer: cp.async.bulk.tensor (128B-swizzled [rows x 64] bf16 boxes of X and W)
//               into a shared-memory ring, completion on mbarriers;
//   warp 1      allocates all 512 TMEM columns (two accumulators), one elected lane issues tcgen05.mma
//               (N=256, K=16, kind::f16, fp32 accumulate in TMEM); tcgen05.commit frees ring slots and
//               signals the epilogue;
//   warps 2..5  epilogue: tcgen05.ld (32 lanes x 32 columns per instruction), activation in fp32 (the
//               operands were bf16 anyway), bf16 tile through swizzled shared memory and TMA stores (hidden
//               layers) or direct stores widened to FP64 (last layer).
// Out-of-range rows/columns of partial tiles are zero-filled by TMA and masked at the store.
//
// CL = 2: two CTAs of a thread-block cluster work on the two 128-row batch tiles that share one 256-row
// block of W.  Each CTA fetches HALF of that W block per k-step and the TMA multicasts it into both
// CTAs' rings (cp.async.bulk.tensor ... .multicast::cluster), so a k-step costs 32 KB of L2 reads per SM
// instead of 48 KB -- the ring was waiting on L2, not on the tensor pipe.  A ring slot is released by
// BOTH CTAs' MMA warps (tcgen05.commit ... .multicast::cluster on a 2-count mbarrier), because a slot
// now also receives the peer's half.  The tensor maps of this kernel cover operands in this GPU's own
// memory only (launch_umma checks the pointers and otherwise takes the CL = 1 kernel): TMA multicast loads
// here read only this GPU's memory, since on B200 a TMA multicast load from another GPU's memory hung the GPU.
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#include "dev_common.cuh"

namespace hpnn {
extern long long g_launches;
void set_error(const char *what, cudaError_t e);
void set_error_msg(const char *what);

namespace umma {

constexpr int BM = 128, BN = 256, BK = 64;       // tile (BK bf16 = one 128-byte swizzle row); N = 256 halves the
                                                 // shared-memory bytes per flop (A is read once per 256 columns)
constexpr int UMMA_K = 16;                       // bf16 MMA depth
constexpr int STAGES = 3;                        // 3 x 48 KB ring (one-CTA and multicast kernels; the pair kernel: 5 x 32 KB)
constexpr int A_BYTES = BM * BK * 2, B_BYTES = BN * BK * 2;
constexpr int STAGE_BYTES = A_BYTES + B_BYTES;   // 16 + 32 = 48 KB
constexpr int TMEM_COLS = 2 * BN;                // two fp32 accumulators of BN columns: the whole TMEM
constexpr int THREADS = 192;
constexpr int OUT_BYTES = BM * BN * 2;           // bf16 output tile staged for the TMA store (64 KB)
constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + OUT_BYTES + 1024 /*alignment slack*/ + 256 /*barriers*/;

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "WAIT_LOOP:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra WAIT_DONE;\n\t"
      "bra WAIT_LOOP;\n\t"
      "WAIT_DONE:\n\t"
      "}\n" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}
__device__ __forceinline__ void tma_load_2d(void *smem_dst, const CUtensorMap *map, uint64_t *bar, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(
          smem_u32(smem_dst)),
      "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
      : "memory");
}
// the same load, delivered to the same shared-memory offset (and mbarrier) of every CTA in `mask`
__device__ __forceinline__ void tma_load_2d_mc(void *smem_dst, const CUtensorMap *map, uint64_t *bar, int c0, int c1,
                                               uint16_t mask) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes.multicast::cluster [%0], [%1, {%3, "
      "%4}], [%2], %5;" ::"r"(smem_u32(smem_dst)),
      "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "h"(mask)
      : "memory");
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
// 2/(1+e^-z) - 1 == tanh(z/2): one MUFU operation instead of two (the bf16-output epilogue is MUFU-bound;
// tanh.approx is good to about 2^-11, the bf16 result keeps 2^-8)
__device__ __forceinline__ float ann_act_f32(float z) {
  float y;
  asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(0.5f * z));
  return y;
}
// K-major operand tile, 128-byte swizzle: rows of 128 B, 8-row groups 1024 B apart
__device__ __forceinline__ uint64_t smem_desc_sw128(uint32_t saddr) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr >> 4) & 0x3FFF);       // start address
  d |= (uint64_t)1 << 16;                       // leading byte offset (unused with swizzle)
  d |= (uint64_t)(1024 >> 4) << 32;             // stride byte offset between 8-row groups
  d |= (uint64_t)1 << 46;                       // descriptor version (Blackwell)
  d |= (uint64_t)2 << 61;                       // SWIZZLE_128B
  return d;
}
__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t acc) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t"
      "}\n" ::"r"(tmem_d),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(acc)
      : "memory");
}
__device__ __forceinline__ void umma_commit(uint64_t *bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
               : "memory");
}
// arrives on the mbarrier at this offset in every CTA of `mask`
__device__ __forceinline__ void umma_commit_mc(uint64_t *bar, uint16_t mask) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(
                   smem_u32(bar)),
               "h"(mask)
               : "memory");
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&v)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
        "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]),
        "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]),
        "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
      : "r"(taddr)
      : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

// instruction descriptor: D=f32, A=B=bf16, both K-major, N=256, M=128
constexpr uint32_t kIdesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);

__device__ __forceinline__ void mbar_arrive(uint64_t *bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tma_store_2d(const CUtensorMap *map, const void *smem_src, int c0, int c1) {
  asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];" ::"l"(
                   reinterpret_cast<uint64_t>(map)),
               "r"(smem_u32(smem_src)), "r"(c0), "r"(c1)
               : "memory");
}
__device__ __forceinline__ void epi_bar() { asm volatile("bar.sync 1, 128;" ::: "memory"); }  // the 4 epilogue warps

__device__ __forceinline__ void mbar_arrive_u32(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
// arrival on an mbarrier of another CTA of the cluster (address from mapa)
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t bar_cluster_addr) {
  asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(bar_cluster_addr) : "memory");
}
__device__ __forceinline__ uint32_t mapa_u32(uint32_t cta_addr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(cta_addr), "r"(rank));
  return r;
}
// waits for arrivals that may come from the peer CTA
__device__ __forceinline__ void mbar_wait_cluster(uint64_t *bar, uint32_t parity) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "WAIT_LOOP_C:\n\t"
      "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra WAIT_DONE_C;\n\t"
      "bra WAIT_LOOP_C;\n\t"
      "WAIT_DONE_C:\n\t"
      "}\n" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}

// One output tile's epilogue (4 warps, 128 threads): TMEM -> registers -> activation -> bf16 tile staged in
// shared memory and stored by TMA, one 64-column box at a time (OUT_BF16; `ldy` is then the padded width of the
// output: the pad columns receive f(0) = 0, W's rows beyond `rows` being zero-filled by the loads), or FP64
// values stored directly.  `empty_bar` is the shared-memory
// address of the mbarrier that hands the accumulator back to the MMA warp -- of this CTA (REMOTE = false) or,
// as a shared::cluster address, of the pair's leader CTA (REMOTE = true).
template <bool OUT_BF16, bool REMOTE>
__device__ __forceinline__ void epilogue_tile(uint32_t tmem_acc, uint32_t empty_bar, int q, int lane, int et, int m0, int n0,
                                              int batch, int rows, double *__restrict__ Y, int ldy, int act_kind,
                                              uint8_t *stage_out, const CUtensorMap *map_y) {
  const int r = q * 32 + lane;  // row of the tile held by this thread
  if (OUT_BF16) {
    if (et == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");  // staging free again
    epi_bar();
  }
#pragma unroll 1
  for (int c0 = 0; c0 < BN; c0 += 32) {
    uint32_t v[32];
    tmem_ld32(tmem_acc + ((uint32_t)(q * 32) << 16) + (uint32_t)c0, v);
    if (c0 == BN - 32) {  // everything of this accumulator is in registers: hand it back to the MMA warp
      asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
      __syncwarp();
      if (lane == 0) {
        if (REMOTE) mbar_arrive_cluster(empty_bar);
        else mbar_arrive_u32(empty_bar);
      }
    }
    if (OUT_BF16) {
      // 128-byte-swizzle layout of a [128 x 64] bf16 box: 16-byte chunk j of row r sits at chunk j ^ (r & 7)
      uint8_t *box = stage_out + (c0 >> 6) * (BM * 128) + r * 128;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        uint32_t pk[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          float z0 = __uint_as_float(v[j * 8 + 2 * e]), z1 = __uint_as_float(v[j * 8 + 2 * e + 1]);
          if (act_kind == 0) {
            z0 = ann_act_f32(z0);
            z1 = ann_act_f32(z1);
          }
          __nv_bfloat162 h = __floats2bfloat162_rn(z0, z1);
          pk[e] = *reinterpret_cast<uint32_t *>(&h);
        }
        const int chunk = ((c0 & 63) >> 3) + j;
        *reinterpret_cast<uint4 *>(box + ((chunk ^ (r & 7)) << 4)) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
      }
      if (c0 & 32) {  // a 64-column box is complete: store it now, so the TMA unit never sees one 64 KB burst
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // staging writes visible to the TMA engine
        epi_bar();
        const int bx = c0 >> 6;  // columns beyond the padded width `ldy` are clipped by the tensor map
        if (et == 0 && n0 + 64 * bx < ldy) {
          tma_store_2d(map_y, stage_out + bx * (BM * 128), n0 + 64 * bx, m0);
          asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        }
      }
    } else {
      const int b = m0 + r;
      if (b < batch) {
        double *yrow = Y + (size_t)b * ldy;
#pragma unroll
        for (int j = 0; j < 32; ++j) {
          const int n = n0 + c0 + j;
          if (n < rows) {
            // reduced-precision path: the activation is evaluated in fp32 (operands were bf16)
            const float z = __uint_as_float(v[j]);
            yrow[n] = (double)(act_kind == 0 ? __fdividef(2.0f, 1.0f + __expf(-z)) - 1.0f : z);
          }
        }
      }
    }
  }
}

// optional time stamps of CTA 0 (bench/kt_trace.py): [0] entry, [1] set-up done, [8 + it] producer past the ring
// wait of k-step `it`, [72 + it] MMA thread past the full wait, [136 + 2 i] / [137 + 2 i] epilogue of tile i
__device__ __forceinline__ void kt_stamp(long long *trace, int slot) {
  if (trace != nullptr && blockIdx.x == 0) trace[slot] = clock64();
}

// Persistent kernel: CTA c works on output tiles c, c + gridDim.x, ...  Two 128-column accumulators in
// TMEM: while the epilogue warps drain tile i (tcgen05.ld, activation, store) the MMA warp already
// accumulates tile i+1, and the TMA producer runs ahead through the shared-memory ring across tile
// boundaries.  OUT_BF16: hidden layers hand bf16 activations straight to the next layer's GEMM -- the tile is
// staged in shared memory in the 128-byte-swizzle layout (conflict-free 16-byte stores) and leaves
// through two TMA stores; otherwise FP64 results are stored directly (the last layer, 10..230 wide).
template <bool OUT_BF16, int CL>
__global__ void __launch_bounds__(THREADS, 1)
    k_gemm_bf16_umma(const __grid_constant__ CUtensorMap map_x, const __grid_constant__ CUtensorMap map_w,
                     const __grid_constant__ CUtensorMap map_y, int batch, int rows, int k_total, double *__restrict__ Y,
                     int ldy, int act_kind, long long *trace) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t *smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t *stage_out = smem + STAGES * STAGE_BYTES;  // 2 x [128 rows x 128 B] (OUT_BF16 only)
  uint64_t *bars = reinterpret_cast<uint64_t *>(stage_out + OUT_BYTES);
  uint64_t *full = bars, *empty = bars + STAGES, *tmem_full = bars + 2 * STAGES, *tmem_empty = tmem_full + 2;
  uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(tmem_empty + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) kt_stamp(trace, 0);
  const int nk = (k_total + BK - 1) / BK;
  const int ntn = (rows + BN - 1) / BN, ntm = (batch + BM - 1) / BM;
  // CL == 2: the cluster walks over pairs of batch tiles; CTA `crank` takes tile 2 * pair + crank
  const int crank = CL == 2 ? (int)cluster_ctarank() : 0;
  const int ntiles = CL == 2 ? ntn * ((ntm + 1) / 2) : ntn * ntm;
  const int first = CL == 2 ? (int)blockIdx.x / 2 : (int)blockIdx.x;
  const int step = CL == 2 ? (int)gridDim.x / 2 : (int)gridDim.x;

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&map_x)) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&map_w)) : "memory");
    if (OUT_BF16) asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&map_y)) : "memory");
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], CL);  // released by the MMA warp of every CTA that reads (part of) this slot
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(&tmem_full[a], 1);
      mbar_init(&tmem_empty[a], 4);  // one arrival per epilogue warp
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {  // whole warp: TMEM allocation (two accumulators)
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                 "r"((uint32_t)TMEM_COLS)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (CL == 2) cluster_sync_all();  // the peer's barriers exist before anything is multicast to them
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem_base = *tmem_slot;
  if (threadIdx.x == 0) kt_stamp(trace, 1);

  if (warp == 0) {
    if (lane == 0) {  // TMA producer: one ring for all tiles of this CTA
      int it = 0;
      for (int t = first; t < ntiles; t += step) {
        const int m0 = ((t / ntn) * CL + crank) * BM, n0 = (t % ntn) * BN;
        for (int kb = 0; kb < nk; ++kb, ++it) {
          const int s = it % STAGES;
          const uint32_t ph = (it / STAGES) & 1;
          mbar_wait(&empty[s], ph ^ 1);
          if (it < 64) kt_stamp(trace, 8 + it);
          mbar_expect_tx(&full[s], STAGE_BYTES);
          uint8_t *a = smem + s * STAGE_BYTES;
          tma_load_2d(a, &map_x, &full[s], kb * BK, m0);
          if (CL == 2)  // my half of the W block, into both CTAs
            tma_load_2d_mc(a + A_BYTES + crank * (B_BYTES / 2), &map_w, &full[s], kb * BK, n0 + crank * (BN / 2), 0x3);
          else
            tma_load_2d(a + A_BYTES, &map_w, &full[s], kb * BK, n0);
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {  // MMA issuer
      int it = 0, i = 0;
      for (int t = first; t < ntiles; t += step, ++i) {
        const int acc = i & 1;
        mbar_wait(&tmem_empty[acc], ((i >> 1) & 1) ^ 1);  // the epilogue has drained this accumulator
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        const uint32_t d = tmem_base + (uint32_t)(acc * BN);
        for (int kb = 0; kb < nk; ++kb, ++it) {
          const int s = it % STAGES;
          const uint32_t ph = (it / STAGES) & 1;
          mbar_wait(&full[s], ph);
          if (it < 64) kt_stamp(trace, 72 + it);
          asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
          const uint32_t a_addr = smem_u32(smem + s * STAGE_BYTES);
          const uint64_t adesc = smem_desc_sw128(a_addr), bdesc = smem_desc_sw128(a_addr + A_BYTES);
#pragma unroll
          for (int k = 0; k < BK / UMMA_K; ++k) {
            // advancing K inside the swizzle atom = +32 bytes on the start address field
            umma_bf16(d, adesc + (uint64_t)(2 * k), bdesc + (uint64_t)(2 * k), kIdesc, (kb | k) ? 1u : 0u);
          }
          if (CL == 2) umma_commit_mc(&empty[s], 0x3);     // slot reusable once BOTH CTAs' MMAs have read it
          else umma_commit(&empty[s]);                     // slot reusable once these MMAs have read it
          if (kb == nk - 1) umma_commit(&tmem_full[acc]);  // accumulator complete
        }
      }
    }
  } else {
    // epilogue: warp w may touch TMEM lanes 32*(w%4) .. +31
    const int q = warp & 3;
    const int et = threadIdx.x - 64;  // 0..127
    int i = 0;
    for (int t = first; t < ntiles; t += step, ++i) {
      const int m0 = ((t / ntn) * CL + crank) * BM, n0 = (t % ntn) * BN;
      const int acc = i & 1;
      mbar_wait(&tmem_full[acc], (i >> 1) & 1);
      if (et == 0 && i < 8) kt_stamp(trace, 136 + 2 * i);
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      epilogue_tile<OUT_BF16, false>(tmem_base + (uint32_t)(acc * BN), smem_u32(&tmem_empty[acc]), q, lane, et, m0, n0, batch,
                                     rows, Y, ldy, act_kind, stage_out, &map_y);
      if (et == 0 && i < 8) kt_stamp(trace, 137 + 2 * i);
    }
    if (OUT_BF16 && et == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (CL == 2) cluster_sync_all();  // nobody leaves while the peer may still signal its barriers
  if (warp == 1) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"((uint32_t)TMEM_COLS)
                 : "memory");
  }
}

// ---- the same GEMM on CTA PAIRS: tcgen05.mma.cta_group::2 ---------------------------------------------------
// Two CTAs of a cluster (two SMs of one TPC) compute one 256 x 256 output tile.  Each CTA stages its own 128
// rows of X and HALF of the 256-row W block (32 KB per k-step instead of 48 KB: the one-CTA kernel waits on
// the L2 -> shared-memory traffic, not on the tensor pipe); the leader's MMA instruction reads both CTAs'
// shared memory and writes both CTAs' TMEM (128 lanes x 256 columns each).  Synchronisation:
//   full[s]        in the LEADER: expect_tx of both CTAs' bytes; both CTAs' TMA loads (.cta_group::2) complete on it;
//   empty[s]       in each CTA: tcgen05.commit.cta_group::2 ... multicast::cluster from the leader's MMA thread;
//   tmem_full[a]   in each CTA: the same multicast commit after the tile's last k-step;
//   tmem_empty[a]  in the LEADER: 8 arrivals, one per epilogue warp of both CTAs (remote mbarrier.arrive).
// No TMA multicast is involved: the two halves of W are different data.
namespace pair {
constexpr int P_STAGES = 5;
constexpr int P_A_BYTES = BM * BK * 2, P_B_BYTES = (BN / 2) * BK * 2;
constexpr int P_STAGE_BYTES = P_A_BYTES + P_B_BYTES;  // 32 KB per CTA
constexpr int P_SMEM_BYTES = P_STAGES * P_STAGE_BYTES + OUT_BYTES + 1024 + 256;
static_assert(P_SMEM_BYTES <= 227 * 1024, "shared memory of the pair kernel");
// D=f32, A=B=bf16, K-major, N=256, M=256 (across the pair)
constexpr uint32_t kIdescPair = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)((2 * BM) >> 4) << 24);

__device__ __forceinline__ void tma_load_2d_cg2(void *smem_dst, const CUtensorMap *map, uint32_t bar_cluster_addr, int c0,
                                                int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::
          "r"(smem_u32(smem_dst)),
      "l"(reinterpret_cast<uint64_t>(map)), "r"(bar_cluster_addr), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void umma2_bf16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t acc) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t"
      "}\n" ::"r"(tmem_d),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(acc)
      : "memory");
}
__device__ __forceinline__ void umma2_commit_mc(uint64_t *bar, uint16_t mask) {
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(
                   smem_u32(bar)),
               "h"(mask)
               : "memory");
}
}  // namespace pair

template <bool OUT_BF16>
__global__ void __launch_bounds__(THREADS, 1)
    k_gemm_bf16_umma_pair(const __grid_constant__ CUtensorMap map_x, const __grid_constant__ CUtensorMap map_w,
                          const __grid_constant__ CUtensorMap map_y, int batch, int rows, int k_total,
                          double *__restrict__ Y, int ldy, int act_kind, long long *trace) {
  using namespace pair;
  extern __shared__ uint8_t smem_raw[];
  uint8_t *smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t *stage_out = smem + P_STAGES * P_STAGE_BYTES;
  uint64_t *bars = reinterpret_cast<uint64_t *>(stage_out + OUT_BYTES);
  uint64_t *full = bars, *empty = bars + P_STAGES, *tmem_full = bars + 2 * P_STAGES, *tmem_empty = tmem_full + 2;
  uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(tmem_empty + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) kt_stamp(trace, 0);
  const int nk = (k_total + BK - 1) / BK;
  const int ntn = (rows + BN - 1) / BN, ntm = (batch + BM - 1) / BM;
  const int crank = (int)cluster_ctarank();
  const int ntiles = ntn * ((ntm + 1) / 2);  // 256-row tiles
  const int first = (int)blockIdx.x / 2, step = (int)gridDim.x / 2;

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&map_x)) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&map_w)) : "memory");
    if (OUT_BF16) asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&map_y)) : "memory");
    for (int s = 0; s < P_STAGES; ++s) {
      mbar_init(&full[s], 1);   // used in the leader: its producer's expect_tx arrival
      mbar_init(&empty[s], 1);  // one multicast commit per use
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(&tmem_full[a], 1);
      mbar_init(&tmem_empty[a], 8);  // used in the leader: four epilogue warps of each CTA
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {  // one warp of EACH CTA of the pair: all 512 columns in both
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                 "r"((uint32_t)TMEM_COLS)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  cluster_sync_all();  // both CTAs' barriers and TMEM exist before anything crosses
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem_base = *tmem_slot;
  if (threadIdx.x == 0) kt_stamp(trace, 1);

  if (warp == 0) {
    if (lane == 0) {  // TMA producer of THIS CTA's share; completion is counted by the leader's full[s]
      int it = 0;
      for (int t = first; t < ntiles; t += step) {
        const int m0 = ((t / ntn) * 2 + crank) * BM, n0 = (t % ntn) * BN + crank * (BN / 2);
        for (int kb = 0; kb < nk; ++kb, ++it) {
          const int s = it % P_STAGES;
          const uint32_t ph = (it / P_STAGES) & 1;
          mbar_wait(&empty[s], ph ^ 1);
          if (it < 64) kt_stamp(trace, 8 + it);
          if (crank == 0) mbar_expect_tx(&full[s], 2 * P_STAGE_BYTES);
          const uint32_t bar = mapa_u32(smem_u32(&full[s]), 0);
          uint8_t *a = smem + s * P_STAGE_BYTES;
          tma_load_2d_cg2(a, &map_x, bar, kb * BK, m0);
          tma_load_2d_cg2(a + P_A_BYTES, &map_w, bar, kb * BK, n0);
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0 && crank == 0) {  // the pair's MMA issuer
      int it = 0, i = 0;
      for (int t = first; t < ntiles; t += step, ++i) {
        const int acc = i & 1;
        mbar_wait_cluster(&tmem_empty[acc], ((i >> 1) & 1) ^ 1);  // both CTAs' epilogues have drained it
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        const uint32_t d = tmem_base + (uint32_t)(acc * BN);
        for (int kb = 0; kb < nk; ++kb, ++it) {
          const int s = it % P_STAGES;
          const uint32_t ph = (it / P_STAGES) & 1;
          mbar_wait_cluster(&full[s], ph);
          if (it < 64) kt_stamp(trace, 72 + it);
          asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
          const uint32_t a_addr = smem_u32(smem + s * P_STAGE_BYTES);
          const uint64_t adesc = smem_desc_sw128(a_addr), bdesc = smem_desc_sw128(a_addr + P_A_BYTES);
#pragma unroll
          for (int k = 0; k < BK / UMMA_K; ++k)
            umma2_bf16(d, adesc + (uint64_t)(2 * k), bdesc + (uint64_t)(2 * k), kIdescPair, (kb | k) ? 1u : 0u);
          umma2_commit_mc(&empty[s], 0x3);                          // the slot is free in both CTAs
          if (kb == nk - 1) umma2_commit_mc(&tmem_full[acc], 0x3);  // both halves of the accumulator are complete
        }
      }
    }
  } else {
    const int q = warp & 3;
    const int et = threadIdx.x - 64;
    int i = 0;
    for (int t = first; t < ntiles; t += step, ++i) {
      const int m0 = ((t / ntn) * 2 + crank) * BM, n0 = (t % ntn) * BN;
      const int acc = i & 1;
      mbar_wait(&tmem_full[acc], (i >> 1) & 1);
      if (et == 0 && i < 8) kt_stamp(trace, 136 + 2 * i);
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      epilogue_tile<OUT_BF16, true>(tmem_base + (uint32_t)(acc * BN), mapa_u32(smem_u32(&tmem_empty[acc]), 0), q, lane, et,
                                    m0, n0, batch, rows, Y, ldy, act_kind, stage_out, &map_y);
      if (et == 0 && i < 8) kt_stamp(trace, 137 + 2 * i);
    }
    if (OUT_BF16 && et == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  cluster_sync_all();  // nobody frees TMEM or leaves while the pair still works
  if (warp == 1) {
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"((uint32_t)TMEM_COLS)
                 : "memory");
  }
}

// fp64 -> bf16 (round to nearest even), row-major with independent leading dimensions
__global__ void k_cast_f64_bf16(const double *__restrict__ src, int ld_src, int rows, int cols, __nv_bfloat16 *dst,
                                int ld_dst) {
  const long long total = (long long)rows * ld_dst;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int r = (int)(i / ld_dst), c = (int)(i % ld_dst);
    dst[i] = __float2bfloat16_rn(c < cols ? (float)src[(size_t)r * ld_src + c] : 0.0f);
  }
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                  const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void *p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  return fn;
}

static bool make_map(CUtensorMap *m, const void *base, int rows, int cols, int ld, int box_rows) {
  EncodeTiledFn fn = encode_fn();
  if (!fn) return false;
  cuuint64_t gdim[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t gstride[1] = {(cuuint64_t)ld * 2};
  cuuint32_t box[2] = {(cuuint32_t)BK, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  return fn(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void *>(base), gdim, gstride, box, estr,
            CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

}  // namespace umma
}  // namespace hpnn

using namespace hpnn;

extern "C" {

int hpnn_dev_cast_bf16(const double *src, int ld_src, int rows, int cols, void *dst_bf16, int ld_dst, void *stream) {
  if (ld_dst < cols) {
    set_error_msg("hpnn_dev_cast_bf16: ld_dst < cols");
    return -1;
  }
  long long blocks = ((long long)rows * ld_dst + 255) / 256;
  if (blocks > 148 * 16) blocks = 148 * 16;
  if (blocks < 1) blocks = 1;
  umma::k_cast_f64_bf16<<<(int)blocks, 256, 0, (cudaStream_t)stream>>>(src, ld_src, rows, cols,
                                                                       (__nv_bfloat16 *)dst_bf16, ld_dst);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_error("k_cast_f64_bf16", e);
    return -1;
  }
  ++g_launches;
  return 0;
}

// Tensor maps are cached: encoding one costs a driver call, and an inference loop uses the same few
// (pointer, shape) combinations over and over.
struct MapKey {
  const void *base;
  int rows, cols, ld, box;
  bool operator==(const MapKey &o) const {
    return base == o.base && rows == o.rows && cols == o.cols && ld == o.ld && box == o.box;
  }
};
struct MapEntry {
  CUtensorMap map;
  bool local;  // the operand is an allocation of the current device (not a peer mapping, not host memory)
};
// cudaMalloc / torch allocations of the current device only; anything else (peer mappings through IPC or
// the VMM API, managed or host memory) answers false and keeps the multicast kernel away from it.
static bool memory_of_this_device(const void *p) {
  cudaPointerAttributes a;
  int dev = -1;
  if (cudaGetDevice(&dev) != cudaSuccess || cudaPointerGetAttributes(&a, p) != cudaSuccess) {
    cudaGetLastError();
    return false;
  }
  return a.type == cudaMemoryTypeDevice && a.device == dev;
}
static const MapEntry *cached_map(const void *base, int rows, int cols, int ld, int box_rows) {
  constexpr int kSlots = 64;
  static MapKey keys[kSlots];
  static MapEntry maps[kSlots];
  static int used = 0, next = 0;
  const MapKey k{base, rows, cols, ld, box_rows};
  for (int i = 0; i < used; ++i)
    if (keys[i] == k) return &maps[i];
  const int slot = used < kSlots ? used++ : (next++ % kSlots);
  keys[slot] = MapKey{nullptr, 0, 0, 0, 0};
  if (!umma::make_map(&maps[slot].map, base, rows, cols, ld, box_rows)) return nullptr;
  maps[slot].local = memory_of_this_device(base);
  keys[slot] = k;
  return &maps[slot];
}

// tile mapping of K-T: 0 one CTA per 128 x 256 tile; 1 pairs of CTAs that share each W block through TMA
// multicast (cta_group::1); 2 pairs of CTAs on one 256 x 256 tile (tcgen05.mma.cta_group::2, the default);
// -1: environment HPNN_KT_CLUSTER
static int g_kt_cluster = -1;
static long long *g_kt_trace = nullptr;  // device buffer of 160 stamps, or null (hpnn_dev_set_kt_trace)

// largest number of co-resident 2-CTA clusters of kernel `kfn` (0: cluster launch unavailable)
static int max_clusters(const void *kfn, int smem_bytes) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(2, 1, 1);
  cfg.blockDim = dim3(umma::THREADS, 1, 1);
  cfg.dynamicSmemBytes = smem_bytes;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = 2;
  at[0].val.clusterDim.y = 1;
  at[0].val.clusterDim.z = 1;
  cfg.attrs = at;
  cfg.numAttrs = 1;
  int n = 0;
  if (cudaOccupancyMaxActiveClusters(&n, kfn, &cfg) != cudaSuccess) {
    cudaGetLastError();
    return 0;
  }
  return n;
}

static int launch_umma(const void *Wb, int ldw, int rows, int cols, const void *Xb, int ldx, int batch, double *Y, int ldy,
                       void *Yb, int ldyb, int act_kind, void *stream) {
  if ((ldw % 8) || (ldx % 8) || ((uintptr_t)Wb & 15) || ((uintptr_t)Xb & 15) || (Yb && ((ldyb % 8) || ((uintptr_t)Yb & 15)))) {
    set_error_msg("hpnn_dev_gemm_bf16: operands must be 16-byte aligned with leading dimensions % 8 == 0");
    return -1;
  }
  if (g_kt_cluster < 0) {
    const char *e = getenv("HPNN_KT_CLUSTER");
    g_kt_cluster = (e && e[0] >= '0' && e[0] <= '2') ? e[0] - '0' : 2;
  }
  static int sms = 0;
  if (!sms) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  }
  const int vi = Yb ? 0 : 1;
  const void *k1 = Yb ? (const void *)umma::k_gemm_bf16_umma<true, 1> : (const void *)umma::k_gemm_bf16_umma<false, 1>;
  const void *k2 = Yb ? (const void *)umma::k_gemm_bf16_umma<true, 2> : (const void *)umma::k_gemm_bf16_umma<false, 2>;
  const void *k3 = Yb ? (const void *)umma::k_gemm_bf16_umma_pair<true> : (const void *)umma::k_gemm_bf16_umma_pair<false>;
  static bool attr_set[2] = {false, false};
  static int clusters_mc[2] = {0, 0}, clusters_pair[2] = {0, 0};
  if (!attr_set[vi]) {
    cudaError_t e = cudaFuncSetAttribute(k1, cudaFuncAttributeMaxDynamicSharedMemorySize, umma::SMEM_BYTES);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(k2, cudaFuncAttributeMaxDynamicSharedMemorySize, umma::SMEM_BYTES);
    if (e == cudaSuccess) e = cudaFuncSetAttribute(k3, cudaFuncAttributeMaxDynamicSharedMemorySize, umma::pair::P_SMEM_BYTES);
    if (e != cudaSuccess) {
      set_error("cudaFuncSetAttribute(k_gemm_bf16_umma)", e);
      return -1;
    }
    clusters_mc[vi] = max_clusters(k2, umma::SMEM_BYTES);
    clusters_pair[vi] = max_clusters(k3, umma::pair::P_SMEM_BYTES);
    attr_set[vi] = true;
  }
  const int ntn = (rows + umma::BN - 1) / umma::BN, ntm = (batch + umma::BM - 1) / umma::BM;
  // a pair kernel pays when there are at least two batch tiles to pair; its operands live on this GPU
  int mode = ntm >= 2 ? g_kt_cluster : 0;
  if (mode == 2 && clusters_pair[vi] < 1) mode = 1;
  if (mode == 1 && clusters_mc[vi] < 1) mode = 0;
  // entries are copied out at once: a later miss may recycle the slot an earlier hit pointed to
  MapEntry ex, ey, ew;
  const MapEntry *q = cached_map(Xb, batch, cols, ldx, umma::BM);
  bool ok = q != nullptr;
  if (ok) ex = *q;
  ey = ex;
  if (ok && Yb) {
    q = cached_map(Yb, batch, ldyb, ldyb, umma::BM);  // the pad columns are written too (zeros)
    ok = q != nullptr;
    if (ok) ey = *q;
  }
  if (ok && mode) {
    q = cached_map(Wb, rows, cols, ldw, umma::BN / 2);
    ok = q != nullptr;
    if (ok) ew = *q;
    if (!(ok && ex.local && ey.local && ew.local)) mode = 0;
  }
  if (ok && !mode) {
    q = cached_map(Wb, rows, cols, ldw, umma::BN);
    ok = q != nullptr;
    if (ok) ew = *q;
  }
  if (!ok) {
    set_error_msg("hpnn_dev_gemm_bf16: cuTensorMapEncodeTiled failed");
    return -1;
  }
  cudaError_t e;
  if (mode) {
    const int npairs = ntn * ((ntm + 1) / 2);
    const int avail = mode == 2 ? clusters_pair[vi] : clusters_mc[vi];
    const int ncl = npairs < avail ? npairs : avail;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(2 * ncl, 1, 1);
    cfg.blockDim = dim3(umma::THREADS, 1, 1);
    cfg.dynamicSmemBytes = mode == 2 ? umma::pair::P_SMEM_BYTES : umma::SMEM_BYTES;
    cfg.stream = (cudaStream_t)stream;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = 2;
    at[0].val.clusterDim.y = 1;
    at[0].val.clusterDim.z = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    double *y = Yb ? nullptr : Y;
    const int ldy_ = Yb ? ldyb : ldy;
    if (mode == 2 && Yb)
      e = cudaLaunchKernelEx(&cfg, umma::k_gemm_bf16_umma_pair<true>, ex.map, ew.map, ey.map, batch, rows, cols, y, ldy_, act_kind, g_kt_trace);
    else if (mode == 2)
      e = cudaLaunchKernelEx(&cfg, umma::k_gemm_bf16_umma_pair<false>, ex.map, ew.map, ey.map, batch, rows, cols, y, ldy_, act_kind, g_kt_trace);
    else if (Yb)
      e = cudaLaunchKernelEx(&cfg, umma::k_gemm_bf16_umma<true, 2>, ex.map, ew.map, ey.map, batch, rows, cols, y, ldy_, act_kind, g_kt_trace);
    else
      e = cudaLaunchKernelEx(&cfg, umma::k_gemm_bf16_umma<false, 2>, ex.map, ew.map, ey.map, batch, rows, cols, y, ldy_, act_kind, g_kt_trace);
    if (e != cudaSuccess) {
      set_error("k_gemm_bf16_umma (cluster launch)", e);
      return -1;
    }
  } else {
    const int ntiles = ntn * ntm;
    const int grid = ntiles < sms ? ntiles : sms;
    if (Yb)
      umma::k_gemm_bf16_umma<true, 1><<<grid, umma::THREADS, umma::SMEM_BYTES, (cudaStream_t)stream>>>(
          ex.map, ew.map, ey.map, batch, rows, cols, nullptr, ldyb, act_kind, g_kt_trace);
    else
      umma::k_gemm_bf16_umma<false, 1><<<grid, umma::THREADS, umma::SMEM_BYTES, (cudaStream_t)stream>>>(
          ex.map, ew.map, ey.map, batch, rows, cols, Y, ldy, act_kind, g_kt_trace);
    e = cudaGetLastError();
    if (e != cudaSuccess) {
      set_error("k_gemm_bf16_umma", e);
      return -1;
    }
  }
  ++g_launches;
  return 0;
}

// 2: CTA pairs on 256 x 256 tiles (cta_group::2, default), 1: CTA pairs sharing W through TMA multicast,
// 0: one CTA per tile, -1: environment HPNN_KT_CLUSTER
// device buffer of at least 160 int64 that CTA 0 of every following K-T launch fills with clock64 stamps; null: off
void hpnn_dev_set_kt_trace(void *buf) { g_kt_trace = (long long *)buf; }
void hpnn_dev_set_kt_cluster(int mode) { g_kt_cluster = mode < 0 ? -1 : (mode > 2 ? 2 : mode); }

// Wb: [rows x ldw] bf16, Xb: [batch x ldx] bf16 (ldw, ldx multiples of 8, base 16-byte aligned); FP64 result
int hpnn_dev_gemm_bf16(const void *Wb, int ldw, int rows, int cols, const void *Xb, int ldx, int batch, double *Y,
                       int ldy, int act_kind, void *stream) {
  return launch_umma(Wb, ldw, rows, cols, Xb, ldx, batch, Y, ldy, nullptr, 0, act_kind, stream);
}
// same GEMM, bf16 result [batch x ldyb] for the next layer's GEMM; the pad columns rows..ldyb-1 are written as
// zeros (ldyb <= the next multiple of 256 of rows).  Leading dimensions that are multiples of 64 elements keep
// every 128-byte box row of the TMA loads and stores inside one cache line.
int hpnn_dev_gemm_bf16_bf16(const void *Wb, int ldw, int rows, int cols, const void *Xb, int ldx, int batch, void *Yb,
                            int ldyb, int act_kind, void *stream) {
  if (!Yb || ldyb < rows || ldyb > ((rows + umma::BN - 1) / umma::BN) * umma::BN) {
    set_error_msg("hpnn_dev_gemm_bf16_bf16: null output or leading dimension outside [rows, next multiple of 256]");
    return -1;
  }
  return launch_umma(Wb, ldw, rows, cols, Xb, ldx, batch, nullptr, 0, Yb, ldyb, act_kind, stream);
}

}  // extern "C"
