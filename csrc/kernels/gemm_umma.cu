// K-T: batched forward on the 5th-generation tensor cores (opt-in reduced precision).
//
//   Y[b, n] = f( sum_k X[b, k] * W[n, k] )     X: [B x K] bf16, W: [N x K] bf16 (both K-major)
//
// The reference has no batched path and no reduced precision (batch = 1, FP64), so
// this kernel is never used for parity: it serves `run_nn --batch` / HpnnNet.forward_batch with
// precision="bf16".  Structure (one 128x128 output tile per CTA, 192 threads):
//   warp 0      TMA producer: cp.async.bulk.tensor (128B-swizzled 128x64 bf16 boxes of X and W)
//               into a 3-stage shared-memory ring, completion on mbarriers;
//   warp 1      allocates 128 TMEM columns, one elected lane issues tcgen05.mma (M=128, N=128,
//               K=16, kind::f16, fp32 accumulate in TMEM); tcgen05.commit frees ring slots and
//               finally signals the epilogue;
//   warps 2..5  epilogue: tcgen05.ld (32 lanes x 32 columns per instruction), activation in
//               fp32 (the operands were bf16 anyway), direct stores widened to FP64.
// Out-of-range rows/columns of partial tiles are zero-filled by TMA and masked at the store.
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "dev_common.cuh"

namespace hpnn {
extern long long g_launches;
void set_error(const char *what, cudaError_t e);
void set_error_msg(const char *what);

namespace umma {

constexpr int BM = 128, BN = 128, BK = 64;       // tile (BK bf16 = one 128-byte swizzle row)
constexpr int UMMA_K = 16;                       // bf16 MMA depth
constexpr int STAGES = 3;                        // 3 x 32 KB: two CTAs fit one SM, so one CTA's
                                                 // epilogue overlaps the other's MMA main loop
constexpr int A_BYTES = BM * BK * 2, B_BYTES = BN * BK * 2;
constexpr int STAGE_BYTES = A_BYTES + B_BYTES;   // 32 KB
constexpr int TMEM_COLS = 128;
constexpr int THREADS = 192;
constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024 /*alignment slack*/ + 256 /*barriers*/;

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

// instruction descriptor: D=f32, A=B=bf16, both K-major, N=128, M=128
constexpr uint32_t kIdesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);

__global__ void __launch_bounds__(THREADS, 2)
    k_gemm_bf16_umma(const __grid_constant__ CUtensorMap map_x, const __grid_constant__ CUtensorMap map_w, int batch,
                     int rows, int k_total, double *__restrict__ Y, int ldy, int act_kind) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t *smem = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint64_t *bars = reinterpret_cast<uint64_t *>(smem + STAGES * STAGE_BYTES);
  uint64_t *full = bars, *empty = bars + STAGES, *tmem_full = bars + 2 * STAGES;
  uint32_t *tmem_slot = reinterpret_cast<uint32_t *>(bars + 2 * STAGES + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n0 = blockIdx.x * BN, m0 = blockIdx.y * BM;
  const int nk = (k_total + BK - 1) / BK;

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&map_x)) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&map_w)) : "memory");
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], 1);
    }
    mbar_init(tmem_full, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {  // whole warp: TMEM allocation
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                 "r"((uint32_t)TMEM_COLS)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    if (lane == 0) {
      for (int kb = 0; kb < nk; ++kb) {
        const int s = kb % STAGES;
        const uint32_t ph = (kb / STAGES) & 1;
        mbar_wait(&empty[s], ph ^ 1);
        mbar_expect_tx(&full[s], STAGE_BYTES);
        uint8_t *a = smem + s * STAGE_BYTES;
        tma_load_2d(a, &map_x, &full[s], kb * BK, m0);
        tma_load_2d(a + A_BYTES, &map_w, &full[s], kb * BK, n0);
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      for (int kb = 0; kb < nk; ++kb) {
        const int s = kb % STAGES;
        const uint32_t ph = (kb / STAGES) & 1;
        mbar_wait(&full[s], ph);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        const uint32_t a_addr = smem_u32(smem + s * STAGE_BYTES);
        const uint64_t adesc = smem_desc_sw128(a_addr), bdesc = smem_desc_sw128(a_addr + A_BYTES);
#pragma unroll
        for (int k = 0; k < BK / UMMA_K; ++k) {
          // advancing K inside the swizzle atom = +32 bytes on the start address field
          umma_bf16(tmem_base, adesc + (uint64_t)(2 * k), bdesc + (uint64_t)(2 * k), kIdesc, (kb | k) ? 1u : 0u);
        }
        umma_commit(&empty[s]);                   // slot reusable once these MMAs have read it
        if (kb == nk - 1) umma_commit(tmem_full); // accumulator complete
      }
    }
  } else {
    // epilogue: warp w may touch TMEM lanes 32*(w%4) .. +31
    const int q = warp & 3;
    mbar_wait(tmem_full, 0);
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const int b = m0 + q * 32 + lane;
#pragma unroll 1
    for (int c0 = 0; c0 < BN; c0 += 32) {
      uint32_t v[32];
      tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)c0, v);
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
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 1) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"((uint32_t)TMEM_COLS)
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

// Wb: [rows x ldw] bf16, Xb: [batch x ldx] bf16 (ldw, ldx multiples of 8, base 16-byte aligned)
int hpnn_dev_gemm_bf16(const void *Wb, int ldw, int rows, int cols, const void *Xb, int ldx, int batch, double *Y,
                       int ldy, int act_kind, void *stream) {
  if ((ldw % 8) || (ldx % 8) || ((uintptr_t)Wb & 15) || ((uintptr_t)Xb & 15)) {
    set_error_msg("hpnn_dev_gemm_bf16: operands must be 16-byte aligned with leading dimensions % 8 == 0");
    return -1;
  }
  CUtensorMap mx, mw;
  if (!umma::make_map(&mx, Xb, batch, cols, ldx, umma::BM) || !umma::make_map(&mw, Wb, rows, cols, ldw, umma::BN)) {
    set_error_msg("hpnn_dev_gemm_bf16: cuTensorMapEncodeTiled failed");
    return -1;
  }
  cudaError_t e = cudaFuncSetAttribute(umma::k_gemm_bf16_umma, cudaFuncAttributeMaxDynamicSharedMemorySize, umma::SMEM_BYTES);
  if (e != cudaSuccess) {
    set_error("cudaFuncSetAttribute(k_gemm_bf16_umma)", e);
    return -1;
  }
  dim3 grid((rows + umma::BN - 1) / umma::BN, (batch + umma::BM - 1) / umma::BM);
  umma::k_gemm_bf16_umma<<<grid, umma::THREADS, umma::SMEM_BYTES, (cudaStream_t)stream>>>(mx, mw, batch, rows, cols, Y, ldy,
                                                                                         act_kind);
  e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_error("k_gemm_bf16_umma", e);
    return -1;
  }
  ++g_launches;
  return 0;
}

}  // extern "C"
