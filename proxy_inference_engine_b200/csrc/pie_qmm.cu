// Prefill quantized GEMM on the 5th-generation tensor cores (tcgen05 + TMEM), sm_100a.
//
// Replaces mx.quantized_matmul(transpose=True) at prefill-sized M (MLX/mlx/ops.cpp:3984-4021; Metal qmm_t,
// MLX/mlx/backend/metal/kernels/quantized.h:954-1078):  y[m, n] = bf16( sum_k x[m, k] * w_hat[n, k] ),
// w_hat = bf16(scale * q + bias): the weight is dequantised ONCE to the activation type into shared memory
// (QuantizedBlockLoader, quantized.h:411-417, 502-512) and the contraction accumulates in fp32 -- the same rounding
// structure here, with the fp32 accumulator in tensor memory.
//
// One CTA computes a 256 x 256 tile of y: two 128-row accumulators of 256 fp32 columns = all 512 TMEM columns.
// Per 64-element K step (= one quantization group) and pipeline stage (3 stages of 32 KB A + 32 KB B):
//   * warps 0-7 (256 threads, thread t <-> activation row t and weight row t of the tile): cp.async the activation
//     row into the 128B-swizzled K-major layout tcgen05 reads; dequantise weight row n0 + t (64 packed values, its
//     scale and bias prefetched into registers one K step ahead) and store it into the swizzled B block;
//     fence.proxy.async; arrive on full[s] one iteration later, when the activation copies have landed.
//   * warp 8, one elected lane: wait full[s]; 2 x 4 tcgen05.mma.kind::f16 (M128 N256 K16, bf16 x bf16 -> fp32) with
//     shared-memory matrix descriptors; tcgen05.commit -> empty[s] hands the stage back to the producers.
//   Tile shape: the bf16 activations are the heavy operand (the 4-bit weights are 1/4 of their bytes), so the tile
//   is made wide in N: 256 x 256 needs 40 KB of L2->SM traffic per 8.4 MFLOP; 256 x 128 measured L2-bound.
//   * epilogue: the 8 warps read (accumulator, TMEM lane quadrant) = (warp / 4, warp % 4) with tcgen05.ld, 32 lanes x
//     32 columns at a time, round to bf16 once and store.
#include "pie_internal.h"
#include "pie_ptx.cuh"
#include "pie_umma.cuh"

namespace pie {

constexpr int kQmmBM = 256, kQmmBN = 256, kQmmBK = 64, kQmmStages = 3;
constexpr int kQmmAStage = kQmmBM * kQmmBK * 2;   // 32 KB
constexpr int kQmmBStage = kQmmBN * kQmmBK * 2;   // 32 KB
constexpr int kQmmSmem = kQmmStages * (kQmmAStage + kQmmBStage) + 1024;  // + alignment slack
constexpr int kQmmTmemCols = 512;
constexpr int kQmmProducers = 256;                   // 8 warps: two threads per tile row (latency hiding for the dequant)
constexpr int kQmmThreads = kQmmProducers + 32;      // + the MMA warp

struct QmmArgs {
  uint16_t* y;
  const uint16_t* x;
  const uint32_t* wq;
  const uint16_t* scales;
  const uint16_t* biases;
  int M, N, K;
};

constexpr uint32_t kQmmIdesc = umma_idesc_bf16_m128(kQmmBN);

template <int BITS>
__global__ void __launch_bounds__(kQmmThreads, 1) qmm_tc_kernel(const __grid_constant__ QmmArgs a) {
  constexpr int S = kQmmStages, LAG = 1;
  constexpr int WPR = kQmmBK * BITS / 32;    // packed words of one row per K step: 8 (q4) / 16 (q8)
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t full_bar[S], empty_bar[S], accum_bar;
  __shared__ uint32_t tmem_slot;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t smA = base, smB = base + S * kQmmAStage;
  const int n0 = blockIdx.x * kQmmBN, m0 = blockIdx.y * kQmmBM;
  const int nk = a.K / kQmmBK;

  if (tid == 0) {
    for (int s = 0; s < S; ++s) { mbar_init(&full_bar[s], kQmmProducers); mbar_init(&empty_bar[s], 1); }
    mbar_init(&accum_bar, 1);
    fence_mbar_init();
  }
  if (warp == kQmmProducers / 32) {  // one warp owns the tensor-memory allocation: 2 accumulators of 256 fp32 columns
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_slot)),
                 "n"(kQmmTmemCols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = tmem_slot;

  if (warp < kQmmProducers / 32) {
    // ---------------- producers: thread t <-> activation row t (of 256) and weight row t (of 256) --------------
    const int t = tid;
    const uint32_t swz = (uint32_t)(t & 7);
    const int nrow = min(n0 + t, a.N - 1);
    const uint4* wsrc = reinterpret_cast<const uint4*>(reinterpret_cast<const uint8_t*>(a.wq) +
                                                       (size_t)nrow * (a.K * BITS / 8));
    const uint16_t* ssrc = a.scales + (size_t)nrow * nk;
    const uint16_t* bsrc = a.biases + (size_t)nrow * nk;
    const uint16_t* xsrc = a.x + (size_t)min(m0 + t, a.M - 1) * a.K;
    // packed words, scale and bias of the NEXT K step travel in registers: one whole iteration hides their latency
    uint4 wn[WPR / 4];
    uint32_t sn, bn;
#pragma unroll
    for (int c = 0; c < WPR / 4; ++c) wn[c] = __ldg(wsrc + c);
    sn = __ldg(ssrc); bn = __ldg(bsrc);
    for (int kb = 0; kb < nk + LAG; ++kb) {
      if (kb < nk) {
        const int s = kb % S, it = kb / S;
        if (it > 0) mbar_wait(&empty_bar[s], (uint32_t)(it - 1) & 1);
        const uint32_t dA = smA + s * kQmmAStage + t * 128;
#pragma unroll
        for (int c = 0; c < 8; ++c) cp_async16(dA + ((c ^ swz) << 4), xsrc + kb * kQmmBK + c * 8);
        cp_async_commit();
        uint4 wc[WPR / 4];
#pragma unroll
        for (int c = 0; c < WPR / 4; ++c) wc[c] = wn[c];
        const float sc = bf16lo(sn), bb = fmaf(-(BITS == 4 ? 128.f : 256.f), sc, bf16lo(bn));
        if (kb + 1 < nk) {
#pragma unroll
          for (int c = 0; c < WPR / 4; ++c) wn[c] = __ldg(wsrc + (size_t)(kb + 1) * (WPR / 4) + c);
          sn = __ldg(ssrc + kb + 1); bn = __ldg(bsrc + kb + 1);
        }
        const uint32_t dB = smB + s * kQmmBStage + t * 128;
#pragma unroll
        for (int c4 = 0; c4 < WPR / 4; ++c4) {
          const uint32_t w[4] = {wc[c4].x, wc[c4].y, wc[c4].z, wc[c4].w};
          if (BITS == 4) {
#pragma unroll
            for (int i = 0; i < 4; ++i) {  // word -> 8 values -> one 16-byte chunk
              uint32_t o[4];
#pragma unroll
              for (int e = 0; e < 4; ++e) {
                const float lo = fmaf(lift_f32<4>(w[i], 8 * e), sc, bb);
                const float hi = fmaf(lift_f32<4>(w[i], 8 * e + 4), sc, bb);
                o[e] = pack_bf16(lo, hi);
              }
              const uint32_t c = (uint32_t)(c4 * 4 + i);
              asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(dB + ((c ^ swz) << 4)), "r"(o[0]), "r"(o[1]),
                           "r"(o[2]), "r"(o[3]) : "memory");
            }
          } else {
#pragma unroll
            for (int i = 0; i < 2; ++i) {  // two words -> 8 values -> one 16-byte chunk
              uint32_t o[4];
#pragma unroll
              for (int e = 0; e < 4; ++e) {
                const uint32_t ww = w[2 * i + (e >> 1)];
                const float lo = fmaf(lift_f32<8>(ww, 16 * (e & 1)), sc, bb);
                const float hi = fmaf(lift_f32<8>(ww, 16 * (e & 1) + 8), sc, bb);
                o[e] = pack_bf16(lo, hi);
              }
              const uint32_t c = (uint32_t)(c4 * 2 + i);
              asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(dB + ((c ^ swz) << 4)), "r"(o[0]), "r"(o[1]),
                           "r"(o[2]), "r"(o[3]) : "memory");
            }
          }
        }
      } else {
        cp_async_commit();
      }
      if (kb >= LAG) {
        const int j = kb - LAG;
        cp_async_wait<LAG>();  // this thread's activation copies of K step j have landed
        fence_proxy_async();   // generic-proxy writes (cp.async, st.shared) -> visible to the tensor core's async proxy
        mbar_arrive(&full_bar[j % S]);
      }
    }
    // ---------------- epilogue: TMEM lane quadrant `warp` -> bf16 rows of y ----------------------------------
    mbar_wait(&accum_bar, 0);
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    {
      const int h = warp >> 2, quad = warp & 3;  // a warp reads TMEM lanes 32 (warp % 4) .. + 31 only
      const int row = m0 + h * 128 + quad * 32 + lane;
#pragma unroll 1
      for (int c = 0; c < kQmmBN / 32; ++c) {
        uint32_t v[32];
        tmem_ld32(tmem + (uint32_t)(h * kQmmBN + c * 32) + ((uint32_t)(quad * 32) << 16), v);
        const int col = n0 + c * 32;
        if (row < a.M) {
          uint16_t* dst = a.y + (size_t)row * a.N + col;
          if (col + 32 <= a.N && (a.N & 7) == 0) {
#pragma unroll
            for (int q = 0; q < 4; ++q) {
              uint4 o;
              o.x = pack_bf16(__uint_as_float(v[8 * q]), __uint_as_float(v[8 * q + 1]));
              o.y = pack_bf16(__uint_as_float(v[8 * q + 2]), __uint_as_float(v[8 * q + 3]));
              o.z = pack_bf16(__uint_as_float(v[8 * q + 4]), __uint_as_float(v[8 * q + 5]));
              o.w = pack_bf16(__uint_as_float(v[8 * q + 6]), __uint_as_float(v[8 * q + 7]));
              reinterpret_cast<uint4*>(dst)[q] = o;
            }
          } else {
            for (int q = 0; q < 32; ++q)
              if (col + q < a.N) dst[q] = f2bf(__uint_as_float(v[q]));
          }
        }
      }
    }
  } else {
    // ---------------- MMA issuer: one elected lane of warp 4 ------------------------------------------------
    for (int kb = 0; kb < nk; ++kb) {
      const int s = kb % S;
      mbar_wait(&full_bar[s], (uint32_t)(kb / S) & 1);
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      if (lane == 0) {
        const uint64_t db0 = umma_desc_sw128(smB + s * kQmmBStage);
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          const uint64_t da0 = umma_desc_sw128(smA + s * kQmmAStage + h * 128 * 128);
#pragma unroll
          for (int k4 = 0; k4 < kQmmBK / 16; ++k4)  // 16 bf16 = 32 bytes along K inside the swizzle atom: +2 units
            umma_bf16(tmem + h * kQmmBN, da0 + 2 * k4, db0 + 2 * k4, kQmmIdesc, (kb > 0 || k4 > 0) ? 1u : 0u);
        }
        umma_commit(&empty_bar[s]);  // arrives when the MMAs above have read the stage
        if (kb == nk - 1) umma_commit(&accum_bar);
      }
      __syncwarp();
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == kQmmProducers / 32)
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(kQmmTmemCols) : "memory");
}

cudaError_t launch_qmm_tc(uint16_t* y, const uint16_t* x, const uint32_t* wq, const uint16_t* s, const uint16_t* b,
                          int M, int N, int K, int bits, cudaStream_t st) {
  QmmArgs a;
  a.y = y; a.x = x; a.wq = wq; a.scales = s; a.biases = b; a.M = M; a.N = N; a.K = K;
  const size_t smem = (size_t)kQmmSmem;
  static unsigned long long configured[2] = {0, 0};   // per device
  const int idx = bits == 4 ? 0 : 1;
  if (first_use_on_device(configured[idx])) {
    cudaError_t e = bits == 4 ? cudaFuncSetAttribute(qmm_tc_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)
                              : cudaFuncSetAttribute(qmm_tc_kernel<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
  }
  const dim3 grid((N + kQmmBN - 1) / kQmmBN, (M + kQmmBM - 1) / kQmmBM);
  count_launch();
  if (bits == 4) qmm_tc_kernel<4><<<grid, kQmmThreads, smem, st>>>(a);
  else qmm_tc_kernel<8><<<grid, kQmmThreads, smem, st>>>(a);
  return cudaGetLastError();
}

}  // namespace pie
