// Batched-decode quantized matmul for 8-bit weights on the INTEGER tensor cores (tcgen05.mma kind::i8), sm_100a.
//
// y[m, n] = sum_g ( s[n, g] * sum_{k in g} q[n, k] x[m, k]  +  b[n, g] * sum_{k in g} x[m, k] ),  9..32 activation rows.
//
// At these batch sizes the bf16 forms (pie_qmv*.cu, pie_qmd.cu) are bound by the instructions that turn packed integers
// into bf16 operands: >= 1 issue slot per weight caps them near 1.2 TB/s of weight traffic.  MLX's 8-bit layout is already
// what an integer MMA wants -- row n of the matrix is K consecutive uint8 -- so here the weights go from HBM to shared
// memory by TMA (128-byte swizzle) and straight into the MMA as its uint8 A operand: no instruction touches a weight.
//   * qmi_prep_kernel (one CTA per activation row): RMS norm (fused, as the GEMV's staging), then the row as two int8
//     planes, x ~= rscale[m] * (128 * hi + lo) with |128 hi + lo| <= 16383 (15-bit fixed point against the row maximum: finer
//     than the bf16 rounding of x for every element above 2^-6 of the maximum, 3e-5 of the maximum below), and the exact
//     per-group sums of x for the bias term.  The planes are the MMA's int8 B operand: 64 "columns" = 32 rows x {hi, lo}.
//   * qmi_kernel, persistent, 6 warps: warp 0 feeds a 6-stage ring (128 weight rows x 128 k = 16 KB + the planes' 64 x 128
//     = 8 KB per stage) by TMA; warp 1 issues, per 64-element quantization group, two tcgen05.mma M128 N64 K32 into one of
//     four TMEM accumulators (int32, exact); warps 2..5 (one TMEM lane quadrant each, a lane = a weight row) read a group's
//     accumulator back, form 128 hi + lo exactly in int32, and apply that group's scale and bias in fp32.
//   * few-tile matrices are split along K over CTAs; partials meet in the workspace and the last slice to arrive sums them
//     in slice order (deterministic), as in pie_qmd.cu.  Epilogues as the GEMV: store / residual add / SwiGLU (a unit then
//     runs the gate tile and the up tile one after the other, same lanes).
#include <cuda.h>

#include <cstdlib>
#include <cstring>
#include <mutex>
#include <unordered_map>

#include "pie_internal.h"
#include "pie_ptx.cuh"
#include "pie_umma.cuh"

namespace pie {

constexpr int kQiRows = 128;     // weight rows per tile (MMA M)
constexpr int kQiBK = 128;       // k per stage: one 128-byte swizzle atom, two quantization groups
constexpr int kQiMb = 32;        // activation rows (MMA N = 2 planes x 32)
constexpr int kQiStages = 6;     // (8 measured no faster: the stream is not depth-limited)
constexpr int kQiABytes = kQiRows * kQiBK, kQiBBytes = 2 * kQiMb * kQiBK, kQiStageBytes = kQiABytes + kQiBBytes;
constexpr int kQiAcc = 4;        // TMEM accumulator sets (one per stage in flight: 2 groups x 64 columns each)
constexpr int kQiThreads = 320;   // warp 0 TMA, warp 1 MMA, warps 2..9 epilogue (two per TMEM lane quadrant: 16 rows of the batch each)
constexpr int kQiMh = kQiMb / 2;
// epilogue operands travel by TMA too, one batch per 4 stages (8 groups): scales and biases of the tile's 128 rows
// (16 B per row each) and the activations' group sums (8 x 32 floats)
constexpr int kQiBatchStages = 4, kQiBatches = 3;
constexpr int kQiSbBytes = kQiRows * 16, kQiGsBytes = 8 * kQiMb * 4, kQiBatchBytes = 2 * kQiSbBytes + kQiGsBytes;
constexpr int kQiSmem = kQiStages * kQiStageBytes + kQiBatches * kQiBatchBytes + 1024;
constexpr int kQiMaxK = 16384;   // the prep kernel keeps a row in registers
// D = s32, A = u8, B = s8, both K-major, N = 64, M = 128   (cute::UMMA::InstrDescriptor)
constexpr uint32_t kQiIdesc = (2u << 4) | (0u << 7) | (1u << 10) | ((64u >> 3) << 17) | ((128u >> 4) << 24);

struct QmiSeg { const uint16_t* s; const uint16_t* b; int N, tile0, out_off; };
struct QmiArgs {
  int nseg, dual, M, K, ntiles, ksplit, nstages, epi, out_ld, dbg;
  QmiSeg seg[kMaxSeg];
  uint16_t* out;
  const uint16_t* residual;
  const float* rscale;   // [32]
  const float* gsum;     // [K / 64][32]
  float* ws_part;        // [ntiles][ksplit][32][128]
  unsigned int* ws_cnt;  // [ntiles][4]
  CUtensorMap wmap[kMaxSeg], smap[kMaxSeg], bmap[kMaxSeg], xmap;
};

__device__ __forceinline__ void umma_i8(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t accum) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}\n" ::"r"(tmem_d), "l"(da), "l"(db), "r"(idesc),
      "r"(accum)
      : "memory");
}

// ---- activations -> int8 planes + row scales + group sums ------------------------------------------------------------
__global__ void __launch_bounds__(1024) qmi_prep_kernel(int8_t* xq, float* rscale, float* gsum, const uint16_t* x, int x_ld,
                                                        const uint16_t* norm_w, float eps, int M, int K) {
  __shared__ float red[64];
  const int m = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  pdl_launch_dependents();   // the matmul's set-up (barriers, TMEM, descriptors) overlaps this kernel; it waits before loading
  pdl_wait();                // launched as a programmatic dependent itself: x is the previous kernel's output
  const int nchunks = K / 8;   // K is a multiple of 128
  uint4 xb[2], wb[2];
  float ss = 0.f;
#pragma unroll
  for (int u = 0; u < 2; ++u) {
    const int c = u * 1024 + tid;
    xb[u] = wb[u] = make_uint4(0, 0, 0, 0);
    if (m < M && c < nchunks) {
      xb[u] = ldcg128(x + (size_t)m * x_ld + (size_t)c * 8);
      if (norm_w) wb[u] = __ldg(reinterpret_cast<const uint4*>(norm_w) + c);
      const uint32_t x4[4] = {xb[u].x, xb[u].y, xb[u].z, xb[u].w};
#pragma unroll
      for (int j = 0; j < 4; ++j) { const float lo = bf16lo(x4[j]), hi = bf16hi(x4[j]); ss = fmaf(lo, lo, ss); ss = fmaf(hi, hi, ss); }
    }
  }
  float inv = 1.f;
  if (norm_w) {
    ss = warp_sum(ss);
    if (lane == 0) red[warp] = ss;
    __syncthreads();
    float tot = 0.f;
    for (int w = 0; w < 32; ++w) tot += red[w];
    inv = 1.0f / sqrtf(tot / (float)K + eps);
  }
  float v[2][8], amax = 0.f;
#pragma unroll
  for (int u = 0; u < 2; ++u) {
    const uint32_t x4[4] = {xb[u].x, xb[u].y, xb[u].z, xb[u].w}, w4[4] = {wb[u].x, wb[u].y, wb[u].z, wb[u].w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      float lo = bf16lo(x4[j]), hi = bf16hi(x4[j]);
      if (norm_w) {   // mx.fast.rms_norm: bf16(bf16(x * inv) * w)   (MLX/mlx/fast.cpp:93-111)
        lo = bfround(bfround(lo * inv) * bf16lo(w4[j]));
        hi = bfround(bfround(hi * inv) * bf16hi(w4[j]));
      }
      v[u][2 * j] = lo; v[u][2 * j + 1] = hi;
      amax = fmaxf(amax, fmaxf(fabsf(lo), fabsf(hi)));
    }
  }
  amax = warp_max(amax);
  if (lane == 0) red[32 + warp] = amax;
  __syncthreads();
  amax = 0.f;
  for (int w = 0; w < 32; ++w) amax = fmaxf(amax, red[32 + w]);
  const bool sane = amax > 0.f && amax < 3.0e38f;
  const float sc = sane ? 16383.f / amax : 0.f;
  if (tid == 0) rscale[m] = sane ? amax / 16383.f : 0.f;
#pragma unroll
  for (int u = 0; u < 2; ++u) {
    const int c = u * 1024 + tid;
    uint32_t hi[2] = {0, 0}, lo[2] = {0, 0};
    float s = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int q = __float2int_rn(v[u][j] * sc);   // |q| <= 16383
      const int h = q >> 7, l = q & 127;            // q = 128 h + l, h in [-128, 127], l in [0, 127]
      hi[j >> 2] |= (uint32_t)(h & 255) << (8 * (j & 3));
      lo[j >> 2] |= (uint32_t)l << (8 * (j & 3));
      s += v[u][j];
    }
    s += __shfl_xor_sync(0xffffffffu, s, 1);
    s += __shfl_xor_sync(0xffffffffu, s, 2);
    s += __shfl_xor_sync(0xffffffffu, s, 4);
    if (c < nchunks) {
      *reinterpret_cast<uint2*>(xq + (size_t)m * K + (size_t)c * 8) = make_uint2(hi[0], hi[1]);
      *reinterpret_cast<uint2*>(xq + (size_t)(kQiMb + m) * K + (size_t)c * 8) = make_uint2(lo[0], lo[1]);
      if ((c & 7) == 0) gsum[(size_t)(c >> 3) * kQiMb + m] = s;
    }
  }
}

__device__ __forceinline__ void qi_tile_seg(const QmiArgs& a, int tile, int pass, int& si, int& nb) {
  if (a.dual) { si = pass; nb = tile * kQiRows; return; }
  si = 0;
  if (a.nseg > 1 && tile >= a.seg[1].tile0) si = 1;
  if (a.nseg > 2 && tile >= a.seg[2].tile0) si = 2;
  nb = (tile - a.seg[si].tile0) * kQiRows;
}

// ---- the matmul --------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kQiThreads, 1) qmi_kernel(const __grid_constant__ QmiArgs a) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  __shared__ uint64_t full[kQiStages], empty[kQiStages], acc_full[kQiAcc], acc_empty[kQiAcc];
  __shared__ uint64_t sb_full[kQiBatches], sb_empty[kQiBatches];
  __shared__ uint32_t tmem_slot;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const uint32_t smA = smem_u32(smem), smB = smA + kQiStages * kQiABytes, smS = smB + kQiStages * kQiBBytes;
  const uint8_t* batch_base = smem + kQiStages * kQiStageBytes;
  const int units = a.ntiles * a.ksplit, npass = a.dual ? 2 : 1;

  if (warp == 0 && lane < 10) {   // descriptors are cold in every launch: fetch them all at once
    const int i = lane / 3, kind = lane - 3 * i;
    if (lane == 9) tma_prefetch_desc(&a.xmap);
    else if (i < a.nseg) tma_prefetch_desc(kind == 0 ? &a.wmap[i] : kind == 1 ? &a.smap[i] : &a.bmap[i]);
  }
  if (tid == 0) {
    for (int i = 0; i < kQiStages; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
    for (int i = 0; i < kQiAcc; ++i) { mbar_init(&acc_full[i], 1); mbar_init(&acc_empty[i], 8); }
    for (int i = 0; i < kQiBatches; ++i) { mbar_init(&sb_full[i], 1); mbar_init(&sb_empty[i], 8); }
    fence_mbar_init();
    fence_proxy_async();
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_slot)),
                 "n"(kQiAcc * 128) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = tmem_slot;
  pdl_wait();   // the planes, row scales and group sums of qmi_prep_kernel are complete and visible
  pdl_launch_dependents();   // the next kernel may be scheduled as SMs free up; it waits for this grid before it reads
  // slice `sl` of a tile covers stages [kb, ke)
  auto k_range = [&](int sl, int& kb, int& ke) {
    const int nb4 = a.nstages / kQiBatchStages;   // whole batches: K is a multiple of 512
    kb = (int)((long long)sl * nb4 / a.ksplit) * kQiBatchStages;
    ke = (int)((long long)(sl + 1) * nb4 / a.ksplit) * kQiBatchStages;
  };

  if (warp == 0) {
    // ================= TMA producer ==================================================================================
    if (lane == 0) {
      const uint64_t pol = l2_policy_evict_first();
      uint64_t pol_keep;   // the planes are re-read by every CTA for every tile: they should stay in L2
      asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol_keep));
      int st = 0, use = 0, bs = 0, buse = 0;
      for (int u = blockIdx.x; u < units; u += gridDim.x) {
        const int tile = u / a.ksplit, sl = u % a.ksplit;
        int kb, ke;
        k_range(sl, kb, ke);
        for (int pass = 0; pass < npass; ++pass) {
          int si, nb;
          qi_tile_seg(a, tile, pass, si, nb);
          for (int ks = kb; ks < ke; ++ks) {
            if (((ks - kb) & (kQiBatchStages - 1)) == 0) {   // the epilogue's operands for stages ks .. ks + 3
              if (buse > 0) mbar_wait(&sb_empty[bs], (uint32_t)(buse - 1) & 1);
              const uint32_t dst = smS + bs * kQiBatchBytes;
              mbar_arrive_expect_tx(&sb_full[bs], kQiBatchBytes);
              tma_load_2d(dst, &a.smap[si], &sb_full[bs], 2 * ks, nb, pol);
              tma_load_2d(dst + kQiSbBytes, &a.bmap[si], &sb_full[bs], 2 * ks, nb, pol);
              asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                               dst + 2 * kQiSbBytes), "l"(a.gsum + (size_t)2 * ks * kQiMb), "r"(kQiGsBytes), "r"(smem_u32(&sb_full[bs]))
                           : "memory");
              if (++bs == kQiBatches) { bs = 0; ++buse; }
            }
            if (use > 0) mbar_wait(&empty[st], (uint32_t)(use - 1) & 1);
            mbar_arrive_expect_tx(&full[st], (a.dbg & 8) ? kQiABytes : (a.dbg & 16) ? kQiBBytes : kQiStageBytes);
            if (!(a.dbg & 16)) tma_load_2d(smA + st * kQiABytes, &a.wmap[si], &full[st], ks * kQiBK, nb, pol);
            if (!(a.dbg & 8)) tma_load_2d(smB + st * kQiBBytes, &a.xmap, &full[st], ks * kQiBK, 0, pol_keep);
            if (++st == kQiStages) { st = 0; ++use; }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ================= MMA issuer ======================================================================================
    int st = 0, use = 0;
    unsigned gc = 0;   // groups issued: accumulator gc % 4, its use number gc / 4
    for (int u = blockIdx.x; u < units; u += gridDim.x) {
      int kb, ke;
      k_range(u % a.ksplit, kb, ke);
      for (int i = 0; i < npass * (ke - kb); ++i) {
        mbar_wait(&full[st], (uint32_t)use & 1);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        if (lane == 0) {
          const uint64_t da = umma_desc_sw128(smA + st * kQiABytes), db = umma_desc_sw128(smB + st * kQiBBytes);
          const unsigned buf = gc & (kQiAcc - 1), n_use = gc / kQiAcc;   // gc counts stages here
          if (!(a.dbg & 32)) {
            if (n_use > 0) mbar_wait(&acc_empty[buf], (n_use - 1) & 1);   // the epilogue has read this accumulator set
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
#pragma unroll
            for (int g = 0; g < 2; ++g) {
              if (a.dbg & 4) continue;
              const uint32_t d = tmem + buf * 128 + g * 64;
              umma_i8(d, da + 4 * g, db + 4 * g, kQiIdesc, 0u);           // k 64 g .. 64 g + 31  (32-byte steps)
              umma_i8(d, da + 4 * g + 2, db + 4 * g + 2, kQiIdesc, 1u);
            }
            umma_commit(&acc_full[buf]);   // one hand-over per stage
          }
          ++gc;
          umma_commit(&empty[st]);
        }
        __syncwarp();
        if (++st == kQiStages) { st = 0; ++use; }
      }
    }
  } else {
    // ================= epilogue: TMEM lane quadrant warp % 4, lane = weight row, half = which 16 rows of the batch ======
    const int quad = warp & 3, half = (warp - 2) >> 2, row = quad * 32 + lane, m0 = half * kQiMh;
    unsigned gc = 0;
    int bs = 0, buse = 0;
    for (int u = blockIdx.x; u < units; u += gridDim.x) {
      const int tile = u / a.ksplit, sl = u % a.ksplit;
      int kb, ke;
      k_range(sl, kb, ke);
      float gate[kQiMh], acc1[kQiMh], acc2[kQiMh];   // after a pass acc1 holds the pass's y
      float (&y)[kQiMh] = acc1;
      for (int pass = 0; pass < npass; ++pass) {
        if (a.dual && pass == 1) {
#pragma unroll
          for (int m = 0; m < kQiMh; ++m) gate[m] = acc1[m];
        }
#pragma unroll
        for (int m = 0; m < kQiMh; ++m) { acc1[m] = 0.f; acc2[m] = 0.f; }
        for (int ks = kb; ks < ke; ks += kQiBatchStages) {
          // this batch's scales / biases of this lane's row (8 groups each) and the group sums, brought by the producer
          mbar_wait(&sb_full[bs], (uint32_t)buse & 1);
          const uint8_t* bb = batch_base + bs * kQiBatchBytes;
          const uint4 sq = *reinterpret_cast<const uint4*>(bb + row * 16);
          const uint4 bq = *reinterpret_cast<const uint4*>(bb + kQiSbBytes + row * 16);
          const uint32_t sw[4] = {sq.x, sq.y, sq.z, sq.w}, bw[4] = {bq.x, bq.y, bq.z, bq.w};
          const float4* gs_b = reinterpret_cast<const float4*>(bb + 2 * kQiSbBytes);
#pragma unroll
          for (int sg = 0; sg < kQiBatchStages; ++sg) {
            const unsigned buf = gc & (kQiAcc - 1), n_use = gc / kQiAcc;   // gc counts stages
            ++gc;
            if (a.dbg & 32) continue;
            mbar_wait(&acc_full[buf], n_use & 1);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            uint32_t hi[2][kQiMh], lo[2][kQiMh];
            if (!(a.dbg & 2)) {
              const uint32_t t0 = tmem + buf * 128 + m0 + ((uint32_t)(quad * 32) << 16);
              tmem_ld16(t0, hi[0]); tmem_ld16(t0 + 32, lo[0]); tmem_ld16(t0 + 64, hi[1]); tmem_ld16(t0 + 96, lo[1]);
              tmem_ld_wait();
            }
            asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
            __syncwarp();
            if (lane == 0) mbar_arrive(&acc_empty[buf]);
            if (a.dbg & 1) continue;
#pragma unroll
            for (int g2 = 0; g2 < 2; ++g2) {
              const int g = 2 * sg + g2;
              const float s = (g & 1) ? bf16hi(sw[g >> 1]) : bf16lo(sw[g >> 1]);
              const float b = (g & 1) ? bf16hi(bw[g >> 1]) : bf16lo(bw[g >> 1]);
#pragma unroll
              for (int m4 = 0; m4 < kQiMh / 4; ++m4) {
                const float4 q = gs_b[g * (kQiMb / 4) + half * (kQiMh / 4) + m4];
                const float qv[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                  const int m = 4 * m4 + j;
                  const int c = (int)hi[g2][m] * 128 + (int)lo[g2][m];   // exact: |c| < 2^29
                  acc1[m] = fmaf(s, (float)c, acc1[m]);
                  acc2[m] = fmaf(b, qv[j], acc2[m]);
                }
              }
            }
          }
          __syncwarp();
          if (lane == 0) mbar_arrive(&sb_empty[bs]);
          if (++bs == kQiBatches) { bs = 0; ++buse; }
        }
#pragma unroll
        for (int m = 0; m < kQiMh; ++m) acc1[m] = fmaf(__ldg(a.rscale + m0 + m), acc1[m], acc2[m]);
      }
      int si, nb;
      qi_tile_seg(a, tile, 0, si, nb);
      const int n = nb + row;
      if (a.ksplit > 1) {   // (never with dual) partial -> workspace; the last slice to arrive sums them in slice order
        float* part = a.ws_part + ((size_t)tile * a.ksplit) * (kQiMb * kQiRows) + row;
#pragma unroll
        for (int m = 0; m < kQiMh; ++m) part[((size_t)sl * kQiMb + m0 + m) * kQiRows] = y[m];
        __threadfence();
        __syncwarp();
        unsigned last = 0;
        if (lane == 0) last = atomicAdd(&a.ws_cnt[tile * 8 + half * 4 + quad], 1u) == (unsigned)a.ksplit - 1;
        last = __shfl_sync(0xffffffffu, last, 0);
        if (!last) continue;
        __threadfence();
#pragma unroll
        for (int m = 0; m < kQiMh; ++m) y[m] = 0.f;
        for (int s2 = 0; s2 < a.ksplit; ++s2) {
#pragma unroll
          for (int m = 0; m < kQiMh; ++m) y[m] += ldcg_f32(part + ((size_t)s2 * kQiMb + m0 + m) * kQiRows);
        }
        if (lane == 0) a.ws_cnt[tile * 8 + half * 4 + quad] = 0;
      }
      if (a.epi == EPI_SILUMUL) {
#pragma unroll
        for (int m = 0; m < kQiMh; ++m) {
          if (m0 + m < a.M) {
            const float gt = bfround(gate[m]), up = bfround(y[m]);
            const float sg = bfround(1.f / (1.f + expf(-gt)));  // nn.silu = x * sigmoid(x), each op rounded
            a.out[(size_t)(m0 + m) * a.out_ld + n] = f2bf(bfround(gt * sg) * up);
          }
        }
      } else {
        const size_t col = (size_t)a.seg[si].out_off + n;
#pragma unroll
        for (int m = 0; m < kQiMh; ++m) {
          if (m0 + m < a.M) {
            float res = bfround(y[m]);
            if (a.epi == EPI_RESADD) res = bf16lo((uint32_t)a.residual[(size_t)(m0 + m) * a.out_ld + col]) + res;
            a.out[(size_t)(m0 + m) * a.out_ld + col] = f2bf(res);
          }
        }
      }
      __syncwarp();
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 1) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(kQiAcc * 128) : "memory");
}

// ---- host side ---------------------------------------------------------------------------------------------------------
typedef CUresult (*QiEncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                               const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                               CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static QiEncodeFn qi_encode_fn() {
  static QiEncodeFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<QiEncodeFn>(p);
  });
  return fn;
}
// uint8 matrix [rows][K] (K contiguous) as a 2-D map, box 128 bytes x `box_rows`, 128-byte swizzle
static bool qi_map(CUtensorMap* out, const void* ptr, long long rows, int K, int box_rows) {
  struct Key { const void* p; long long r; int k, b; bool operator==(const Key& o) const { return p == o.p && r == o.r && k == o.k && b == o.b; } };
  struct Hash { size_t operator()(const Key& k) const { return reinterpret_cast<size_t>(k.p) ^ ((size_t)k.r * 1315423911u) ^ ((size_t)k.k << 20) ^ (size_t)k.b; } };
  static std::unordered_map<Key, CUtensorMap, Hash> cache;
  static std::mutex mu;
  const Key key{ptr, rows, K, box_rows};
  std::lock_guard<std::mutex> lk(mu);
  auto it = cache.find(key);
  if (it != cache.end()) { *out = it->second; return true; }
  QiEncodeFn enc = qi_encode_fn();
  if (!enc) return false;
  const cuuint64_t dims[2] = {(cuuint64_t)K, (cuuint64_t)rows};
  const cuuint64_t strides[1] = {(cuuint64_t)K};
  const cuuint32_t box[2] = {(cuuint32_t)kQiBK, (cuuint32_t)box_rows};
  const cuuint32_t es[2] = {1, 1};
  CUtensorMap m;
  if (enc(&m, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, const_cast<void*>(ptr), dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
          CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
    return false;
  if (cache.size() > 65536) cache.clear();
  cache.emplace(key, m);
  *out = m;
  return true;
}

// bf16 scales / biases [rows][K / 64] as a 2-D map, box 8 groups (16 bytes) x 128 rows, no swizzle
static bool qi_sb_map(CUtensorMap* out, const void* ptr, long long rows, int K) {
  struct Key { const void* p; long long r; int k; bool operator==(const Key& o) const { return p == o.p && r == o.r && k == o.k; } };
  struct Hash { size_t operator()(const Key& k) const { return reinterpret_cast<size_t>(k.p) ^ ((size_t)k.r * 2654435761u) ^ ((size_t)k.k << 24); } };
  static std::unordered_map<Key, CUtensorMap, Hash> cache;
  static std::mutex mu;
  const Key key{ptr, rows, K};
  std::lock_guard<std::mutex> lk(mu);
  auto it = cache.find(key);
  if (it != cache.end()) { *out = it->second; return true; }
  QiEncodeFn enc = qi_encode_fn();
  if (!enc) return false;
  const cuuint64_t dims[2] = {(cuuint64_t)(K / kGroup), (cuuint64_t)rows};
  const cuuint64_t strides[1] = {(cuuint64_t)(K / kGroup) * 2};
  const cuuint32_t box[2] = {8, (cuuint32_t)kQiRows};
  const cuuint32_t es[2] = {1, 1};
  CUtensorMap m;
  if (enc(&m, CU_TENSOR_MAP_DATA_TYPE_UINT16, 2, const_cast<void*>(ptr), dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
          CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
    return false;
  if (cache.size() > 65536) cache.clear();
  cache.emplace(key, m);
  *out = m;
  return true;
}

bool qmi_supported(int mb, int K, int x_ld, int group_size, int bits, const pie_qmv_seg* segs, int nseg, int dual) {
  if (bits != 8 || group_size != 64 || mb < 1 || mb > kQiMb || K % (kQiBK * kQiBatchStages) != 0 || K > kQiMaxK || x_ld % 8 != 0)
    return false;
  if (dual && (nseg != 2 || segs[0].N != segs[1].N)) return false;
  for (int i = 0; i < nseg; ++i) {
    if (segs[i].N % kQiRows != 0) return false;
    if ((reinterpret_cast<uintptr_t>(segs[i].wq) & 15) || (reinterpret_cast<uintptr_t>(segs[i].scales) & 15) ||
        (reinterpret_cast<uintptr_t>(segs[i].biases) & 15))
      return false;
  }
  return qi_encode_fn() != nullptr;
}

// K slices per tile: enough units to fill the SMs, at least 4 stages per unit, at most 8 slices; none with SwiGLU
int qmi_pick_ksplit(int ntiles, int K, int dual) {
  if (dual) return 1;
  const int stages = K / kQiBK, sms = num_sms();
  int best = 1;
  double best_eff = 0.0;
  for (int d = 1; d <= 8; ++d) {
    if (d > 1 && stages / d < 2 * kQiBatchStages) continue;
    const long long units = (long long)ntiles * d;
    const double eff = (double)units / (double)(((units + sms - 1) / sms) * sms);
    if (eff > best_eff + 0.02) { best_eff = eff; best = d; }
  }
  return best;
}

// workspace behind the shared counter region: planes, row scales, group sums, split-K partials
size_t qmi_ws_bytes(int ntiles, int K, int ksplit) {
  size_t b = (size_t)2 * kQiMb * K;                                   // int8 planes
  b += 256;                                                            // row scales
  b += (((size_t)(K / kGroup) * kQiMb * 4) + 255) & ~(size_t)255;      // group sums
  if (ksplit > 1) b += (size_t)ntiles * ksplit * kQiMb * kQiRows * 4;
  return b;
}

cudaError_t launch_qmi(void* out, int out_ld, const void* x, int x_ld, int M, int K, const pie_qmv_seg* segs, int nseg,
                       const void* norm_w, float eps, int epi, const void* residual, int ksplit, uint8_t* ws,
                       unsigned int* ws_cnt, cudaStream_t st) {
  QmiArgs a;
  memset(&a, 0, sizeof(a));
  const int dual = epi == EPI_SILUMUL;
  int tile = 0, off = 0;
  for (int i = 0; i < nseg; ++i) {
    a.seg[i].s = (const uint16_t*)segs[i].scales; a.seg[i].b = (const uint16_t*)segs[i].biases;
    a.seg[i].N = segs[i].N; a.seg[i].tile0 = tile; a.seg[i].out_off = off;
    if (!qi_map(&a.wmap[i], segs[i].wq, segs[i].N, K, kQiRows)) return cudaErrorInvalidValue;
    if (!qi_sb_map(&a.smap[i], segs[i].scales, segs[i].N, K) || !qi_sb_map(&a.bmap[i], segs[i].biases, segs[i].N, K))
      return cudaErrorInvalidValue;
    tile += segs[i].N / kQiRows;
    off += segs[i].N;
  }
  a.nseg = nseg; a.dual = dual; a.M = M; a.K = K; a.epi = epi; a.out_ld = out_ld;
  a.ntiles = dual ? segs[0].N / kQiRows : tile;
  a.ksplit = ksplit; a.nstages = K / kQiBK;
  { const char* e = getenv("PIE_QMI_DBG"); a.dbg = e ? atoi(e) : 0; }
  a.out = (uint16_t*)out; a.residual = (const uint16_t*)residual;
  int8_t* xq = (int8_t*)ws;
  float* rscale = (float*)(ws + (size_t)2 * kQiMb * K);
  float* gsum = rscale + 64;
  a.rscale = rscale; a.gsum = gsum;
  a.ws_part = (float*)((uint8_t*)gsum + ((((size_t)(K / kGroup) * kQiMb * 4) + 255) & ~(size_t)255));
  a.ws_cnt = ws_cnt;
  if (!qi_map(&a.xmap, xq, 2 * kQiMb, K, 2 * kQiMb)) return cudaErrorInvalidValue;
  {  // (per launch, not once per process: the attribute is per device)
    cudaError_t e = cudaFuncSetAttribute(qmi_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kQiSmem);
    if (e != cudaSuccess) return e;
  }
  count_launch();
  cudaError_t e = cudaSuccess;
  if (!(a.dbg & 64))   // (debug: 64 = time the matmul without its prep launch; results are then stale)
    e = launch_pdl(qmi_prep_kernel, dim3(kQiMb), dim3(1024), 0, st, xq, rscale, gsum, (const uint16_t*)x, x_ld,
                   (const uint16_t*)norm_w, eps, M, K);
  if (e != cudaSuccess) return e;
  const int units = a.ntiles * a.ksplit;
  const int grid = units < num_sms() ? units : num_sms();
  count_launch();
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)grid, 1, 1);
  cfg.blockDim = dim3(kQiThreads, 1, 1);
  cfg.dynamicSmemBytes = kQiSmem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, qmi_kernel, a);
}

}  // namespace pie
