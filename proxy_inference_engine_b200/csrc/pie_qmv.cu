// Quantized matrix-vector product for decode (1 <= M <= 8 activation rows), sm_100a.
//
// Replaces mx.quantized_matmul(transpose=True) at small M (QuantizedLinear.__call__,
// MLX/python/mlx/nn/layers/quantized.py:204-216 -> Metal qmv_fast, MLX/mlx/backend/metal/kernels/quantized.h:623-689):
//     y[m, n] = bf16( sum_k x[m, k] * (scales[n, k/64] * q[n, k] + biases[n, k/64]) ),  fp32 accumulation.
//
// Shape of the kernel
//   * a CTA owns one tile of 16 weight rows and one K slice (gridDim.y = split-K); its warps take the
//     slice's 128-byte-per-row stages round-robin, so the CTA walks contiguous 128*W bytes of every row;
//   * packed weights go global -> shared with 16-byte cp.async (whole 128 B lines, L1 bypass) through a
//     private 4-stage ring per warp (no cross-warp barrier in the main loop); the bf16 scales/biases of
//     the stage ride in the same ring;
//   * dequantisation happens in registers: a 4-bit field is lifted into a bf16 mantissa,
//     ((w >> 4i) & 0x000f000f) | 0x43004300 = two bf16 values 128+q, and fed straight to
//     mma.sync.m16n8k16 as the A operand (16 weight rows x 16 k); the activations are the B operand
//     (k x 8 batch columns), staged once per CTA in shared memory in fragment order.  x*(128+q) is exact
//     in fp32, the group's affine is applied once per 64 elements:
//         y += s * acc + (b - 128 s) * sum_x(group)          (the factoring of qmv_fast, quantized.h:225)
//     8-bit weights use two accumulators (low / high nibble): q = lo + 16 hi.
//   * epilogues: plain store, residual add (language.py:150-154), SwiGLU (language.py:126-127) with
//     tiles made of 8 gate rows + 8 up rows; split-K partials are combined in a fixed order by the
//     last CTA to arrive (deterministic).
#include "pie_internal.h"
#include "pie_ptx.cuh"

namespace pie {

template <int BITS>
struct QmvGeom {
  static constexpr int CK = kRowBytes * 8 / BITS;  // K elements per stage: 256 (q4) / 128 (q8)
  static constexpr int GPC = CK / kGroup;          // groups per stage: 4 / 2
  static constexpr int SB_BYTES = kTileRows * GPC * 2;  // scales (or biases) bytes per stage
  static constexpr int STAGE_BYTES = kTileRows * kRowBytes + 2 * SB_BYTES;
};

__host__ __device__ inline int qmv_xrow_bytes(int kslice) { return kslice * 2 + 16; }

size_t qmv_smem_bytes(int M, int kslice, int bits, int warps, int mt) {
  size_t xs = ((size_t)M * qmv_xrow_bytes(kslice) + 127) & ~(size_t)127;
  size_t sumx = (size_t)(kslice / kGroup) * 8 * mt * sizeof(float);
  size_t red = (size_t)warps * kQmvSlots * 128 * mt * sizeof(float);  // exchange slots
  size_t stage = bits == 4 ? QmvGeom<4>::STAGE_BYTES : QmvGeom<8>::STAGE_BYTES;
  return xs + sumx + red + 512 + 128 + (size_t)warps * kQmvStages * stage;
}

// ---- swizzled position (bytes within the 128-byte stage row) of 16-byte chunk c of tile row r ----------
template <int BITS>
__device__ __forceinline__ int swz_chunk(int r, int c) {
  if (BITS == 4) return ((((c >> 1) ^ (r & 3)) << 1) | (c & 1)) * 16;  // 32-byte groups, 4 rows conflict-free
  return ((((c >> 2) ^ (r & 1)) << 2) | (c & 3)) * 16;                 // 64-byte groups, 2 rows conflict-free
}

// two bf16 values 128 + nibble: ((w >> sh) & 0x000f000f) | 0x43004300.  mask and magic arrive in registers (kernel
// arguments) so that the and-or is ONE LOP3; as immediates ptxas needs two.
__device__ __forceinline__ uint32_t lift4(uint32_t w, int sh, uint32_t mask, uint32_t magic) {
  uint32_t d;
  asm("lop3.b32 %0, %1, %2, %3, 0xEA;" : "=r"(d) : "r"(w >> sh), "r"(mask), "r"(magic));
  return d;
}

// ---- stage x (optionally RMS-normalised) into shared memory in B-fragment order + per-group sums ----
// Within every block of 16 k: q4 position 4m+i <-> k = m + 4i ; q8 position 4m+{0,1,2,3} <-> k = 4m+{0,2,1,3}.
template <int BITS, int MT>
__device__ void stage_x(const QmvArgs& a, int k0, int klen, uint8_t* xs, float* sumx, float* red) {
  const int tid = threadIdx.x, nthr = blockDim.x, lane = tid & 31;
  const int xrow = qmv_xrow_bytes(a.kslice);
  // fused RMSNorm (mx.fast.rms_norm, MLX/mlx/fast.cpp:93-111): inverse rms over the FULL row
  float inv[8 * MT];
  if (a.norm_w) {
    for (int m = 0; m < a.M; ++m) {
      float ss = 0.f;
      const uint4* xr = reinterpret_cast<const uint4*>(a.x + (size_t)m * a.x_ld);
      for (int i = tid; i < a.K / 8; i += nthr) {
        uint4 v = ldcg128(xr + i);
        uint32_t u[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) { float lo = bf16lo(u[j]), hi = bf16hi(u[j]); ss += lo * lo + hi * hi; }
      }
      ss = warp_sum(ss);
      if (lane == 0) red[tid >> 5] = ss;
      __syncthreads();
      float tot = 0.f;
      for (int w = 0; w < (nthr >> 5); ++w) tot += red[w];
      inv[m] = 1.0f / sqrtf(tot / (float)a.K + a.eps);
      __syncthreads();
    }
  }
  const int bpr = klen / 16;             // 16-element blocks per row in this slice (multiple of 4)
  const int total = a.M * bpr;
  const int rounds = (total + nthr - 1) / nthr;
  for (int it = 0; it < rounds; ++it) {
    const int idx = it * nthr + tid;
    const bool ok = idx < total;
    const int m = ok ? idx / bpr : 0, blk = ok ? idx % bpr : 0;
    float v[16];
    if (ok) {
      const uint4* src = reinterpret_cast<const uint4*>(a.x + (size_t)m * a.x_ld + k0 + blk * 16);
      uint4 p0 = ldcg128(src), p1 = ldcg128(src + 1);
      uint32_t u[8] = {p0.x, p0.y, p0.z, p0.w, p1.x, p1.y, p1.z, p1.w};
#pragma unroll
      for (int j = 0; j < 8; ++j) { v[2 * j] = bf16lo(u[j]); v[2 * j + 1] = bf16hi(u[j]); }
      if (a.norm_w) {
        const uint4* wp = reinterpret_cast<const uint4*>(a.norm_w + k0 + blk * 16);
        uint4 w0 = __ldg(wp), w1 = __ldg(wp + 1);
        uint32_t wu[8] = {w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w};
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          v[2 * j] = bfround(bfround(v[2 * j] * inv[m]) * bf16lo(wu[j]));
          v[2 * j + 1] = bfround(bfround(v[2 * j + 1] * inv[m]) * bf16hi(wu[j]));
        }
      }
    } else {
#pragma unroll
      for (int j = 0; j < 16; ++j) v[j] = 0.f;
    }
    float s = 0.f;
#pragma unroll
    for (int j = 0; j < 16; ++j) s += v[j];
    s += __shfl_xor_sync(0xffffffffu, s, 1);
    s += __shfl_xor_sync(0xffffffffu, s, 2);
    if (ok) {
      uint32_t o[8];
#pragma unroll
      for (int mm = 0; mm < 4; ++mm) {
        if (BITS == 4) {
          o[2 * mm] = pack_bf16(v[mm], v[mm + 4]);
          o[2 * mm + 1] = pack_bf16(v[mm + 8], v[mm + 12]);
        } else {
          o[2 * mm] = pack_bf16(v[4 * mm], v[4 * mm + 2]);
          o[2 * mm + 1] = pack_bf16(v[4 * mm + 1], v[4 * mm + 3]);
        }
      }
      uint4* dst = reinterpret_cast<uint4*>(xs + (size_t)m * xrow + blk * 32);
      dst[0] = make_uint4(o[0], o[1], o[2], o[3]);
      dst[1] = make_uint4(o[4], o[5], o[6], o[7]);
      if ((blk & 3) == 0) sumx[((blk >> 2) * MT + (m >> 3)) * 8 + (m & 7)] = s;
    }
  }
}

}  // namespace pie
#include "pie_qmv_main.inl"
