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
#include "pie_qmv_body.cuh"

namespace pie {

size_t qmv_smem_bytes(int M, int kslice, int bits, int warps, int mt) {
  size_t xs = ((size_t)M * qmv_xrow_bytes(kslice) + 127) & ~(size_t)127;
  size_t sumx = (size_t)(kslice / kGroup) * 8 * mt * sizeof(float);
  size_t red = (size_t)warps * kQmvSlots * 128 * mt * sizeof(float);  // exchange slots
  size_t stage = bits == 4 ? QmvGeom<4>::STAGE_BYTES : QmvGeom<8>::STAGE_BYTES;
  return kQmvBarBytes + (size_t)warps * kQmvStages * stage + xs + sumx + red;
}

template <int BITS, int MT, bool ONECOL>
__global__ void __launch_bounds__(MT == 4 ? 256 : 512, 1) qmv_kernel(const __grid_constant__ QmvArgs a) {
  extern __shared__ __align__(128) uint8_t smem[];
  qmv_body<BITS, MT, ONECOL, false, QMV_FULL>(a, smem, false, [] { pdl_launch_dependents(); pdl_wait(); });
}

// Experiment (PIE_QMV_SMEMARGS=1, tools/bench_qmv.py): the same matrix with the descriptor read from shared memory
// and the producer words parked in shared memory, i.e. the GEMV exactly as the one-launch decode step runs it.
template <int BITS>
__global__ void __launch_bounds__(512, 1) qmv_kernel_smemargs(const __grid_constant__ QmvArgs a) {
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ QmvArgs sa;
  for (int i = threadIdx.x; i < (int)(sizeof(QmvArgs) / 4); i += blockDim.x)
    reinterpret_cast<uint32_t*>(&sa)[i] = reinterpret_cast<const uint32_t*>(&a)[i];
  __syncthreads();
  qmv_body<BITS, 1, true, true, QMV_FULL>(sa, smem, false, [] { pdl_launch_dependents(); pdl_wait(); });
}

}  // namespace pie
#include "pie_qmv_launch.inl"
