// Probe (run on a B200): where does a 3-D TMA box land in shared memory under CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B?
// Prints, for rows 0..15 of chunk 0, the logical 16-byte piece index found at each physical 16-byte slot of the row.
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
#include <vector>
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
__global__ void probe(const __grid_constant__ CUtensorMap map, uint8_t* out, int rows) {
  extern __shared__ uint8_t raw[];
  uint8_t* sm = (uint8_t*)(((uintptr_t)raw + 1023) & ~(uintptr_t)1023);
  __shared__ uint64_t bar;
  uint32_t bar_s = (uint32_t)__cvta_generic_to_shared(&bar), dst = (uint32_t)__cvta_generic_to_shared(sm);
  if (threadIdx.x == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar_s));
    asm volatile("fence.mbarrier_init.release.cluster;");
    asm volatile("fence.proxy.async.shared::cta;");
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar_s), "r"(128 * rows * 2));
    asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
                 ::"r"(dst), "l"(&map), "r"(bar_s), "r"(0), "r"(0), "r"(0) : "memory");
  }
  __syncthreads();
  uint32_t ok = 0;
  while (!ok) asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(ok) : "r"(bar_s) : "memory");
  for (int i = threadIdx.x; i < 128 * rows * 2; i += blockDim.x) out[i] = sm[i];
}
int main() {
  void* p = nullptr; cudaDriverEntryPointQueryResult q;
  cudaFree(0);
  if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess || !p) { printf("no encode fn\n"); return 1; }
  EncodeTiledFn enc = (EncodeTiledFn)p;
  const int N = 64, K = 1024, RB = K / 2;  // 512 B per row = 4 chunks of 128 B
  std::vector<uint16_t> h(N * RB / 2);
  // every 16-bit word encodes (row, 16-byte piece index within the row, word within piece)
  for (int r = 0; r < N; ++r) for (int w = 0; w < RB / 2; ++w) h[r * RB / 2 + w] = (uint16_t)((r << 8) | ((w / 8) << 3) | (w % 8));
  uint8_t *d, *o; cudaMalloc(&d, N * RB); cudaMalloc(&o, 8192); cudaMemcpy(d, h.data(), N * RB, cudaMemcpyHostToDevice);
  const int modes[4] = {CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B_FLIP_8B};
  for (int rows = 16; rows >= 8; rows -= 8) for (int mi = 0; mi < 4; ++mi) {
    CUtensorMap m;
    const cuuint64_t dims[3] = {128, (cuuint64_t)N, (cuuint64_t)(K / 256)};
    const cuuint64_t strides[2] = {(cuuint64_t)RB, 128};
    const cuuint32_t box[3] = {128, (cuuint32_t)rows, 2}, es[3] = {1, 1, 1};
    CUresult r = enc(&m, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, d, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     (CUtensorMapSwizzle)modes[mi], CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    printf("rows %d mode %d encode -> %d\n", rows, modes[mi], (int)r);
    if (r != CUDA_SUCCESS) continue;
    cudaMemset(o, 0xff, 8192);
    probe<<<1, 128, 8192 + 1024>>>(m, o, rows);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("kernel: %s\n", cudaGetErrorString(e)); return 2; }
    std::vector<uint16_t> g(4096); cudaMemcpy(g.data(), o, 128 * rows * 2, cudaMemcpyDeviceToHost);
    for (int c = 0; c < 2; ++c) for (int r2 = 0; r2 < rows; ++r2) {
      printf("  chunk %d row %2d:", c, r2);
      for (int s = 0; s < 8; ++s) {  // physical 16-byte slot s of this 128-byte row: print (src row, src piece, first word idx)
        uint16_t v = g[(c * rows + r2) * 64 + s * 8];
        uint16_t v4 = g[(c * rows + r2) * 64 + s * 8 + 4];
        printf(" r%dp%d.%d/%d", v >> 8, (v >> 3) & 31, v & 7, v4 & 7);
      }
      printf("\n");
      if (r2 == 7 && rows == 16 && c == 1) {}
    }
  }
  return 0;
}
