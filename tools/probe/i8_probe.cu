// Probe (run on a B200): tcgen05.mma kind::i8 with A = uint8 [128 x K] (raw q8 weights), B = int8 [64 x K], both K-major
// in 128-byte-swizzled shared memory, D = int32 [128 x 64] in TMEM.  Prints the number of mismatches against the host.
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <vector>
constexpr int KK = 128;   // one 128-byte swizzle atom of K
__device__ __forceinline__ uint32_t s32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint64_t desc_sw128(uint32_t a) {
  uint64_t d = 0;
  d |= (uint64_t)((a & 0x3FFFF) >> 4);
  d |= (uint64_t)1 << 16;
  d |= (uint64_t)(1024 >> 4) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}
__global__ void __launch_bounds__(128) probe(const uint8_t* A, const int8_t* B, int32_t* D) {
  extern __shared__ uint8_t raw[];
  uint8_t* sm = (uint8_t*)(((uintptr_t)raw + 1023) & ~(uintptr_t)1023);
  uint8_t* sA = sm;                 // 128 rows x 128 B
  uint8_t* sB = sm + 128 * 128;     // 64 rows x 128 B
  __shared__ uint64_t bar;
  __shared__ uint32_t slot;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  // swizzled K-major tiles: row r, 16-byte chunk c -> (r / 8) * 1024 + (r % 8) * 128 + ((c ^ (r % 8)) * 16)
  for (int i = tid; i < 128 * 8; i += 128) {
    const int r = i / 8, c = i % 8;
    *(uint4*)(sA + (r / 8) * 1024 + (r % 8) * 128 + ((c ^ (r % 8)) * 16)) = *(const uint4*)(A + r * KK + c * 16);
  }
  for (int i = tid; i < 64 * 8; i += 128) {
    const int r = i / 8, c = i % 8;
    *(uint4*)(sB + (r / 8) * 1024 + (r % 8) * 128 + ((c ^ (r % 8)) * 16)) = *(const uint4*)(B + r * KK + c * 16);
  }
  if (tid == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(s32(&bar)));
    asm volatile("fence.mbarrier_init.release.cluster;");
  }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 64;" ::"r"(s32(&slot)) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic-proxy writes of the tiles -> async proxy (MMA)
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = slot;
  // idesc: D = s32 (2 << 4), A = u8 (0 << 7), B = s8 (1 << 10), K-major both, N = 64, M = 128
  const uint32_t idesc = (2u << 4) | (0u << 7) | (1u << 10) | ((64u >> 3) << 17) | ((128u >> 4) << 24);
  if (tid == 0) {
    const uint64_t da = desc_sw128(s32(sA)), db = desc_sw128(s32(sB));
    for (int k = 0; k < KK / 32; ++k) {   // K = 32 bytes per instruction: +2 in 16-byte units
      const uint32_t acc = k > 0;
      asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                   "tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n\t}\n" ::"r"(tmem), "l"(da + 2 * k), "l"(db + 2 * k),
                   "r"(idesc), "r"(acc) : "memory");
    }
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(s32(&bar)) : "memory");
  }
  uint32_t ok = 0;
  while (!ok) asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(ok) : "r"(s32(&bar)) : "memory");
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  uint32_t v[32];
  for (int half = 0; half < 2; ++half) {
    const uint32_t taddr = tmem + half * 32 + ((uint32_t)(warp * 32) << 16);
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
        "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]),
        "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]),
        "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31]) : "r"(taddr) : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
    for (int j = 0; j < 32; ++j) D[(warp * 32 + lane) * 64 + half * 32 + j] = (int32_t)v[j];
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 64;" ::"r"(tmem) : "memory");
}
int main() {
  std::vector<uint8_t> A(128 * KK); std::vector<int8_t> B(64 * KK);
  srand(1);
  for (auto& x : A) x = (uint8_t)(rand() & 255);
  for (auto& x : B) x = (int8_t)((rand() & 255) - 128);
  uint8_t* dA; int8_t* dB; int32_t* dD;
  cudaMalloc(&dA, A.size()); cudaMalloc(&dB, B.size()); cudaMalloc(&dD, 128 * 64 * 4);
  cudaMemcpy(dA, A.data(), A.size(), cudaMemcpyHostToDevice); cudaMemcpy(dB, B.data(), B.size(), cudaMemcpyHostToDevice);
  cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 32768);
  probe<<<1, 128, 32768>>>(dA, dB, dD);
  cudaError_t e = cudaDeviceSynchronize();
  printf("cuda: %s\n", cudaGetErrorString(e));
  std::vector<int32_t> D(128 * 64);
  cudaMemcpy(D.data(), dD, D.size() * 4, cudaMemcpyDeviceToHost);
  int bad = 0;
  for (int m = 0; m < 128; ++m) for (int n = 0; n < 64; ++n) {
    int32_t r = 0;
    for (int k = 0; k < KK; ++k) r += (int32_t)A[m * KK + k] * (int32_t)B[n * KK + k];
    if (r != D[m * 64 + n]) { if (bad < 5) printf("mismatch m %d n %d ref %d got %d\n", m, n, r, D[m * 64 + n]); ++bad; }
  }
  printf("i8 probe: %d mismatches of %d\n", bad, 128 * 64);
  return bad != 0;
}
