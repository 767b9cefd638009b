// Launcher of the decode GEMV and the slow generic kernel for shapes off the fast path.
#include <cstdlib>
namespace pie {

int qmv_mt(int M) { return M <= 8 ? 1 : (M <= 16 ? 2 : 4); }

int num_sms() {
  static int n = 0;
  if (n == 0) {
    int dev = 0, v = 0;
    if (cudaGetDevice(&dev) == cudaSuccess &&
        cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess && v > 0)
      n = v;
    else
      return 148;  // no device visible (host-side sizing calls): the B200 count
  }
  return n;
}

template <int BITS, int MT, bool ONECOL>
static cudaError_t launch_qmv_t(const QmvArgs& a, int warps, int grid, bool pdl, cudaStream_t st) {
  const size_t smem = qmv_smem_bytes(a.M, a.kslice, BITS, warps, MT);
  static unsigned long long configured = 0;  // opt in to the full 227 KB of dynamic shared memory, once per device
  if (first_use_on_device(configured)) {
    cudaError_t e = cudaFuncSetAttribute(qmv_kernel<BITS, MT, ONECOL>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         kQmvMaxSmem);
    if (e != cudaSuccess) return e;
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)grid, 1, 1);
  cfg.blockDim = dim3((unsigned)warps * 32, 1, 1);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl ? 1 : 0;
  count_launch();
  return cudaLaunchKernelEx(&cfg, qmv_kernel<BITS, MT, ONECOL>, a);
}

static cudaError_t launch_qmv_smemargs(const QmvArgs& a, int warps, int grid, bool pdl, cudaStream_t st) {
  const size_t smem = qmv_smem_bytes(a.M, a.kslice, 4, warps, 1) + (size_t)warps * 32 * kQmvProdWords * 4;
  static unsigned long long configured = 0;
  if (first_use_on_device(configured)) {
    cudaError_t e = cudaFuncSetAttribute(qmv_kernel_smemargs<4>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         kQmvMaxSmem - 1024);
    if (e != cudaSuccess) return e;
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)grid, 1, 1);
  cfg.blockDim = dim3((unsigned)warps * 32, 1, 1);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl ? 1 : 0;
  count_launch();
  return cudaLaunchKernelEx(&cfg, qmv_kernel_smemargs<4>, a);
}

cudaError_t launch_qmv(const QmvArgs& a, int bits, int warps, int grid, bool pdl, cudaStream_t st) {
  const int mt = qmv_mt(a.M);
  static const bool smemargs = getenv("PIE_QMV_SMEMARGS") && atoi(getenv("PIE_QMV_SMEMARGS")) != 0;
  if (smemargs && bits == 4 && a.M == 1 && warps == 16) return launch_qmv_smemargs(a, warps, grid, pdl, st);
  if (bits == 4) {
    if (a.M == 1) return launch_qmv_t<4, 1, true>(a, warps, grid, pdl, st);
    if (mt == 1) return launch_qmv_t<4, 1, false>(a, warps, grid, pdl, st);
    if (mt == 2) return launch_qmv_t<4, 2, false>(a, warps, grid, pdl, st);
    return launch_qmv_t<4, 4, false>(a, warps, grid, pdl, st);
  }
  if (a.M == 1) return launch_qmv_t<8, 1, true>(a, warps, grid, pdl, st);
  if (mt == 1) return launch_qmv_t<8, 1, false>(a, warps, grid, pdl, st);
  if (mt == 2) return launch_qmv_t<8, 2, false>(a, warps, grid, pdl, st);
  return launch_qmv_t<8, 4, false>(a, warps, grid, pdl, st);
}

// One warp per output row, any M; bits in {2,4,8}, group_size in {32,64,128}.  fp32, k-ordered per lane.
__global__ void qmv_generic_kernel(uint16_t* y, const uint16_t* x, const uint32_t* wq, const uint16_t* s,
                                   const uint16_t* b, int M, int N, int K, int gs, int bits) {
  const int n = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (n >= N) return;
  const int per = 32 / bits;
  const uint32_t mask = (1u << bits) - 1u;
  const uint32_t* wr = wq + (size_t)n * (K / per);
  const int gpr = K / gs;
  for (int m = 0; m < M; ++m) {
    float acc = 0.f;
    for (int wi = lane; wi < K / per; wi += 32) {
      const uint32_t w = wr[wi];
      const int kb = wi * per;
      const int grp = kb / gs;  // a word never straddles a group (gs % per == 0)
      const float sc = bf16lo((uint32_t)s[(size_t)n * gpr + grp]), bi = bf16lo((uint32_t)b[(size_t)n * gpr + grp]);
      for (int e = 0; e < per; ++e) {
        const float q = (float)((w >> (e * bits)) & mask);
        acc += bf16lo((uint32_t)x[(size_t)m * K + kb + e]) * (sc * q + bi);
      }
    }
    acc = warp_sum(acc);
    if (lane == 0) y[(size_t)m * N + n] = f2bf(acc);
  }
}

cudaError_t launch_qmv_generic(uint16_t* y, const uint16_t* x, const uint32_t* wq, const uint16_t* s,
                               const uint16_t* b, int M, int N, int K, int gs, int bits, cudaStream_t st) {
  const int wpb = 8;
  count_launch();
  qmv_generic_kernel<<<(N + wpb - 1) / wpb, wpb * 32, 0, st>>>(y, x, wq, s, b, M, N, K, gs, bits);
  return cudaGetLastError();
}

}  // namespace pie
