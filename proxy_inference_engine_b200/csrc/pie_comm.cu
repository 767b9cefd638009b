// One-shot tensor-parallel exchanges over NVLink peer memory (no NCCL on the decode path).
//
// The exchange step of MLX's sharded linears -- `all_sum` after the row-split matrix
// (QuantizedShardedToAllLinear, MLX/python/mlx/nn/layers/distributed.py:459-566, all_sum at :566) and the gather of the
// vocabulary-split logits -- at decode sizes is 8-32 KB per GPU: pure latency.  Instead of a ring / tree collective,
// every rank PUSHES its partial row straight into a slot of every peer's buffer (peer-mapped symmetric memory; NVSwitch
// gives each pair full bandwidth), raises a flag there, waits for its peers' flags in its own memory and reduces the
// `world` slots locally in rank order: one kernel, one NVLink traversal, deterministic, identical bits on all ranks.
//
// Symmetric buffer of a rank (same layout everywhere; peers see it through the pointers of `pie_tp_peers`):
//   data   [2 sets][world][max_elems] bf16      slot r of set s = rank r's contribution to exchange number e, s = e & 1
//   flags  [2 sets][world][kTpMaxCtas] u32      flag (s, r, c) = e once rank r's chunk c of exchange e has landed
// Two sets: a rank can run at most one exchange ahead of a peer (it needs the peer's flag of exchange e to finish e), so
// when it writes set s again (exchange e + 2) every peer has finished reading exchange e.  The exchange number lives in
// device memory and is advanced by the kernel itself: a captured CUDA graph replays without host involvement.
#include <cstring>

#include "pie_internal.h"
#include "pie_ptx.cuh"

namespace pie {

constexpr int kTpChunk = 4096;  // bf16 elements per CTA (512 threads x 16 B)

struct TpArgs {
  uint8_t* peer[kTpMaxWorld];   // base of every rank's symmetric buffer, as mapped in THIS process
  int rank, world, max_elems;
  unsigned int* epoch;          // local: [0] exchange number (starts at 1), [1] CTA departure counter
  const uint16_t* src;          // local contribution
  const uint16_t* residual;     // all-reduce: optional, added after the sum (language.py:150-154)
  uint16_t* out;                // all-reduce: result; all-gather: unused (peers write the destination directly)
  int n;                        // elements of this exchange (all-reduce: rows * width)
  // all-gather of row slices: src is [rows, width]; destination [rows, world * width] lives at dst_off bytes in the
  // symmetric buffer of every rank
  int rows, width;
  long long dst_off;
};

__device__ __forceinline__ void tp_depart(const TpArgs& a) {
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    if (atomicAdd(&a.epoch[1], 1u) == gridDim.x - 1) {  // last CTA out: every CTA has read the exchange number
      a.epoch[1] = 0;
      __threadfence();
      a.epoch[0] += 1;
    }
  }
}

// out = bf16( residual + bf16( sum_r partial_r ) ), fp32 sum in rank order
__global__ void __launch_bounds__(512) tp_allreduce_kernel(const __grid_constant__ TpArgs a) {
  pdl_launch_dependents();   // the next GEMV may start prefetching its weights now; it waits for this grid before x
  pdl_wait();                // the producer of `src` (and of the exchange number) has completed
  const unsigned int e = *reinterpret_cast<volatile unsigned int*>(a.epoch);
  const int set = e & 1, c = blockIdx.x, tid = threadIdx.x;
  const int i0 = c * kTpChunk + tid * 8;   // 8 bf16 = 16 B per thread
  const size_t foff = tp_flags_off(a.world, a.max_elems);
  const bool live = i0 < a.n;              // n is a multiple of 8
  if (live) {
    const uint4 v = *reinterpret_cast<const uint4*>(a.src + i0);
    for (int p = 0; p < a.world; ++p) {
      uint16_t* slot = reinterpret_cast<uint16_t*>(a.peer[p]) + ((size_t)set * a.world + a.rank) * a.max_elems;
      *reinterpret_cast<uint4*>(slot + i0) = v;
    }
  }
  // the CTA barrier orders every thread's pushes before the signalling threads' release (cumulative at system scope)
  __syncthreads();
  if (tid < a.world)
    st_release_sys(reinterpret_cast<unsigned int*>(a.peer[tid] + foff) + ((size_t)set * a.world + a.rank) * kTpMaxCtas + c, e);
  if (tid < a.world)
    wait_flag(reinterpret_cast<unsigned int*>(a.peer[a.rank] + foff) + ((size_t)set * a.world + tid) * kTpMaxCtas + c, e);
  __syncthreads();
  if (live) {
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    const uint16_t* mine = reinterpret_cast<const uint16_t*>(a.peer[a.rank]) + (size_t)set * a.world * a.max_elems;
    for (int r = 0; r < a.world; ++r) {
      const uint4 v = ldcg128(mine + (size_t)r * a.max_elems + i0);
      const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
      for (int j = 0; j < 4; ++j) { acc[2 * j] += bf16lo(w[j]); acc[2 * j + 1] += bf16hi(w[j]); }
    }
    uint32_t res[4] = {0, 0, 0, 0};
    if (a.residual) {
      const uint4 v = *reinterpret_cast<const uint4*>(a.residual + i0);
      res[0] = v.x; res[1] = v.y; res[2] = v.z; res[3] = v.w;
    }
    uint32_t o[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      float lo = bfround(acc[2 * j]), hi = bfround(acc[2 * j + 1]);
      if (a.residual) { lo += bf16lo(res[j]); hi += bf16hi(res[j]); }
      o[j] = pack_bf16(lo, hi);
    }
    *reinterpret_cast<uint4*>(a.out + i0) = make_uint4(o[0], o[1], o[2], o[3]);
  }
  tp_depart(a);
}

// every rank's [rows, width] slice lands at columns [rank * width, (rank + 1) * width) of every rank's destination
__global__ void __launch_bounds__(512) tp_allgather_kernel(const __grid_constant__ TpArgs a) {
  pdl_launch_dependents();
  pdl_wait();
  const unsigned int e = *reinterpret_cast<volatile unsigned int*>(a.epoch);
  const int set = e & 1, c = blockIdx.x, tid = threadIdx.x;
  const size_t foff = tp_flags_off(a.world, a.max_elems);
  const int per_row = a.width / 8;  // 16-byte pieces per row (width is a multiple of 8)
  for (int i = c * 512 + tid; i < a.rows * per_row; i += gridDim.x * 512) {
    const int row = i / per_row, col = (i - row * per_row) * 8;
    const uint4 v = *reinterpret_cast<const uint4*>(a.src + (size_t)row * a.width + col);
    for (int p = 0; p < a.world; ++p) {
      uint16_t* dst = reinterpret_cast<uint16_t*>(a.peer[p] + a.dst_off) + (size_t)row * a.world * a.width +
                      (size_t)a.rank * a.width + col;
      *reinterpret_cast<uint4*>(dst) = v;
    }
  }
  __syncthreads();
  if (tid < a.world)
    st_release_sys(reinterpret_cast<unsigned int*>(a.peer[tid] + foff) + ((size_t)set * a.world + a.rank) * kTpMaxCtas + c, e);
  if (tid < a.world)
    wait_flag(reinterpret_cast<unsigned int*>(a.peer[a.rank] + foff) + ((size_t)set * a.world + tid) * kTpMaxCtas + c, e);
  tp_depart(a);
}

template <class K>
static cudaError_t tp_launch(K kernel, int grid, const TpArgs& a, cudaStream_t st) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)grid, 1, 1);
  cfg.blockDim = dim3(512, 1, 1);
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  count_launch();
  return cudaLaunchKernelEx(&cfg, kernel, a);
}

}  // namespace pie

using namespace pie;

extern "C" {

size_t pie_tp_buffer_bytes(int world, int max_elems, size_t gather_bytes) {
  if (world < 1 || world > kTpMaxWorld || max_elems <= 0) return 0;
  return ((tp_flags_off(world, max_elems) + (size_t)2 * world * kTpMaxCtas * 4 + 255) & ~(size_t)255) + gather_bytes;
}
size_t pie_tp_gather_offset(int world, int max_elems) {
  if (world < 1 || world > kTpMaxWorld || max_elems <= 0) return 0;
  return (tp_flags_off(world, max_elems) + (size_t)2 * world * kTpMaxCtas * 4 + 255) & ~(size_t)255;
}

static int tp_fill(TpArgs& a, const pie_tp_peers* peers, const void* src, int n) {
  if (!peers || !src || !peers->epoch) return PIE_ENULL;
  if (peers->world < 1 || peers->world > kTpMaxWorld || peers->rank < 0 || peers->rank >= peers->world) return PIE_EINVAL_SHAPE;
  if (n <= 0 || n % 8 != 0) return PIE_EINVAL_SHAPE;
  memset(&a, 0, sizeof(a));
  for (int i = 0; i < peers->world; ++i) {
    if (!peers->base[i]) return PIE_ENULL;
    a.peer[i] = (uint8_t*)peers->base[i];
  }
  a.rank = peers->rank; a.world = peers->world; a.max_elems = peers->max_elems;
  a.epoch = (unsigned int*)peers->epoch;
  a.src = (const uint16_t*)src; a.n = n;
  return PIE_OK;
}

int pie_tp_allreduce(void* out, const void* partial, const void* residual, int rows, int width,
                     const pie_tp_peers* peers_host, void* stream) {
  TpArgs a;
  const long long n = (long long)rows * width;
  if (n > (long long)kTpMaxCtas * kTpChunk) return PIE_EUNSUPPORTED;
  const int st = tp_fill(a, peers_host, partial, (int)n);
  if (st != PIE_OK) return st;
  if (!out) return PIE_ENULL;
  if (n > peers_host->max_elems) return PIE_EWORKSPACE;
  a.out = (uint16_t*)out; a.residual = (const uint16_t*)residual;
  const int grid = (int)((n + kTpChunk - 1) / kTpChunk);
  return status_of(tp_launch(tp_allreduce_kernel, grid, a, (cudaStream_t)stream));
}

int pie_tp_allgather(const void* slice, int rows, int width, const pie_tp_peers* peers_host, void* stream) {
  TpArgs a;
  const int st = tp_fill(a, peers_host, slice, rows * width);
  if (st != PIE_OK) return st;
  if (width % 8 != 0) return PIE_EINVAL_SHAPE;
  a.rows = rows; a.width = width; a.dst_off = (long long)pie_tp_gather_offset(peers_host->world, peers_host->max_elems);
  long long pieces = (long long)rows * width / 8;
  int grid = (int)((pieces + 511) / 512);
  if (grid > kTpMaxCtas) grid = kTpMaxCtas;
  return status_of(tp_launch(tp_allgather_kernel, grid, a, (cudaStream_t)stream));
}

}  // extern "C"
