// pie_decode_attn_static: the split-K flash-decode of pie_decode_attn with its scratch (split counters + per-split
// partial outputs) in a file-scope device array instead of a caller-owned workspace.  No counterpart in the reference;
// added for callers that cannot allocate device memory (code under CUDA graph
// capture, tight serving loops, harnesses that forbid cudaMalloc).  Nothing here allocates: no cudaMalloc / cudaFree, no
// allocating library routine.
//
// The array costs PIE_STATIC_SCRATCH_BYTES (4 MiB) of device memory on EVERY GPU that loads the library, until exit.
// All calls share it and it keeps the last call's contents: the counter region is cleared on the caller's stream before
// every launch (the partials are fully rewritten by the kernel before they are read), and two calls that overlap on the
// GPU (different streams, no ordering between them) RACE on it -- serialise them, or use pie_decode_attn with one
// workspace per stream.
#include <mutex>

#include "../../include/pie_b200.h"
#include "pie_internal.h"

namespace {

// In host code this name is only a stand-in for the symbol, not a device address: the kernel gets the address that
// cudaGetSymbolAddress returns for the current device (below).
static __device__ __align__(256) unsigned char g_static_scratch[PIE_STATIC_SCRATCH_BYTES];

constexpr int kMaxDevices = 64;
void* g_scratch_ptr[kMaxDevices];   // per device, filled on first use with that device current
std::mutex g_scratch_mu;

// device address of the array on the CURRENT device (looked up once per device)
int static_scratch(void** out) {
  int dev = -1;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kMaxDevices) return PIE_ECUDA;
  std::lock_guard<std::mutex> lk(g_scratch_mu);
  if (!g_scratch_ptr[dev]) {
    void* p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(&p, g_static_scratch);
    if (e != cudaSuccess || !p) return pie::status_of(e == cudaSuccess ? cudaErrorInvalidSymbol : e);
    g_scratch_ptr[dev] = p;
  }
  *out = g_scratch_ptr[dev];
  return PIE_OK;
}

}  // namespace

extern "C" int pie_decode_attn_static(void* out, const void* q, const void* kcache, const void* vcache,
                                      const pie_kv_layout* kv_host, const int32_t* positions, int B, int Hq, int Hkv,
                                      int D, float scale, int nsplit, void* stream) {
  if (B == 0) return PIE_OK;
  if (B < 0 || Hq <= 0 || Hkv <= 0 || D <= 0) return PIE_EINVAL_SHAPE;
  if (nsplit <= 0) nsplit = 1;
  // size check first, in 64-bit arithmetic: a request beyond the array is refused before anything is touched
  const unsigned long long cnt = (((unsigned long long)B * Hq * 4) + 255ull) & ~255ull;
  const unsigned long long need = cnt + (unsigned long long)B * Hq * nsplit * (D + 2) * sizeof(float);
  if (nsplit > 1 && need > (unsigned long long)PIE_STATIC_SCRATCH_BYTES) return PIE_ESTATIC_LIMIT;
  void* ws = nullptr;
  if (int s = static_scratch(&ws)) return s;
  if (nsplit > 1) {
    // the split counters must start at zero (a call that was cut short may have left them otherwise)
    const cudaError_t e = cudaMemsetAsync(ws, 0, (size_t)cnt, (cudaStream_t)stream);
    if (e != cudaSuccess) return pie::status_of(e);
  }
  return pie_decode_attn(out, q, kcache, vcache, kv_host, positions, B, Hq, Hkv, D, scale, nsplit, ws,
                         (size_t)PIE_STATIC_SCRATCH_BYTES, stream);
}
