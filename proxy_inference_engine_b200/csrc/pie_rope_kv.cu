// RoPE (mx.fast.rope, MLX/mlx/fast.cpp:353-511; kernels/rope.metal:6-47) and the KV-cache writes of
// ReusableKVCache.update_and_fetch (PIE/cache/kv_cache/reusable.py:96-142).
// theta = (pos + offset) * scale * (1 / freqs[i]) in fp32, range-reduced sine / cosine (rope_sincos, pie_ptx.cuh), one bf16 rounding at the store.
#include "pie_internal.h"
#include "pie_ptx.cuh"

namespace pie {

__device__ __forceinline__ size_t kv_addr(const KvLayout& kv, int b, int h, int t) {
  if (kv.page_table) {
    const int page = kv.page_table[(size_t)b * kv.pt_stride + t / kv.page_size];
    return ((size_t)page * kv.page_size + t % kv.page_size) * kv.tok_stride + (size_t)h * kv.head_stride;
  }
  return (size_t)b * kv.batch_stride + (size_t)h * kv.head_stride + (size_t)t * kv.tok_stride;
}

__device__ __forceinline__ float rope_inv_freq(const float* freqs, float base, int i, int half) {
  if (freqs) return 1.0f / freqs[i];
  return expf((float)(-i) * (logf(base) / (float)half));
}

// general form: x, out [n_batch, L, D]; rotates the first `dims` features
__global__ void rope_kernel(uint16_t* out, const uint16_t* x, long long rows, int L, int D, int dims,
                            int traditional, float base, float scale, int offset, const float* freqs) {
  const int half = dims / 2;
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const int per_row = D - half;  // `half` rotating pairs + (D - dims) pass-through features
  if (idx >= rows * per_row) return;
  const long long row = idx / per_row;
  const int i = (int)(idx % per_row);
  const uint16_t* xr = x + row * D;
  uint16_t* o = out + row * D;
  if (i >= half) { o[dims + (i - half)] = xr[dims + (i - half)]; return; }
  const int p = (int)(row % L);
  const float theta = ((float)p + (float)offset) * scale * rope_inv_freq(freqs, base, i, half);
  float s, c;
  rope_sincos(theta, &s, &c);
  const int i1 = traditional ? 2 * i : i, i2 = traditional ? 2 * i + 1 : i + half;
  const float x1 = bf16lo((uint32_t)xr[i1]), x2 = bf16lo((uint32_t)xr[i2]);
  o[i1] = f2bf(__fsub_rn(__fmul_rn(x1, c), __fmul_rn(x2, s)));
  o[i2] = f2bf(__fadd_rn(__fmul_rn(x1, s), __fmul_rn(x2, c)));
}

cudaError_t launch_rope(uint16_t* out, const uint16_t* x, long long n_batch, int L, int D, int dims,
                        int traditional, float base, float scale, int offset, const float* freqs,
                        cudaStream_t st) {
  const long long n = n_batch * L * (D - dims / 2);
  if (n == 0) return cudaSuccess;
  count_launch();
  rope_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(out, x, n_batch * L, L, D, dims, traditional, base,
                                                            scale, offset, freqs);
  return cudaGetLastError();
}

// decode (L == 1): one CTA per (head, batch row) of the fused q/k/v projection output
// (Attention.__call__, PIE/models/llama/language.py:84-98): q heads are rotated into q_out, k heads are
// rotated into the cache at positions[b], v heads are copied into the cache.
__global__ void rope_kv_append_kernel(uint16_t* q_out, const uint16_t* qkv, uint16_t* kc, uint16_t* vc,
                                      const KvLayout kv, const int32_t* pos, const float* freqs, int Hq,
                                      int Hkv, int D) {
  pdl_launch_dependents();
  pdl_wait();
  const int h = blockIdx.x, b = blockIdx.y, i = threadIdx.x, half = D / 2;
  const int p = pos[b];
  const uint16_t* src = qkv + ((size_t)b * (Hq + 2 * Hkv) + h) * D;
  if (h >= Hq + Hkv) {  // v: plain copy
    uint16_t* dst = vc + kv_addr(kv, b, h - Hq - Hkv, p);
    dst[i] = src[i];
    dst[i + half] = src[i + half];
    return;
  }
  const float theta = (float)p * (1.0f / freqs[i]);
  float s, c;
  rope_sincos(theta, &s, &c);
  const float x1 = bf16lo((uint32_t)src[i]), x2 = bf16lo((uint32_t)src[i + half]);
  const uint16_t o1 = f2bf(__fsub_rn(__fmul_rn(x1, c), __fmul_rn(x2, s)));
  const uint16_t o2 = f2bf(__fadd_rn(__fmul_rn(x1, s), __fmul_rn(x2, c)));
  uint16_t* dst = h < Hq ? q_out + ((size_t)b * Hq + h) * D : kc + kv_addr(kv, b, h - Hq, p);
  dst[i] = o1;
  dst[i + half] = o2;
}

cudaError_t launch_rope_kv_append(uint16_t* q_out, const uint16_t* qkv, uint16_t* kc, uint16_t* vc,
                                  const KvLayout& kv, const int32_t* pos, const float* freqs, int B, int Hq,
                                  int Hkv, int D, cudaStream_t st) {
  if (B == 0) return cudaSuccess;
  count_launch();
  return launch_pdl(rope_kv_append_kernel, dim3(Hq + 2 * Hkv, B), dim3(D / 2), 0, st, q_out, qkv, kc, vc, kv, pos, freqs, Hq,
                    Hkv, D);
}

// prefill: k, v [B, Hkv, L, D] -> cache rows offset .. offset + L - 1
__global__ void kv_write_kernel(uint16_t* kc, uint16_t* vc, const KvLayout kv, const uint16_t* k, const uint16_t* v,
                                int Hkv, int L, int D, int offset) {
  const int l = blockIdx.x, h = blockIdx.y, b = blockIdx.z;
  const size_t src = (((size_t)b * Hkv + h) * L + l) * D;
  const size_t dst = kv_addr(kv, b, h, offset + l);
  for (int i = threadIdx.x; i < D; i += blockDim.x) { kc[dst + i] = k[src + i]; vc[dst + i] = v[src + i]; }
}

cudaError_t launch_kv_write(uint16_t* kc, uint16_t* vc, const KvLayout& kv, const uint16_t* k,
                            const uint16_t* v, int B, int Hkv, int L, int D, int offset, cudaStream_t st) {
  if (B == 0 || L == 0) return cudaSuccess;
  count_launch();
  kv_write_kernel<<<dim3(L, Hkv, B), 64, 0, st>>>(kc, vc, kv, k, v, Hkv, L, D, offset);
  return cudaGetLastError();
}

}  // namespace pie
