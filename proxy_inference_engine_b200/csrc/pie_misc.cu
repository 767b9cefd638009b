// Small operators of the decode path: quantize / dequantize, RMSNorm, SwiGLU glue, residual add, arg-max.
// Each cites the reference operator it stands in for (MLX = src/pie_core/external/mlx, PIE = src/proxy_inference_engine).
#include <cstdlib>

#include "pie_internal.h"
#include "pie_ptx.cuh"

namespace pie {

__device__ __forceinline__ float ld_as_f32(const void* p, long long i, int dtype) {
  return dtype == 0 ? bf16lo((uint32_t) reinterpret_cast<const uint16_t*>(p)[i]) : reinterpret_cast<const float*>(p)[i];
}
__device__ __forceinline__ void st_from_f32(void* p, long long i, int dtype, float v) {
  if (dtype == 0) reinterpret_cast<uint16_t*>(p)[i] = f2bf(v);
  else reinterpret_cast<float*>(p)[i] = v;
}

// ---- a1 mx.quantize (MLX/mlx/backend/cpu/quantized.cpp:603-661): one thread per group, same op order ----
__global__ void quantize_kernel(const void* w, int dtype, uint32_t* wq, void* scales, void* biases,
                                long long n_groups, int gs, int bits) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_groups) return;
  const float n_bins = (float)((1 << bits) - 1);
  const int per = 32 / bits, ipg = gs / per;
  const long long w0 = i * gs;
  float w_min = INFINITY, w_max = -INFINITY;
  for (int j = 0; j < gs; ++j) {
    const float v = ld_as_f32(w, w0 + j, dtype);
    w_max = (w_max < v) ? v : w_max;
    w_min = (v < w_min) ? v : w_min;
  }
  const bool mask = fabsf(w_min) > fabsf(w_max);
  float scale = __fdiv_rn(w_max - w_min, n_bins);
  scale = (scale < 1e-7f) ? 1e-7f : scale;
  scale = mask ? scale : -scale;
  const float edge = mask ? w_min : w_max;
  const float q0 = rintf(__fdiv_rn(edge, scale));
  float bias = 0.f;
  if (q0 != 0.f) { scale = __fdiv_rn(edge, q0); bias = edge; }
  for (int j = 0; j < ipg; ++j) {
    uint32_t word = 0;
    for (int k = 0; k < per; ++k) {
      float e = ld_as_f32(w, w0 + (long long)j * per + k, dtype);
      e = rintf(__fdiv_rn(__fsub_rn(e, bias), scale));
      e = fminf(fmaxf(e, 0.f), n_bins);
      word |= ((uint32_t)e) << (k * bits);
    }
    wq[i * ipg + j] = word;
  }
  st_from_f32(scales, i, dtype, scale);
  st_from_f32(biases, i, dtype, bias);
}

cudaError_t launch_quantize(const void* w, int dtype, uint32_t* wq, void* s, void* b, long long rows, int K,
                            int gs, int bits, cudaStream_t st) {
  const long long ng = rows * (K / gs);
  if (ng == 0) return cudaSuccess;
  count_launch();
  quantize_kernel<<<(unsigned)((ng + 127) / 128), 128, 0, st>>>(w, dtype, wq, s, b, ng, gs, bits);
  return cudaGetLastError();
}

// ---- a4 mx.dequantize + row gather (MLX/mlx/fast.cpp:972-1017; quantized.py:100-107) ---------------------
// w_hat = bf16(bf16(q * scale) + bias); one thread per packed word.
__global__ void dequantize_kernel(uint16_t* out, const uint32_t* wq, const uint16_t* s, const uint16_t* b,
                                  const int32_t* ids, long long n_rows, long long table_rows, int K, int gs,
                                  int bits) {
  const int per = 32 / bits, wpr = K / per;
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= n_rows * wpr) return;
  const long long i = idx / wpr;
  const int wi = (int)(idx % wpr);
  long long r = ids ? (long long)ids[i] : i;
  r = r < 0 ? 0 : (r >= table_rows ? table_rows - 1 : r);
  const uint32_t w = wq[r * wpr + wi];
  const int grp = wi * per / gs;
  const float sc = bf16lo((uint32_t)s[r * (K / gs) + grp]), bi = bf16lo((uint32_t)b[r * (K / gs) + grp]);
  const uint32_t mask = (1u << bits) - 1u;
  uint16_t* o = out + i * K + (long long)wi * per;
  for (int e = 0; e < per; ++e) o[e] = f2bf(bfround((float)((w >> (e * bits)) & mask) * sc) + bi);
}

cudaError_t launch_dequantize(uint16_t* out, const uint32_t* wq, const uint16_t* s, const uint16_t* b,
                              const int32_t* ids, long long n_rows, long long table_rows, int K, int gs,
                              int bits, int*, cudaStream_t st) {
  const long long n = n_rows * (K / (32 / bits));
  if (n == 0) return cudaSuccess;
  count_launch();
  dequantize_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(out, wq, s, b, ids, n_rows, table_rows, K, gs, bits);
  return cudaGetLastError();
}

// ---- a6 mx.fast.rms_norm (MLX/mlx/fast.cpp:93-111; kernels/rms_norm.metal:13-81) --------------------------
// fp32 sum of squares; out = bf16( bf16(x * inv) * w ): the cast to T comes BEFORE the weight multiply.
__global__ void __launch_bounds__(256) rms_norm_kernel(uint16_t* out, const uint16_t* x, const uint16_t* w,
                                                       int dims, float eps) {
  __shared__ float red[8];
  const uint16_t* xr = x + (size_t)blockIdx.x * dims;
  uint16_t* o = out + (size_t)blockIdx.x * dims;
  float ss = 0.f;
  for (int i = threadIdx.x; i < dims; i += 256) { const float v = bf16lo((uint32_t)xr[i]); ss += v * v; }
  ss = warp_sum(ss);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = ss;
  __syncthreads();
  float tot = 0.f;
  for (int i = 0; i < 8; ++i) tot += red[i];
  const float inv = 1.0f / sqrtf(tot / (float)dims + eps);
  for (int i = threadIdx.x; i < dims; i += 256) {
    float v = bfround(bf16lo((uint32_t)xr[i]) * inv);
    if (w) v = v * bf16lo((uint32_t)w[i]);
    o[i] = f2bf(v);
  }
}

cudaError_t launch_rms_norm(uint16_t* out, const uint16_t* x, const uint16_t* w, long long rows, int dims,
                            float eps, cudaStream_t st) {
  if (rows == 0) return cudaSuccess;
  count_launch();
  rms_norm_kernel<<<(unsigned)rows, 256, 0, st>>>(out, x, w, dims, eps);
  return cudaGetLastError();
}

// ---- a11 glue (PIE/models/llama/language.py:126-127,150-154; nn.silu activations.py:125-132) ----------------
__global__ void silu_mul_kernel(uint16_t* out, const uint16_t* g, const uint16_t* u, long long n) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const float gt = bf16lo((uint32_t)g[i]);
    const float sg = bfround(1.f / (1.f + expf(-gt)));
    out[i] = f2bf(bfround(gt * sg) * bf16lo((uint32_t)u[i]));
  }
}
__global__ void add_kernel(uint16_t* out, const uint16_t* a, const uint16_t* b, long long n) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    out[i] = f2bf(bf16lo((uint32_t)a[i]) + bf16lo((uint32_t)b[i]));
}
static unsigned ew_grid(long long n) { return (unsigned)((n + 255) / 256 < 148 * 16 ? (n + 255) / 256 : 148 * 16); }
cudaError_t launch_silu_mul(uint16_t* out, const uint16_t* g, const uint16_t* u, long long n, cudaStream_t st) {
  if (n == 0) return cudaSuccess;
  count_launch();
  silu_mul_kernel<<<ew_grid(n), 256, 0, st>>>(out, g, u, n);
  return cudaGetLastError();
}
cudaError_t launch_add(uint16_t* out, const uint16_t* a, const uint16_t* b, long long n, cudaStream_t st) {
  if (n == 0) return cudaSuccess;
  count_launch();
  add_kernel<<<ew_grid(n), 256, 0, st>>>(out, a, b, n);
  return cudaGetLastError();
}

// ---- greedy sampler: first index of the row maximum (mx.argmax; PIE/samplers/__init__.py, temp == 0) --------
__global__ void __launch_bounds__(1024) argmax_kernel(int32_t* out, const uint16_t* logits, int vocab,
                                                      int32_t* pos_adv, int32_t* seq_out, int seq_ld) {
  __shared__ float sv[32];
  __shared__ int si[32];
  const uint16_t* row = logits + (size_t)blockIdx.x * vocab;
  float best = -INFINITY;
  int bi = 0x7fffffff;
  for (int i = threadIdx.x; i < vocab; i += 1024) {
    const float v = bf16lo((uint32_t)row[i]);
    if (v > best || (v == best && i < bi)) { best = v; bi = i; }
  }
  for (int o = 16; o > 0; o >>= 1) {
    const float ov = __shfl_xor_sync(0xffffffffu, best, o);
    const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
    if (ov > best || (ov == best && oi < bi)) { best = ov; bi = oi; }
  }
  if ((threadIdx.x & 31) == 0) { sv[threadIdx.x >> 5] = best; si[threadIdx.x >> 5] = bi; }
  __syncthreads();
  if (threadIdx.x < 32) {
    best = sv[threadIdx.x]; bi = si[threadIdx.x];
    for (int o = 16; o > 0; o >>= 1) {
      const float ov = __shfl_xor_sync(0xffffffffu, best, o);
      const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
      if (ov > best || (ov == best && oi < bi)) { best = ov; bi = oi; }
    }
    if (threadIdx.x == 0) {
      out[blockIdx.x] = bi == 0x7fffffff ? 0 : bi;
      if (pos_adv) {  // the token just sampled occupies the next cache slot
        const int np = pos_adv[blockIdx.x] + 1;
        pos_adv[blockIdx.x] = np;
        if (seq_out && np < seq_ld) seq_out[(size_t)blockIdx.x * seq_ld + np] = bi == 0x7fffffff ? 0 : bi;
      }
    }
  }
}
// The same over a thread-block cluster of 8 CTAs per row (one 1024-thread CTA reading 256 KB of logits is a 16 us chain
// of L2 round trips at the end of every decode step): each CTA scans an eighth of the row with 16-byte loads, CTAs 1..7
// push their candidate into CTA 0's shared memory (st.async counted by CTA 0's mbarrier), CTA 0 merges in index order
// -- first maximum, as above -- and does the bookkeeping.  No global scratch, so concurrent calls cannot interfere.
constexpr int kAmCluster = 8;
__global__ void __launch_bounds__(1024) argmax_cluster_kernel(int32_t* out, const uint16_t* logits, int vocab,
                                                              int32_t* pos_adv, int32_t* seq_out, int seq_ld) {
  __shared__ float sv[32];
  __shared__ int si[32];
  __shared__ __align__(8) uint32_t cand[kAmCluster][2];   // CTA 0: (value bits, index) of every part
  __shared__ uint64_t bar;
  const int part = blockIdx.x % kAmCluster, r = blockIdx.x / kAmCluster, tid = threadIdx.x;
  if (part == 0 && tid == 0) {
    mbar_init(&bar, 1);
    mbar_arrive_expect_tx(&bar, (kAmCluster - 1) * 8);
    fence_mbar_init();
  }
  asm volatile("barrier.cluster.arrive.relaxed.aligned;" ::: "memory");
  const uint16_t* row = logits + (size_t)r * vocab;
  const int per = ((vocab + kAmCluster - 1) / kAmCluster + 7) & ~7;   // multiple of 8 elements: 16-byte pieces
  const int i0 = part * per, i1 = min(vocab, i0 + per);
  float best = -INFINITY;
  int bi = 0x7fffffff;
  auto take = [&](float v, int i) { if (v > best || (v == best && i < bi)) { best = v; bi = i; } };
  if ((reinterpret_cast<uintptr_t>(row + i0) & 15) == 0) {
    const int nv = i1 > i0 ? (i1 - i0) / 8 : 0;
    for (int j = tid; j < nv; j += 1024) {
      const uint4 q = ldcg128(row + i0 + 8 * j);
      const uint32_t w[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
      for (int e = 0; e < 4; ++e) { take(bf16lo(w[e]), i0 + 8 * j + 2 * e); take(bf16hi(w[e]), i0 + 8 * j + 2 * e + 1); }
    }
    for (int i = i0 + 8 * nv + tid; i < i1; i += 1024) take(bf16lo((uint32_t)row[i]), i);
  } else {
    for (int i = i0 + tid; i < i1; i += 1024) take(bf16lo((uint32_t)row[i]), i);
  }
  auto merge = [&](float ov, int oi) { if (ov > best || (ov == best && oi < bi)) { best = ov; bi = oi; } };
  for (int o = 16; o > 0; o >>= 1) merge(__shfl_xor_sync(0xffffffffu, best, o), __shfl_xor_sync(0xffffffffu, bi, o));
  if ((tid & 31) == 0) { sv[tid >> 5] = best; si[tid >> 5] = bi; }
  __syncthreads();
  asm volatile("barrier.cluster.wait.aligned;" ::: "memory");   // CTA 0's mbarrier is set up
  if (tid < 32) {
    best = sv[tid]; bi = si[tid];
    for (int o = 16; o > 0; o >>= 1) merge(__shfl_xor_sync(0xffffffffu, best, o), __shfl_xor_sync(0xffffffffu, bi, o));
    if (tid == 0) {
      if (part != 0) {
        uint32_t ra, rb;
        asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(ra) : "r"(smem_u32(&cand[part][0])), "r"(0));
        asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(rb) : "r"(smem_u32(&bar)), "r"(0));
        asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.v2.b32 [%0], {%1, %2}, [%3];" ::"r"(ra),
                     "r"(__float_as_uint(best)), "r"((uint32_t)bi), "r"(rb) : "memory");
      } else {
        mbar_wait(&bar, 0);
        for (int p2 = 1; p2 < kAmCluster; ++p2) merge(__uint_as_float(cand[p2][0]), (int)cand[p2][1]);
        const int tok = bi == 0x7fffffff ? 0 : bi;
        out[r] = tok;
        if (pos_adv) {  // the token just sampled occupies the next cache slot
          const int np = pos_adv[r] + 1;
          pos_adv[r] = np;
          if (seq_out && np < seq_ld) seq_out[(size_t)r * seq_ld + np] = tok;
        }
      }
    }
  }
}

cudaError_t launch_argmax(int32_t* out, const uint16_t* logits, int rows, int vocab, int32_t* pos_adv,
                          int32_t* seq_out, int seq_ld, cudaStream_t st) {
  if (rows == 0) return cudaSuccess;
  count_launch();
  static const bool no_cluster = getenv("PIE_ARGMAX_CLUSTER") && atoi(getenv("PIE_ARGMAX_CLUSTER")) == 0;
  if (!no_cluster && vocab >= 8192 && rows <= 64) {   // few long rows: the end of a decode step
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)(rows * kAmCluster), 1, 1);
    cfg.blockDim = dim3(1024, 1, 1);
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = kAmCluster; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    return cudaLaunchKernelEx(&cfg, argmax_cluster_kernel, out, logits, vocab, pos_adv, seq_out, seq_ld);
  }
  argmax_kernel<<<rows, 1024, 0, st>>>(out, logits, vocab, pos_adv, seq_out, seq_ld);
  return cudaGetLastError();
}

}  // namespace pie
