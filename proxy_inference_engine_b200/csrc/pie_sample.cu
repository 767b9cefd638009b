// Device-side sampler of the decode loop: log-probabilities + temperature / top-k / top-p / min-p /
// categorical sampling in ONE launch per step, token ids never leave the GPU.
//
// Replaces, per step of generate_step (PIE/engine/inference_engine.py:268-271):
//     logprobs = f32(logits) - logsumexp(f32(logits));  next = sampler(logprobs)
// with the samplers of PIE/samplers/__init__.py:11-45 (make_sampler):
//     temp == 0          -> argmax                                  (pie_argmax, pie_misc.cu)
//     0 < top_p < 1      -> top_p_sampling   (samplers/top_p.py)     keep a token iff the mass of strictly larger ones < top_p
//     min_p != 0         -> min_p_sampling   (samplers/min_p.py)     keep x >= max(x) + log(min_p), and the top min_tokens_to_keep
//     top_k > 0          -> top_k_sampling   (samplers/top_k.py)     keep the k largest
//     else               -> categorical_sampling (samplers/categorical.py)
// where x = logprobs / temp and the draw is mx.random.categorical = arg-max of x + Gumbel noise over the kept tokens.
// MLX's own random stream (threefry keyed by mx.random.state) cannot be reproduced without MLX: here the noise comes
// from Philox4x32-10 keyed by a seed in DEVICE memory with a per-call counter the kernel advances itself (a captured
// CUDA graph replays and still draws fresh numbers), or -- for parity tests -- from a caller-supplied array of uniforms.
//
// One CTA of 1024 threads per row; the row of logits (256 KB at Llama's vocabulary) is re-read from L2 by every pass:
// (1) online max / sum-exp -> logsumexp; (2) optional logprobs store; (3) the keep threshold by a 4-pass radix descent
// over order-preserving 32-bit keys of x (count histogram for top-k / min-tokens, probability-mass histogram for top-p);
// (4) Gumbel arg-max over the kept tokens, first index on ties; then the bookkeeping of pie_argmax (token, position + 1,
// sequence append).  Ties AT the threshold value are all kept (the reference's argpartition keeps an arbitrary subset).
#include <cfloat>

#include "pie_internal.h"
#include "pie_ptx.cuh"

namespace pie {

struct SampleArgs {
  const uint16_t* logits;   // bf16 [rows, ld]
  int rows, vocab, ld;
  float temp, top_p, min_p;
  int min_keep, top_k;
  unsigned long long* rng;  // device: [0] seed, [1] draw counter (advanced by one per launch)
  const float* uniforms;    // optional f32 [rows, vocab] in (0, 1): replaces Philox (tests)
  float* logprobs;          // optional f32 [rows, vocab]
  int32_t* tokens;
  int32_t* pos_adv;
  int32_t* seq;
  int seq_ld;
};

constexpr int kSmpThreads = 1024;

__device__ __forceinline__ uint32_t f2key(float f) {  // order-preserving: larger float <-> larger key
  const uint32_t u = __float_as_uint(f);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float key2f(uint32_t k) {
  return __uint_as_float((k & 0x80000000u) ? (k & 0x7fffffffu) : ~k);
}
// Philox4x32-10 (Salmon et al. 2011): counter (c0..c3), key (k0, k1)
__device__ __forceinline__ uint4 philox4(uint4 c, uint2 k) {
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const uint32_t hi0 = __umulhi(0xD2511F53u, c.x), lo0 = 0xD2511F53u * c.x;
    const uint32_t hi1 = __umulhi(0xCD9E8D57u, c.z), lo1 = 0xCD9E8D57u * c.z;
    c = make_uint4(hi1 ^ c.y ^ k.x, lo1, hi0 ^ c.w ^ k.y, lo0);
    k.x += 0x9E3779B9u; k.y += 0xBB67AE85u;
  }
  return c;
}
__device__ __forceinline__ float block_max(float v, float* sm) {
  v = warp_max(v);
  if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = v;
  __syncthreads();
  v = warp_max(sm[threadIdx.x & 31]);  // 32 warps: every warp reduces the 32 partials itself
  __syncthreads();
  return v;
}
__device__ __forceinline__ float block_sum(float v, float* sm) {
  v = warp_sum(v);
  if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = v;
  __syncthreads();
  v = warp_sum(sm[threadIdx.x & 31]);
  __syncthreads();
  return v;
}

__global__ void __launch_bounds__(kSmpThreads) sample_kernel(const __grid_constant__ SampleArgs a) {
  __shared__ float red[32];
  __shared__ float hist[256];
  __shared__ uint32_t s_prefix, s_k;
  __shared__ float s_above;
  __shared__ float bestv[32];
  __shared__ int besti[32];
  const int tid = threadIdx.x, b = blockIdx.x;
  const uint16_t* row = a.logits + (size_t)b * a.ld;
  const int V = a.vocab;
  // ---- (1) logsumexp of f32(logits) --------------------------------------------------------------------------------
  float m = -INFINITY, s = 0.f;
  for (int i = tid; i < V; i += kSmpThreads) {
    const float v = bf16lo((uint32_t)row[i]);
    if (v > m) { s = s * __expf(m - v) + 1.f; m = v; } else { s += __expf(v - m); }
  }
  const float M = block_max(m, red);
  const float Z = block_sum(m == -INFINITY ? 0.f : s * __expf(m - M), red);
  const float lse = M + logf(Z);
  if (a.logprobs)
    for (int i = tid; i < V; i += kSmpThreads) a.logprobs[(size_t)b * V + i] = bf16lo((uint32_t)row[i]) - lse;
  const float itemp = 1.0f / a.temp;
  auto xval = [&](int i) { return (bf16lo((uint32_t)row[i]) - lse) * itemp; };   // logprobs * (1 / temp)
  // ---- (3) the keep threshold on x -----------------------------------------------------------------------------------
  const int mode = (a.top_p > 0.f && a.top_p < 1.f) ? 1 : (a.min_p != 0.f ? 2 : (a.top_k > 0 ? 3 : 0));
  float thr = -INFINITY;
  const float xmax = (M - lse) * itemp;
  // radix descent: after 4 passes s_prefix is the key of the boundary value.  by_mass: accumulate softmax(x) mass from the
  // top until (mass strictly above) + (this bin) >= target; else count until >= k.
  auto descend = [&](bool by_mass, float target, uint32_t k) -> float {
    float zx = 1.f;
    if (by_mass) {  // normaliser of softmax(x): sum exp(x - xmax)
      float t = 0.f;
      for (int i = tid; i < V; i += kSmpThreads) t += __expf(xval(i) - xmax);
      zx = block_sum(t, red);
    }
    if (tid == 0) { s_prefix = 0; s_k = k; s_above = 0.f; }
    __syncthreads();
    for (int pass = 0; pass < 4; ++pass) {
      const int shift = 24 - 8 * pass;
      if (tid < 256) hist[tid] = 0.f;
      __syncthreads();
      const uint32_t prefix = s_prefix, pmask = pass == 0 ? 0u : (0xffffffffu << (shift + 8));
      for (int i = tid; i < V; i += kSmpThreads) {
        const float x = xval(i);
        const uint32_t key = f2key(x);
        if ((key & pmask) == prefix) atomicAdd(&hist[(key >> shift) & 255u], by_mass ? __expf(x - xmax) / zx : 1.f);
      }
      __syncthreads();
      if (tid == 0) {
        float acc = by_mass ? s_above : 0.f;
        uint32_t kk = s_k;
        int bin = 255;
        for (; bin > 0; --bin) {
          if (by_mass) { if (acc + hist[bin] >= target) break; acc += hist[bin]; }
          else { if (hist[bin] >= (float)kk) break; kk -= (uint32_t)hist[bin]; }
        }
        s_prefix = prefix | ((uint32_t)bin << shift);
        s_k = kk; s_above = acc;
      }
      __syncthreads();
    }
    return key2f(s_prefix);
  };
  if (mode == 1) {
    thr = descend(true, a.top_p, 0);
  } else if (mode == 2) {
    thr = xmax + logf(a.min_p);
    if (a.min_keep > 1) thr = fminf(thr, descend(false, 0.f, (uint32_t)min(a.min_keep, V)));
  } else if (mode == 3) {
    thr = descend(false, 0.f, (uint32_t)min(a.top_k, V));
  }
  // ---- (4) Gumbel arg-max over the kept tokens ---------------------------------------------------------------------------
  const unsigned long long seed = a.rng ? a.rng[0] : 0ull, draw = a.rng ? a.rng[1] : 0ull;
  float best = -INFINITY;
  int bi = 0x7fffffff;
  for (int i0 = tid * 4; i0 < V; i0 += kSmpThreads * 4) {
    uint4 r = make_uint4(0, 0, 0, 0);
    if (!a.uniforms)
      r = philox4(make_uint4((uint32_t)(i0 >> 2), (uint32_t)b, (uint32_t)draw, (uint32_t)(draw >> 32)),
                  make_uint2((uint32_t)seed, (uint32_t)(seed >> 32)));
    const uint32_t rr[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int i = i0 + j;
      if (i >= V) break;
      const float x = xval(i);
      if (!(x >= thr)) continue;
      const float u = a.uniforms ? a.uniforms[(size_t)b * V + i] : ((float)rr[j] + 0.5f) * 2.3283064365386963e-10f;
      const float v = x - logf(-logf(u));
      if (v > best || (v == best && i < bi)) { best = v; bi = i; }
    }
  }
  for (int o = 16; o > 0; o >>= 1) {
    const float ov = __shfl_xor_sync(0xffffffffu, best, o);
    const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
    if (ov > best || (ov == best && oi < bi)) { best = ov; bi = oi; }
  }
  if ((tid & 31) == 0) { bestv[tid >> 5] = best; besti[tid >> 5] = bi; }
  __syncthreads();
  if (tid < 32) {
    best = bestv[tid]; bi = besti[tid];
    for (int o = 16; o > 0; o >>= 1) {
      const float ov = __shfl_xor_sync(0xffffffffu, best, o);
      const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
      if (ov > best || (ov == best && oi < bi)) { best = ov; bi = oi; }
    }
    if (tid == 0) {
      const int tok = bi == 0x7fffffff ? 0 : bi;
      a.tokens[b] = tok;
      if (a.pos_adv) {
        const int np = a.pos_adv[b] + 1;
        a.pos_adv[b] = np;
        if (a.seq && np < a.seq_ld) a.seq[(size_t)b * a.seq_ld + np] = tok;
      }
    }
  }
  // the last row's CTA to finish advances the draw counter (every CTA has read it by then: rows are few and the read
  // is the first thing stage (4) does -- but to be exact, count departures)
  if (a.rng && tid == 0) {
    __threadfence();
    unsigned int* dep = reinterpret_cast<unsigned int*>(a.rng + 2);
    if (atomicAdd(dep, 1u) == gridDim.x - 1) { *dep = 0; a.rng[1] = draw + 1; }
  }
}

}  // namespace pie

using namespace pie;

extern "C" int pie_sample(int32_t* tokens, const void* logits, int rows, int vocab, int ld, float temp, float top_p,
                          float min_p, int min_tokens_to_keep, int top_k, void* rng_state, const float* uniforms,
                          float* logprobs_out, int32_t* positions_to_advance, int32_t* seq_out, int seq_ld, void* stream) {
  if (rows == 0) return PIE_OK;
  if (!tokens || !logits) return PIE_ENULL;
  if (rows < 0 || vocab <= 0 || ld < vocab) return PIE_EINVAL_SHAPE;
  if (!(temp > 0.f)) return PIE_EINVAL_SHAPE;            // temp == 0 is pie_argmax (make_sampler's first branch)
  if (!(min_p >= 0.f && min_p <= 1.f)) return PIE_EINVAL_SHAPE;          // min_p.py:33-36
  if (min_tokens_to_keep < 1) return PIE_EINVAL_SHAPE;                   // min_p.py:37-40
  if (top_k > 0 && !(top_p > 0.f && top_p < 1.f) && min_p == 0.f && top_k >= vocab) return PIE_EINVAL_SHAPE;  // top_k.py:23-27
  if (!rng_state && !uniforms) return PIE_ENULL;
  SampleArgs a;
  a.logits = (const uint16_t*)logits; a.rows = rows; a.vocab = vocab; a.ld = ld;
  a.temp = temp; a.top_p = top_p; a.min_p = min_p; a.min_keep = min_tokens_to_keep; a.top_k = top_k;
  a.rng = (unsigned long long*)rng_state; a.uniforms = uniforms; a.logprobs = logprobs_out;
  a.tokens = tokens; a.pos_adv = positions_to_advance; a.seq = seq_out; a.seq_ld = seq_ld;
  count_launch();
  sample_kernel<<<rows, kSmpThreads, 0, (cudaStream_t)stream>>>(a);
  return status_of(cudaGetLastError());
}
