// Decode attention (L == 1, mask None): split-K flash-decode over a contiguous or paged bf16 KV cache.
// Replaces mx.fast.scaled_dot_product_attention as called from PIE/models/base.py:92-113 (Metal sdpa_vector /
// sdpa_vector_2pass, MLX/mlx/backend/metal/kernels/sdpa_vector.h:14-392): everything fp32 inside, GQA by
// reading each K/V row once for the R query heads that share it, online softmax, one bf16 rounding at the store.
//
// grid (nsplit, Hkv * RG, B); 4 warps; a warp handles 4 tokens at a time (8 lanes per token, 16-byte loads so
// that every load instruction covers whole 128-byte lines).  Partials (m, l, acc[D]) of the splits are
// merged in split order by the last CTA to arrive (deterministic).
#include "pie_internal.h"
#include "pie_ptx.cuh"

namespace pie {

__device__ __forceinline__ size_t kv_addr_a(const KvLayout& kv, int b, int h, int t) {
  if (kv.page_table) {
    const int page = __ldg(kv.page_table + (size_t)b * kv.pt_stride + t / kv.page_size);
    return ((size_t)page * kv.page_size + t % kv.page_size) * kv.tok_stride + (size_t)h * kv.head_stride;
  }
  return (size_t)b * kv.batch_stride + (size_t)h * kv.head_stride + (size_t)t * kv.tok_stride;
}

struct DecAttnArgs {
  uint16_t* out;
  const uint16_t* q;
  const uint16_t* kc;
  const uint16_t* vc;
  KvLayout kv;
  const int32_t* pos;
  int Hq, Hkv, nsplit, rg;  // rg = query-head groups (of R heads) per kv head
  float scale_log2e;
  float* ws_part;           // [B*Hq][nsplit][D + 2]
  unsigned int* ws_cnt;     // [B * Hkv * rg]
};

__device__ __forceinline__ void unpack8f(const uint4& v, float* f) {
  f[0] = bf16lo(v.x); f[1] = bf16hi(v.x); f[2] = bf16lo(v.y); f[3] = bf16hi(v.y);
  f[4] = bf16lo(v.z); f[5] = bf16hi(v.z); f[6] = bf16lo(v.w); f[7] = bf16hi(v.w);
}

template <int D, int R>
__global__ void __launch_bounds__(128) decode_attn_kernel(const DecAttnArgs a) {
  constexpr int C = D / 64;       // 16-byte chunks per lane (8 lanes cover one token row)
  constexpr int E = 8 * C;        // elements per lane
  constexpr int PW = D + 2;       // partial record: acc[D], m, l
  __shared__ float sm[4][R][PW];
  __shared__ int s_last;
  pdl_launch_dependents();
  pdl_wait();
  const int split = blockIdx.x, b = blockIdx.z;
  const int kvh = blockIdx.y / a.rg, rgi = blockIdx.y % a.rg;
  const int qh0 = kvh * (a.rg * R) + rgi * R;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, sub = lane & 7, tg = lane >> 3;
  const int ctx = a.pos[b] + 1;
  const int per = (ctx + a.nsplit - 1) / a.nsplit;
  const int t0 = split * per, t1 = min(ctx, t0 + per);

  float q[R][E];
#pragma unroll
  for (int r = 0; r < R; ++r) {
    const uint4* qp = reinterpret_cast<const uint4*>(a.q + ((size_t)b * a.Hq + qh0 + r) * D);
#pragma unroll
    for (int c = 0; c < C; ++c) {
      unpack8f(ldcg128(qp + sub + 8 * c), &q[r][8 * c]);
#pragma unroll
      for (int j = 0; j < 8; ++j) q[r][8 * c + j] *= a.scale_log2e;
    }
  }
  float m[R], l[R], acc[R][E];
#pragma unroll
  for (int r = 0; r < R; ++r) {
    m[r] = -INFINITY; l[r] = 0.f;
#pragma unroll
    for (int e = 0; e < E; ++e) acc[r][e] = 0.f;
  }
  for (int tb = t0 + warp * 4; tb < t1; tb += 16) {  // warp-uniform trip count; tail lanes idle
    const int t = tb + tg;
    const bool ok = t < t1;
    float kf[E], vf[E];
    if (ok) {
      const size_t off = kv_addr_a(a.kv, b, kvh, t);
      const uint4* kp = reinterpret_cast<const uint4*>(a.kc + off);
      const uint4* vp = reinterpret_cast<const uint4*>(a.vc + off);
      uint4 kr[C], vr[C];
#pragma unroll
      for (int c = 0; c < C; ++c) { kr[c] = ldcg128(kp + sub + 8 * c); vr[c] = ldcg128(vp + sub + 8 * c); }
#pragma unroll
      for (int c = 0; c < C; ++c) { unpack8f(kr[c], &kf[8 * c]); unpack8f(vr[c], &vf[8 * c]); }
    } else {
#pragma unroll
      for (int e = 0; e < E; ++e) { kf[e] = 0.f; vf[e] = 0.f; }
    }
#pragma unroll
    for (int r = 0; r < R; ++r) {
      float s = 0.f;
#pragma unroll
      for (int e = 0; e < E; ++e) s = fmaf(q[r][e], kf[e], s);
      s += __shfl_xor_sync(0xffffffffu, s, 1);
      s += __shfl_xor_sync(0xffffffffu, s, 2);
      s += __shfl_xor_sync(0xffffffffu, s, 4);
      if (ok) {
        const float mn = fmaxf(m[r], s);
        const float corr = exp2f(m[r] - mn), p = exp2f(s - mn);
        l[r] = l[r] * corr + p;
#pragma unroll
        for (int e = 0; e < E; ++e) acc[r][e] = fmaf(p, vf[e], acc[r][e] * corr);
        m[r] = mn;
      }
    }
  }
  // merge the 4 token slots of the warp (lanes with equal `sub` hold the same feature slice)
#pragma unroll
  for (int o = 8; o <= 16; o <<= 1) {
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const float mo = __shfl_xor_sync(0xffffffffu, m[r], o), lo = __shfl_xor_sync(0xffffffffu, l[r], o);
      const float mn = fmaxf(m[r], mo);
      const float ca = mn == -INFINITY ? 0.f : exp2f(m[r] - mn), cb = mn == -INFINITY ? 0.f : exp2f(mo - mn);
      l[r] = l[r] * ca + lo * cb;
#pragma unroll
      for (int e = 0; e < E; ++e) acc[r][e] = acc[r][e] * ca + __shfl_xor_sync(0xffffffffu, acc[r][e], o) * cb;
      m[r] = mn;
    }
  }
  if (tg == 0) {
#pragma unroll
    for (int r = 0; r < R; ++r) {
#pragma unroll
      for (int c = 0; c < C; ++c)
#pragma unroll
        for (int j = 0; j < 8; ++j) sm[warp][r][(sub + 8 * c) * 8 + j] = acc[r][8 * c + j];
      if (sub == 0) { sm[warp][r][D] = m[r]; sm[warp][r][D + 1] = l[r]; }
    }
  }
  __syncthreads();
  // merge the 4 warps; thread -> (r, d)
  const size_t rec0 = ((size_t)b * a.Hq + qh0) * a.nsplit;
  for (int i = threadIdx.x; i < R * D; i += 128) {
    const int r = i / D, d = i % D;
    float mn = -INFINITY;
    for (int w = 0; w < 4; ++w) mn = fmaxf(mn, sm[w][r][D]);
    float lt = 0.f, at = 0.f;
    for (int w = 0; w < 4; ++w) {
      const float cw = mn == -INFINITY ? 0.f : exp2f(sm[w][r][D] - mn);
      lt += sm[w][r][D + 1] * cw;
      at += sm[w][r][d] * cw;
    }
    if (a.nsplit == 1) {
      a.out[((size_t)b * a.Hq + qh0 + r) * D + d] = f2bf(at / lt);
    } else {
      float* rec = a.ws_part + (rec0 + (size_t)r * a.nsplit + split) * PW;
      rec[d] = at;
      if (d == 0) { rec[D] = mn; rec[D + 1] = lt; }
    }
  }
  if (a.nsplit == 1) return;
  __threadfence();
  __syncthreads();
  const int cidx = (b * a.Hkv + kvh) * a.rg + rgi;
  if (threadIdx.x == 0) s_last = (atomicAdd(&a.ws_cnt[cidx], 1u) == (unsigned)a.nsplit - 1);
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  for (int i = threadIdx.x; i < R * D; i += 128) {
    const int r = i / D, d = i % D;
    const float* rec = a.ws_part + (rec0 + (size_t)r * a.nsplit) * PW;
    float mn = -INFINITY;
    for (int s = 0; s < a.nsplit; ++s) mn = fmaxf(mn, ldcg_f32(rec + (size_t)s * PW + D));
    float lt = 0.f, at = 0.f;
    for (int s = 0; s < a.nsplit; ++s) {
      const float ms = ldcg_f32(rec + (size_t)s * PW + D);
      const float cw = ms == -INFINITY ? 0.f : exp2f(ms - mn);
      lt += ldcg_f32(rec + (size_t)s * PW + D + 1) * cw;
      at += ldcg_f32(rec + (size_t)s * PW + d) * cw;
    }
    a.out[((size_t)b * a.Hq + qh0 + r) * D + d] = f2bf(at / lt);
  }
  if (threadIdx.x == 0) a.ws_cnt[cidx] = 0;
}

}  // namespace pie
#include "pie_attn_launch.inl"
