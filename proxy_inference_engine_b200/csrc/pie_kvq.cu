// Quantised KV cache: QuantizedKVCache.update_and_fetch (PIE/cache/kv_cache/quantized.py:38-108:
// mx.quantize of the new keys / values, 8 bits, groups of 64 along head_dim) and quantized_scaled_dot_product_attention
// for L == 1 (PIE/models/base.py:56-89: scores = quantized_matmul(q * scale, K, transpose=True), softmax,
// out = quantized_matmul(scores, V, transpose=False)).
//
// HBM layout = the reference's state tuple per K and V: quants u32 [slots, D / 4] (byte e of a row = element e, as
// mx.quantize packs them, MLX/mlx/backend/cpu/quantized.cpp:603-661), scales and biases bf16 [slots, D / 64]; slot of
// (row b, kv head h, token t) = (b * Hkv + h) * capacity + t, or with a page table (page * Hkv + h) * page_size +
// t % page_size.  A token-head costs 2 * (D + 4 * D / 64) = 272 bytes at D = 128 instead of 512 in bf16.
//
// The attention kernel never materialises K, V or the scores: per 64-group  q.k = s * sum(q_e * kq_e) + b * sum(q_e),
// and  sum_t p_t * v_t = sum_t (p_t s_t) * vq_t + sum_t p_t b_t  (the second term is one scalar per group, added at the
// end).  Bytes become floats with one PRMT + one FADD (0x4B000000 | byte = 2^23 + byte); everything fp32, online softmax
// whose running maximum only moves when a score exceeds it by 8 (log2 units): the result is the same sum, the
// rescaling multiplies leave the inner loop.
#include <cuda_fp16.h>
#include <stdlib.h>

#include "pie_internal.h"
#include "pie_ptx.cuh"

namespace pie {

struct KvqArgs {
  uint32_t* kq; uint16_t* ks; uint16_t* kb;
  uint32_t* vq; uint16_t* vs; uint16_t* vb;
  const int32_t* page_table;
  int pt_stride, page_size, page_shift;   // page_size = 1 << page_shift
  long long cap;
  int Hkv;
};

__device__ __forceinline__ size_t kvq_slot(const KvqArgs& c, int b, int h, int t) {
  if (c.page_table) {
    const int page = __ldg(c.page_table + (size_t)b * c.pt_stride + (t >> c.page_shift));
    return (((size_t)page * c.Hkv + h) << c.page_shift) + (t & (c.page_size - 1));
  }
  return ((size_t)b * c.Hkv + h) * (size_t)c.cap + t;
}

// mx.quantize of one group of 64 by one warp (lane holds elements 2*lane, 2*lane + 1), 8 bits: the operations and
// their order are those of quantize_kernel (pie_misc.cu) = MLX/mlx/backend/cpu/quantized.cpp:612-659; scale and bias
// quantise in fp32 and are stored rounded to bf16, as the reference's quantize<bfloat16_t, uint32_t> does.
__device__ __forceinline__ void kvq_store_group(float e0, float e1, int lane, uint32_t* dq, uint16_t* ds, uint16_t* db) {
  float w_max = fmaxf(e0, e1), w_min = fminf(e0, e1);
#pragma unroll
  for (int o = 16; o; o >>= 1) {
    w_max = fmaxf(w_max, __shfl_xor_sync(0xffffffffu, w_max, o));
    w_min = fminf(w_min, __shfl_xor_sync(0xffffffffu, w_min, o));
  }
  const bool mask = fabsf(w_min) > fabsf(w_max);
  float scale = __fdiv_rn(__fsub_rn(w_max, w_min), 255.f);
  scale = (scale < 1e-7f) ? 1e-7f : scale;
  scale = mask ? scale : -scale;
  const float edge = mask ? w_min : w_max;
  const float q0 = rintf(__fdiv_rn(edge, scale));
  float bias = 0.f;
  if (q0 != 0.f) { scale = __fdiv_rn(edge, q0); bias = edge; }
  const float a = fminf(fmaxf(rintf(__fdiv_rn(__fsub_rn(e0, bias), scale)), 0.f), 255.f);
  const float c = fminf(fmaxf(rintf(__fdiv_rn(__fsub_rn(e1, bias), scale)), 0.f), 255.f);
  const uint32_t two = (uint32_t)a | ((uint32_t)c << 8);
  const uint32_t hi = __shfl_down_sync(0xffffffffu, two, 1);
  if (!(lane & 1)) dq[lane >> 1] = two | (hi << 16);
  if (lane == 0) { *ds = f2bf(scale); *db = f2bf(bias); }
}

// decode (L == 1): one CTA per (head, row) of the fused q/k/v projection output, D / 2 threads.  q heads are rotated
// into q_out; k heads are rotated, rounded to bf16 (the reference quantises the bf16 output of rope) and quantised into
// the cache at positions[b]; v heads are quantised into the cache.
__global__ void kvq_append_kernel(uint16_t* q_out, const uint16_t* qkv, const KvqArgs c, const int32_t* pos,
                                  const float* freqs, int Hq, int D) {
  __shared__ float row[128];
  pdl_launch_dependents();
  pdl_wait();
  const int h = blockIdx.x, b = blockIdx.y, i = threadIdx.x, half = D / 2;
  const int lane = i & 31, warp = i >> 5;
  const int p = pos[b];
  const uint16_t* src = qkv + ((size_t)b * (Hq + 2 * c.Hkv) + h) * D;
  const bool is_v = h >= Hq + c.Hkv;
  if (is_v) {
    row[i] = bf16lo((uint32_t)src[i]);
    row[i + half] = bf16lo((uint32_t)src[i + half]);
  } else {
    const float theta = (float)p * (1.0f / freqs[i]);
    float s, co;
    rope_sincos(theta, &s, &co);
    const float x1 = bf16lo((uint32_t)src[i]), x2 = bf16lo((uint32_t)src[i + half]);
    const uint16_t o1 = f2bf(__fsub_rn(__fmul_rn(x1, co), __fmul_rn(x2, s)));
    const uint16_t o2 = f2bf(__fadd_rn(__fmul_rn(x1, s), __fmul_rn(x2, co)));
    if (h < Hq) {
      uint16_t* dst = q_out + ((size_t)b * Hq + h) * D;
      dst[i] = o1;
      dst[i + half] = o2;
      return;
    }
    row[i] = bf16lo((uint32_t)o1);
    row[i + half] = bf16lo((uint32_t)o2);
  }
  __syncthreads();
  const int kvh = is_v ? h - Hq - c.Hkv : h - Hq;
  const size_t slot = kvq_slot(c, b, kvh, p);
  const int G = D / 64;
  uint32_t* dq = (is_v ? c.vq : c.kq) + slot * (D / 4) + warp * 16;
  uint16_t* ds = (is_v ? c.vs : c.ks) + slot * G + warp;
  uint16_t* db = (is_v ? c.vb : c.kb) + slot * G + warp;
  kvq_store_group(row[warp * 64 + 2 * lane], row[warp * 64 + 2 * lane + 1], lane, dq, ds, db);
}

// prefill: k, v bf16 [B, Hkv, L, D] (k already rotated) -> cache rows offset .. offset + L - 1; one warp per group
__global__ void kvq_write_kernel(const KvqArgs c, const uint16_t* k, const uint16_t* v, int L, int D, int offset) {
  const int l = blockIdx.x, h = blockIdx.y, b = blockIdx.z;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, G = D / 64;
  const size_t src = (((size_t)b * c.Hkv + h) * L + l) * D + warp * 64 + 2 * lane;
  const size_t slot = kvq_slot(c, b, h, offset + l);
  const uint32_t kk = *reinterpret_cast<const uint32_t*>(k + src), vv = *reinterpret_cast<const uint32_t*>(v + src);
  kvq_store_group(bf16lo(kk), bf16hi(kk), lane, c.kq + slot * (D / 4) + warp * 16, c.ks + slot * G + warp,
                  c.kb + slot * G + warp);
  kvq_store_group(bf16lo(vv), bf16hi(vv), lane, c.vq + slot * (D / 4) + warp * 16, c.vs + slot * G + warp,
                  c.vb + slot * G + warp);
}

struct DecAttnQArgs {
  uint16_t* out;
  const uint16_t* q;
  KvqArgs c;
  const int32_t* pos;
  int Hq, nsplit, rg;
  float scale_log2e;
  float* ws_part;        // [B*Hq][nsplit][D + 2]
  unsigned int* ws_cnt;  // [B * Hkv * rg]
};

// packed fp32 pairs (FFMA2 / FADD2 / FMUL2 of sm_100): one instruction, two lanes of fp32 arithmetic
__device__ __forceinline__ uint64_t pk2(float lo, float hi) {
  uint64_t r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
  return r;
}
__device__ __forceinline__ uint64_t pk2u(uint32_t lo, uint32_t hi) {
  uint64_t r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "r"(lo), "r"(hi));
  return r;
}
__device__ __forceinline__ void upk2(uint64_t v, float& lo, float& hi) {
  asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ uint64_t ffma2(uint64_t a, uint64_t b, uint64_t c) {
  uint64_t d;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
  return d;
}
__device__ __forceinline__ uint64_t fadd2(uint64_t a, uint64_t b) {
  uint64_t d;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  return d;
}
__device__ __forceinline__ uint64_t fmul2(uint64_t a, uint64_t b) {
  uint64_t d;
  asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
  return d;
}

// 16 bytes -> 8 pairs of floats: 0x4B000000 | byte = 2^23 + byte (PRMT), minus 2^23 (one FADD2 per pair)
__device__ __forceinline__ void bytes16_to_f32x2(const uint4& v, uint64_t* f, uint64_t neg) {
  const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    f[2 * j] = fadd2(pk2u(__byte_perm(w[j], 0x4B000000u, 0x7540), __byte_perm(w[j], 0x4B000000u, 0x7541)), neg);
    f[2 * j + 1] = fadd2(pk2u(__byte_perm(w[j], 0x4B000000u, 0x7542), __byte_perm(w[j], 0x4B000000u, 0x7543)), neg);
  }
}

// grid (nsplit, Hkv * rg, B); 4 warps; a lane owns 16 features (one 16-byte load of quants), D / 16 lanes cover a
// token, a warp handles 32 / (D / 16) tokens at a time.  Partials of the splits are merged in split order by the last
// CTA to arrive, as in decode_attn_kernel (pie_attn.cu), whose workspace layout this kernel shares.
template <int D, int R>
__global__ void __launch_bounds__(128, 2) decode_attn_q8_kernel(const DecAttnQArgs a) {
  constexpr int LPT = D / 16;     // lanes per token
  constexpr int TPW = 32 / LPT;   // tokens per warp pass
  constexpr int E = 16;
  constexpr int PW = D + 2;
  constexpr int G = D / 64;
  constexpr int kStages = 4;
  constexpr int kStageBytes = 1024 + 16 * TPW;   // K quants, V quants (16 bytes per lane each), 16 bytes per token
  __shared__ float sm[4][R][PW];
  __shared__ __align__(16) uint8_t ring[4][kStages][kStageBytes];
  __shared__ int s_last;
  pdl_launch_dependents();
  pdl_wait();
  const int split = blockIdx.x, b = blockIdx.z;
  const int kvh = blockIdx.y / a.rg, rgi = blockIdx.y % a.rg;
  const int qh0 = kvh * (a.rg * R) + rgi * R;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, sub = lane % LPT, tg = lane / LPT;
  const int grp = sub >> 2;       // the 64-group of this lane's features
  const int ctx = a.pos[b] + 1;
  const int per = (ctx + a.nsplit - 1) / a.nsplit;
  const int t0 = split * per, t1 = min(ctx, t0 + per);

  constexpr int E2 = E / 2;
  uint64_t q2[R][E2];   // pairs of features
  float qs[R];
#pragma unroll
  for (int r = 0; r < R; ++r) {
    const uint4* qp = reinterpret_cast<const uint4*>(a.q + ((size_t)b * a.Hq + qh0 + r) * D + sub * E);
    const uint4 lo = ldcg128(qp), hi = ldcg128(qp + 1);
    const uint32_t w[8] = {lo.x, lo.y, lo.z, lo.w, hi.x, hi.y, hi.z, hi.w};
    float t = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float x = bf16lo(w[j]) * a.scale_log2e, y = bf16hi(w[j]) * a.scale_log2e;
      q2[r][j] = pk2(x, y);
      t += x + y;
    }
    asm volatile("" : "+f"(t));   // a value, not a recipe: keep the 16 additions out of the token loop
    qs[r] = t;
  }
  float m[R], l[R], bacc[R];
  uint64_t acc2[R][E2];
#pragma unroll
  for (int r = 0; r < R; ++r) {
    m[r] = -INFINITY; l[r] = 0.f; bacc[r] = 0.f;
#pragma unroll
    for (int e = 0; e < E2; ++e) acc2[r][e] = 0ull;
  }
  const uint64_t neg2 = pk2(-8388608.f, -8388608.f);
  // per-warp ring of kStages passes fed by cp.async: a lane copies (and later reads) its own 16 bytes of K and of V
  // quants; lanes sub 0..3 of a token copy the 4-byte words holding its k-scale / k-bias / v-scale / v-bias.
  const uint32_t ring0 = smem_u32(&ring[warp][0][0]);
  const int32_t* ptab = a.c.page_table ? a.c.page_table + (size_t)b * a.c.pt_stride : nullptr;
  const int pshift = a.c.page_shift, pmask = a.c.page_size - 1, hkv = a.c.Hkv;
  const size_t slot0 = ((size_t)b * hkv + kvh) * (size_t)a.c.cap;
  const uint32_t* kq_l = a.c.kq + sub * 4;
  const uint32_t* vq_l = a.c.vq + sub * 4;
  const char* sc_l = reinterpret_cast<const char*>(sub == 0 ? a.c.ks : sub == 1 ? a.c.kb : sub == 2 ? a.c.vs : a.c.vb);
  const uint32_t ring_l = ring0 + lane * 16, ring_s = ring0 + 1024 + tg * 16 + sub * 4;
  auto issue = [&](int stage, int tb) {
    const int t = tb + tg;
    if (t < t1) {
      const size_t slot = ptab ? ((((size_t)__ldg(ptab + (t >> pshift)) * hkv + kvh) << pshift) + (t & pmask)) : slot0 + t;
      const uint32_t so = stage * kStageBytes;
      cp_async16(ring_l + so, kq_l + slot * (D / 4));
      cp_async16(ring_l + so + 512, vq_l + slot * (D / 4));
      if (sub < 4) cp_async4(ring_s + so, sc_l + ((slot * (G * 2)) & ~(size_t)3));
    }
    cp_async_commit();
  };
  constexpr int kStep = 4 * TPW;
  int tb_issue = t0 + warp * TPW;
#pragma unroll
  for (int s = 0; s < kStages - 1; ++s) { issue(s, tb_issue); tb_issue += kStep; }
  int stage = 0;
  for (int tb = t0 + warp * TPW; tb < t1; tb += kStep) {  // warp-uniform trip count; tail lanes idle
    __syncwarp();   // every lane is done with the stage about to be refilled
    issue(stage == 0 ? kStages - 1 : stage - 1, tb_issue);
    tb_issue += kStep;
    cp_async_wait<kStages - 1>();
    __syncwarp();
    const int t = tb + tg;
    const bool ok = t < t1;
    // a pass beyond t1 reads whatever the ring holds: its probability is forced to 0 and its scales are not used
    const uint32_t src = ring0 + stage * kStageBytes;
    const uint4 kr = lds128(src + lane * 16), vr = lds128(src + 512 + lane * 16);
    const uint4 sc = lds128(src + 1024 + tg * 16);
    const int sh = (G == 2 ? grp : (t & 1)) * 16;   // G == 1: the word holds the scales of slots 2j, 2j + 1
    const float sk = __uint_as_float((sc.x >> sh) << 16), bk = __uint_as_float((sc.y >> sh) << 16);
    const float sv = ok ? __uint_as_float((sc.z >> sh) << 16) : 0.f;
    const float bv = ok ? __uint_as_float((sc.w >> sh) << 16) : 0.f;
    uint64_t kf2[E2], vf2[E2];
    bytes16_to_f32x2(kr, kf2, neg2);
    bytes16_to_f32x2(vr, vf2, neg2);
    // scores of the R heads first (independent chains), then one rare branch, then the accumulation: no per-head
    // branches, so the compiler interleaves the heads
    float s[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
      uint64_t d2 = 0ull;
#pragma unroll
      for (int e = 0; e < E2; ++e) d2 = ffma2(q2[r][e], kf2[e], d2);
      float dx, dy;
      upk2(d2, dx, dy);
      s[r] = fmaf(sk, dx + dy, bk * qs[r]);
    }
#pragma unroll
    for (int o = 1; o < LPT; o <<= 1) {
#pragma unroll
      for (int r = 0; r < R; ++r) s[r] += __shfl_xor_sync(0xffffffffu, s[r], o);
    }
    bool need = false;
#pragma unroll
    for (int r = 0; r < R; ++r) need |= s[r] > m[r] + 8.f;
    if (ok && need) {   // rare after the first tokens: move the reference point of the running sums
#pragma unroll
      for (int r = 0; r < R; ++r) {
        if (s[r] > m[r] + 8.f) {
          const float corr = exp2f(m[r] - s[r]);
          const uint64_t c2 = pk2(corr, corr);
          l[r] *= corr; bacc[r] *= corr;
#pragma unroll
          for (int e = 0; e < E2; ++e) acc2[r][e] = fmul2(acc2[r][e], c2);
          m[r] = s[r];
        }
      }
    }
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const float p = ok ? exp2f(s[r] - m[r]) : 0.f;
      l[r] += p;
      bacc[r] = fmaf(p, bv, bacc[r]);
      const float ps = p * sv;
      const uint64_t ps2 = pk2(ps, ps);
#pragma unroll
      for (int e = 0; e < E2; ++e) acc2[r][e] = ffma2(ps2, vf2[e], acc2[r][e]);
    }
    stage = stage == kStages - 1 ? 0 : stage + 1;
  }
  cp_async_wait<0>();
  float acc[R][E];
#pragma unroll
  for (int r = 0; r < R; ++r)
#pragma unroll
    for (int e = 0; e < E2; ++e) {
      upk2(acc2[r][e], acc[r][2 * e], acc[r][2 * e + 1]);
      acc[r][2 * e] += bacc[r];
      acc[r][2 * e + 1] += bacc[r];
    }
  // merge the token slots of the warp (lanes with equal `sub` hold the same feature slice)
#pragma unroll
  for (int o = LPT; o < 32; o <<= 1) {
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
      for (int e = 0; e < E; ++e) sm[warp][r][sub * E + e] = acc[r][e];
      if (sub == 0) { sm[warp][r][D] = m[r]; sm[warp][r][D + 1] = l[r]; }
    }
  }
  __syncthreads();
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
  const int cidx = (b * a.c.Hkv + kvh) * a.rg + rgi;
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

// ---- the same attention on tensor cores (D = 128, up to 8 query heads per kv head) ---------------------------------
// mma.sync.m16n8k16 f16: the M rows are the query heads of a kv head, a CTA of 4 warps shares 64-key stages (K quants
// 8 KB + V quants 8 KB + 16 bytes of scales per key, cp.async, XOR-swizzled rows), warp w owns keys 16w .. 16w + 15.
// The quants never become fp32: 0x6400 | byte IS the f16 number 1024 + byte, one PRMT per two elements, and the
// 1024 * sum(q) (scores) / 1024 * sum(p') (values) excess is subtracted from the fp32 accumulators afterwards.  Per
// 64-group: scores = sk * (acc - 1024 qsum) + bk * qsum; values: the A operand is p * sv (f16), the bias term one
// scalar per group.  The k-slots of a thread ({2j, 2j+1, 2j+8, 2j+9}) are mapped to 4 consecutive features so a B
// fragment of K comes from one 32-bit word.  The value product is transposed (V as the A operand, features x keys; p * sv
// as B, keys x heads -- the score accumulators ARE that B fragment): every accumulator row is live, and rows g, g + 8 of
// value tile (G, q) are features 64 G + 8 g + 2 q + {0, 1}, so a thread's bytes of a V row are one 64-bit word.
constexpr int kQtKeys = 64, kQtStages = 4, kQtAhead = 2, kQtThreads = 128, kQtStageBytes = 2 * kQtKeys * 128 + kQtKeys * 16;

__device__ __forceinline__ void mma_f16_16816(float (&d)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3,
                                              uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
__device__ __forceinline__ uint32_t pack_f16(float lo, float hi) {
  uint32_t r;
  asm("cvt.rn.f16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
  return r;
}
__device__ __forceinline__ float f16lo(uint32_t v) { return __half2float(__ushort_as_half((unsigned short)(v & 0xffffu))); }
__device__ __forceinline__ float f16hi(uint32_t v) { return __half2float(__ushort_as_half((unsigned short)(v >> 16))); }

__global__ void __launch_bounds__(kQtThreads) decode_attn_q8_tc_kernel(const __grid_constant__ DecAttnQArgs a) {
  constexpr int D = 128, PW = D + 2, NS = kQtStages;
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ int s_last;
  pdl_launch_dependents();
  pdl_wait();
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, g = lane >> 2, j = lane & 3;
  const int split = blockIdx.x, kvh = blockIdx.y, b = blockIdx.z;
  const int n_rep = a.Hq / a.c.Hkv, qh0 = kvh * n_rep;
  const int p = a.pos[b], ctx = p + 1;
  const int tiles = (ctx + kQtKeys - 1) / kQtKeys, per = (tiles + a.nsplit - 1) / a.nsplit;
  const int j0 = split * per, j1 = min(tiles, j0 + per);
  const uint32_t s0 = smem_u32(smem);

  // a stage: thread copies K / V chunk (tid & 7) of rows (tid >> 3) + 16 k, and two 4-byte scale words of one key.
  // The 64 keys of a stage are 64 consecutive slots (one page, or a contiguous cache): one slot computation per stage.
  const int cc = tid & 7, rr = tid >> 3, skey = tid >> 1;
  const bool one_run = !a.c.page_table || (a.c.page_size % kQtKeys) == 0;
  const uint32_t off_kv = rr * 128 + ((cc ^ (rr & 7)) << 4);   // (rr + 16 k) & 7 == rr & 7
  const uint32_t off_sc = 2 * kQtKeys * 128 + skey * 16 + (tid & 1) * 8;
  const uint16_t* sc_a = (tid & 1) ? a.c.vs : a.c.ks;
  const uint16_t* sc_b = (tid & 1) ? a.c.vb : a.c.kb;
  auto issue = [&](int jt) {   // always commits one group (empty past the end)
    if (jt < j1) {
      const uint32_t dst = s0 + ((jt - j0) % NS) * kQtStageBytes;
      const int t0 = jt * kQtKeys, rmax = p - t0;   // rows beyond the last key re-read it (masked later)
      if (one_run) {
        const size_t slot0 = kvq_slot(a.c, b, kvh, t0);
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const size_t slot = slot0 + min(rr + 16 * k, rmax);
          cp_async16(dst + off_kv + k * 2048, a.c.kq + slot * 32 + cc * 4);
          cp_async16(dst + kQtKeys * 128 + off_kv + k * 2048, a.c.vq + slot * 32 + cc * 4);
        }
        const size_t slot = slot0 + min(skey, rmax);
        cp_async4(dst + off_sc, sc_a + slot * 2);
        cp_async4(dst + off_sc + 4, sc_b + slot * 2);
      } else {
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const size_t slot = kvq_slot(a.c, b, kvh, t0 + min(rr + 16 * k, rmax));
          cp_async16(dst + off_kv + k * 2048, a.c.kq + slot * 32 + cc * 4);
          cp_async16(dst + kQtKeys * 128 + off_kv + k * 2048, a.c.vq + slot * 32 + cc * 4);
        }
        const size_t slot = kvq_slot(a.c, b, kvh, t0 + min(skey, rmax));
        cp_async4(dst + off_sc, sc_a + slot * 2);
        cp_async4(dst + off_sc + 4, sc_b + slot * 2);
      }
    }
    cp_async_commit();
  };
  // kQtAhead stages in flight, one more than that in the ring: the stage refilled at step jt was read at step jt - 2,
  // and the barrier of step jt - 1 already orders every warp behind that -- one barrier per stage, not two
  for (int s = 0; s < kQtAhead; ++s) issue(j0 + s);

  // Q as f16 A fragments: row g = head qh0 + g; k-slots {2j, 2j+1 | 2j+8, 2j+9} of step kc in group G hold features
  // 64 G + 16 j + 4 kc + {0, 1 | 2, 3}
  uint32_t qf[2][4][2];
  float qs[2];   // sum of the head's q over each 64-group
  {
    const bool valid = g < n_rep;
    const uint16_t* qrow = a.q + ((size_t)b * a.Hq + qh0 + (valid ? g : 0)) * D;
#pragma unroll
    for (int G = 0; G < 2; ++G) {
      const uint4 lo = ldcg128(qrow + 64 * G + 16 * j), hi = ldcg128(qrow + 64 * G + 16 * j + 8);
      const uint32_t w[8] = {lo.x, lo.y, lo.z, lo.w, hi.x, hi.y, hi.z, hi.w};
      float t = 0.f;
#pragma unroll
      for (int kc = 0; kc < 4; ++kc) {
        const float x0 = valid ? bf16lo(w[2 * kc]) : 0.f, x1 = valid ? bf16hi(w[2 * kc]) : 0.f;
        const float x2 = valid ? bf16lo(w[2 * kc + 1]) : 0.f, x3 = valid ? bf16hi(w[2 * kc + 1]) : 0.f;
        qf[G][kc][0] = pack_f16(x0, x1);
        qf[G][kc][1] = pack_f16(x2, x3);
        t += (f16lo(qf[G][kc][0]) + f16hi(qf[G][kc][0])) + (f16lo(qf[G][kc][1]) + f16hi(qf[G][kc][1]));
      }
      t += __shfl_xor_sync(0xffffffffu, t, 1);
      t += __shfl_xor_sync(0xffffffffu, t, 2);
      qs[G] = t;
    }
  }
  // value accumulators, transposed: tile mt = 16 features (rows g, g + 8 -> features 64 G + 8 g + 2 (mt & 3) + {0, 1},
  // G = mt / 4) x 8 heads (columns 2j, 2j + 1): every row of every tile is live
  float o[8][4];
#pragma unroll
  for (int i = 0; i < 8; ++i) { o[i][0] = o[i][1] = o[i][2] = o[i][3] = 0.f; }
  float m = -INFINITY, l = 0.f, psum[2] = {0.f, 0.f}, bacc[2] = {0.f, 0.f};
  const float sl2e = a.scale_log2e;

  for (int jt = j0; jt < j1; ++jt) {
    issue(jt + kQtAhead);
    cp_async_wait<kQtAhead>();
    __syncthreads();
    const uint32_t tk = s0 + ((jt - j0) % NS) * kQtStageBytes, tv = tk + kQtKeys * 128, ts = tv + kQtKeys * 128;
    // ---- scores of this warp's 16 keys: [2 key tiles of 8][2 groups]
    float sacc[2][2][4];
#pragma unroll
    for (int nt = 0; nt < 2; ++nt)
#pragma unroll
      for (int G = 0; G < 2; ++G) {
        sacc[nt][G][0] = sacc[nt][G][1] = sacc[nt][G][2] = sacc[nt][G][3] = 0.f;
        const int key = warp * 16 + nt * 8 + g;
        const uint4 w4 = lds128(tk + key * 128 + (((4 * G + j) ^ (key & 7)) << 4));
        const uint32_t w[4] = {w4.x, w4.y, w4.z, w4.w};
#pragma unroll
        for (int kc = 0; kc < 4; ++kc)
          mma_f16_16816(sacc[nt][G], qf[G][kc][0], 0u, qf[G][kc][1], 0u, __byte_perm(w[kc], 0x64646464u, 0x4140),
                        __byte_perm(w[kc], 0x64646464u, 0x4342));
      }
    // ---- this thread's 4 keys (head g): keys 16 w + 8 nt + 2 j + e
    float sc[2][2], svv[2][2][2], bvv[2][2][2];
    float mx = -INFINITY;
#pragma unroll
    for (int nt = 0; nt < 2; ++nt)
#pragma unroll
      for (int e = 0; e < 2; ++e) {
        const int key = warp * 16 + nt * 8 + 2 * j + e;
        const uint4 z = lds128(ts + key * 16);   // {sk, bk, sv, bv} x {group 0 (low half), group 1 (high half)}
        const float S0 = sacc[nt][0][e] - 1024.f * qs[0], S1 = sacc[nt][1][e] - 1024.f * qs[1];
        float v = bf16lo(z.x) * S0 + bf16lo(z.y) * qs[0] + bf16hi(z.x) * S1 + bf16hi(z.y) * qs[1];
        v *= sl2e;
        if (jt * kQtKeys + key > p) v = -INFINITY;
        sc[nt][e] = v;
        mx = fmaxf(mx, v);
        svv[nt][e][0] = bf16lo(z.z); svv[nt][e][1] = bf16hi(z.z);
        bvv[nt][e][0] = bf16lo(z.w); bvv[nt][e][1] = bf16hi(z.w);
      }
    mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 1));
    mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 2));
    const bool mv = mx > m + 8.f;
    if (__any_sync(0xffffffffu, mv)) {   // move the reference point of the running sums (rare after the first stage)
      const float c = mv ? exp2f(m - mx) : 1.f;   // m == -inf: 0, and everything it multiplies is 0
      l *= c; psum[0] *= c; psum[1] *= c; bacc[0] *= c; bacc[1] *= c;
      if (mv) m = mx;
      // the accumulators of this thread belong to heads 2j and 2j + 1 (lanes 8j and 8j + 4 hold their factors)
      const float ca = __shfl_sync(0xffffffffu, c, 8 * j), cb = __shfl_sync(0xffffffffu, c, 8 * j + 4);
#pragma unroll
      for (int i = 0; i < 8; ++i) { o[i][0] *= ca; o[i][1] *= cb; o[i][2] *= ca; o[i][3] *= cb; }
    }
    uint32_t pa[2][2];   // [group][key tile] = A registers a0 / a2: p * sv as f16 pairs (keys 2j, 2j + 1)
    {
      float pr[2][2];
#pragma unroll
      for (int nt = 0; nt < 2; ++nt)
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          pr[nt][e] = sc[nt][e] == -INFINITY ? 0.f : exp2f(sc[nt][e] - m);
          l += pr[nt][e];
        }
#pragma unroll
      for (int G = 0; G < 2; ++G)
#pragma unroll
        for (int nt = 0; nt < 2; ++nt) {
          pa[G][nt] = pack_f16(pr[nt][0] * svv[nt][0][G], pr[nt][1] * svv[nt][1][G]);
          psum[G] += f16lo(pa[G][nt]) + f16hi(pa[G][nt]);
          bacc[G] = fmaf(pr[nt][0], bvv[nt][0][G], fmaf(pr[nt][1], bvv[nt][1][G], bacc[G]));
        }
    }
    // ---- values, V as the A operand (features x keys) and p * sv as B (keys x heads): a thread's A registers are the
    // bytes of features 64 G + 8 g + {2q, 2q + 1} of its four keys 2j, 2j+1, 2j+8, 2j+9
#pragma unroll
    for (int G = 0; G < 2; ++G) {
      uint32_t vw[4][2];
#pragma unroll
      for (int r4 = 0; r4 < 4; ++r4) {
        const int key = warp * 16 + 2 * j + (r4 & 1) + (r4 >> 1) * 8;
        const uint32_t ad = tv + key * 128 + (((4 * G + (g >> 1)) ^ (key & 7)) << 4) + (g & 1) * 8;
        asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(vw[r4][0]), "=r"(vw[r4][1]) : "r"(ad));
      }
#pragma unroll
      for (int q4 = 0; q4 < 4; ++q4) {   // bytes 2 q4, 2 q4 + 1 of the 8-byte run
        const int wd = q4 >> 1;
        const uint32_t sel = (q4 & 1) ? 0x7362u : 0x5140u;
        const uint32_t xab = __byte_perm(vw[0][wd], vw[1][wd], sel), xcd = __byte_perm(vw[2][wd], vw[3][wd], sel);
        mma_f16_16816(o[4 * G + q4], __byte_perm(xab, 0x64646464u, 0x4140), __byte_perm(xab, 0x64646464u, 0x4342),
                      __byte_perm(xcd, 0x64646464u, 0x4140), __byte_perm(xcd, 0x64646464u, 0x4342), pa[G][0], pa[G][1]);
      }
    }
  }
  cp_async_wait<0>();
  __syncthreads();   // every warp is done with the stages: they become the merge records
#pragma unroll
  for (int sft = 1; sft <= 2; sft <<= 1) {
    l += __shfl_xor_sync(0xffffffffu, l, sft);
#pragma unroll
    for (int G = 0; G < 2; ++G) {
      psum[G] += __shfl_xor_sync(0xffffffffu, psum[G], sft);
      bacc[G] += __shfl_xor_sync(0xffffffffu, bacc[G], sft);
    }
  }
  // per-warp partials -> shared memory; thread (g, j) holds heads 2j, 2j + 1 of features 64 G + 8 g + 2 q + {0, 1}
  float* rec = reinterpret_cast<float*>(smem);  // [4 warps][n_rep][PW]
#pragma unroll
  for (int G = 0; G < 2; ++G) {
    const float fix = bacc[G] - 1024.f * psum[G];   // of head g: fetch those of heads 2j, 2j + 1
    const float fa = __shfl_sync(0xffffffffu, fix, 8 * j), fb = __shfl_sync(0xffffffffu, fix, 8 * j + 4);
#pragma unroll
    for (int q4 = 0; q4 < 4; ++q4) {
      const int f = 64 * G + 8 * g + 2 * q4;
      if (2 * j < n_rep) {
        float* ra = rec + ((size_t)warp * n_rep + 2 * j) * PW;
        ra[f] = o[4 * G + q4][0] + fa;
        ra[f + 1] = o[4 * G + q4][2] + fa;
      }
      if (2 * j + 1 < n_rep) {
        float* rb = rec + ((size_t)warp * n_rep + 2 * j + 1) * PW;
        rb[f] = o[4 * G + q4][1] + fb;
        rb[f + 1] = o[4 * G + q4][3] + fb;
      }
    }
  }
  if (g < n_rep && j == 0) {
    float* r0 = rec + ((size_t)warp * n_rep + g) * PW;
    r0[D] = m; r0[D + 1] = l;
  }
  __syncthreads();
  const size_t rec0 = ((size_t)b * a.Hq + qh0) * a.nsplit;
  for (int i = tid; i < n_rep * D; i += kQtThreads) {
    const int r = i / D, d = i % D;
    float mn = -INFINITY;
    for (int w = 0; w < 4; ++w) mn = fmaxf(mn, rec[((size_t)w * n_rep + r) * PW + D]);
    float lt = 0.f, at = 0.f;
    for (int w = 0; w < 4; ++w) {
      const float* rw = rec + ((size_t)w * n_rep + r) * PW;
      const float cw = rw[D] == -INFINITY ? 0.f : exp2f(rw[D] - mn);
      lt += rw[D + 1] * cw;
      at += rw[d] * cw;
    }
    if (a.nsplit == 1) {
      a.out[((size_t)b * a.Hq + qh0 + r) * D + d] = f2bf(at / lt);
    } else {
      float* pr = a.ws_part + (rec0 + (size_t)r * a.nsplit + split) * PW;
      pr[d] = at;
      if (d == 0) { pr[D] = mn; pr[D + 1] = lt; }
    }
  }
  if (a.nsplit == 1) return;
  __threadfence();
  __syncthreads();
  const int cidx = b * a.c.Hkv + kvh;
  if (tid == 0) s_last = (atomicAdd(&a.ws_cnt[cidx], 1u) == (unsigned)a.nsplit - 1);
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  for (int i = tid; i < n_rep * D; i += kQtThreads) {
    const int r = i / D, d = i % D;
    const float* pr = a.ws_part + (rec0 + (size_t)r * a.nsplit) * PW;
    float mn = -INFINITY;
    for (int s = 0; s < a.nsplit; ++s) mn = fmaxf(mn, ldcg_f32(pr + (size_t)s * PW + D));
    float lt = 0.f, at = 0.f;
    for (int s = 0; s < a.nsplit; ++s) {
      const float ms = ldcg_f32(pr + (size_t)s * PW + D);
      const float cw = ms == -INFINITY ? 0.f : exp2f(ms - mn);
      lt += ldcg_f32(pr + (size_t)s * PW + D + 1) * cw;
      at += ldcg_f32(pr + (size_t)s * PW + d) * cw;
    }
    a.out[((size_t)b * a.Hq + qh0 + r) * D + d] = f2bf(at / lt);
  }
  if (tid == 0) a.ws_cnt[cidx] = 0;
}

static cudaError_t launch_decode_attn_q8_tc(const DecAttnQArgs& a, int B, cudaStream_t st) {
  size_t smem = (size_t)kQtStages * kQtStageBytes;
  {
    // Spread the grid: the block scheduler fills an SM to its residency limit before it moves on, so a grid of fewer
    // CTAs than the machine can hold leaves SMs idle.  Asking for more shared memory than the kernel needs caps the
    // residency at ceil(CTAs / SMs) and the grid covers every SM.
    const long long ctas = (long long)a.nsplit * a.c.Hkv * B;
    const int sms = num_sms();
    long long per_sm = (ctas + sms - 1) / sms;
    static const int force = [] { const char* e = getenv("PIE_KVQ_PER_SM"); return e ? atoi(e) : 0; }();
    if (force > 0) per_sm = force;
    if (per_sm >= 1 && per_sm <= 2) {
      const size_t want = (size_t)(227 * 1024) / (size_t)per_sm - 1024 - 1152;   // 1 KB reserved per CTA + static
      if (want > smem) smem = want & ~(size_t)127;
    }
  }
  {  // per launch, not once per process: the attribute is per device
    cudaError_t e = cudaFuncSetAttribute(decode_attn_q8_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
  }
  count_launch();
  return launch_pdl(decode_attn_q8_tc_kernel, dim3(a.nsplit, a.c.Hkv, B), dim3(kQtThreads), smem, st, a);
}

template <int D>
static cudaError_t launch_decode_attn_q8_d(const DecAttnQArgs& a0, int B, int n_rep, cudaStream_t st) {
  DecAttnQArgs a = a0;
  const int R = n_rep % 4 == 0 ? 4 : (n_rep % 2 == 0 ? 2 : 1);
  a.rg = n_rep / R;
  const dim3 grid(a.nsplit, a.c.Hkv * a.rg, B);
  count_launch();
  if (R == 4) decode_attn_q8_kernel<D, 4><<<grid, 128, 0, st>>>(a);
  else if (R == 2) decode_attn_q8_kernel<D, 2><<<grid, 128, 0, st>>>(a);
  else decode_attn_q8_kernel<D, 1><<<grid, 128, 0, st>>>(a);
  return cudaGetLastError();
}

static int kvq_check(const pie_kvq_cache* c, int Hkv, int D, KvqArgs* o) {
  if (!c) return PIE_ENULL;
  if (!c->k_quant || !c->k_scales || !c->k_biases || !c->v_quant || !c->v_scales || !c->v_biases) return PIE_ENULL;
  if (c->group_size != 64 || c->bits != 8) return PIE_EUNSUPPORTED;
  if (D != 64 && D != 128) return PIE_EUNSUPPORTED;
  // pages: a power of two of at least 2 tokens; capacities even (a 4-byte copy fetches the scales of a pair of slots)
  if (c->page_table ? (c->page_size < 2 || c->pt_stride <= 0 || (c->page_size & (c->page_size - 1)))
                    : (c->capacity <= 0 || (c->capacity & 1)))
    return PIE_EINVAL_SHAPE;
  if (((uintptr_t)c->k_quant | (uintptr_t)c->v_quant) & 15) return PIE_EALIGN;
  o->kq = c->k_quant; o->ks = (uint16_t*)c->k_scales; o->kb = (uint16_t*)c->k_biases;
  o->vq = c->v_quant; o->vs = (uint16_t*)c->v_scales; o->vb = (uint16_t*)c->v_biases;
  o->page_table = c->page_table; o->pt_stride = c->pt_stride; o->page_size = c->page_size;
  o->page_shift = 0;
  while (c->page_table && (1 << o->page_shift) < c->page_size) ++o->page_shift;
  o->cap = c->capacity; o->Hkv = Hkv;
  return PIE_OK;
}

}  // namespace pie

using namespace pie;

extern "C" {

int pie_kvq_append(void* q_out, const void* qkv, const pie_kvq_cache* cache, const int32_t* positions,
                   const float* freqs, int B, int Hq, int Hkv, int D, void* stream) {
  if (B == 0) return PIE_OK;
  if (!q_out || !qkv || !positions || !freqs) return PIE_ENULL;
  if (B < 0 || Hq <= 0 || Hkv <= 0) return PIE_EINVAL_SHAPE;
  KvqArgs c;
  if (int s = kvq_check(cache, Hkv, D, &c)) return s;
  count_launch();
  return status_of(launch_pdl(kvq_append_kernel, dim3(Hq + 2 * Hkv, B), dim3(D / 2), 0, (cudaStream_t)stream,
                              (uint16_t*)q_out, (const uint16_t*)qkv, c, positions, freqs, Hq, D));
}

int pie_kvq_write(const pie_kvq_cache* cache, const void* k, const void* v, int B, int Hkv, int L, int D, int offset,
                  void* stream) {
  if (B == 0 || L == 0) return PIE_OK;
  if (!k || !v) return PIE_ENULL;
  if (B < 0 || Hkv <= 0 || L < 0 || offset < 0) return PIE_EINVAL_SHAPE;
  KvqArgs c;
  if (int s = kvq_check(cache, Hkv, D, &c)) return s;
  if (!c.page_table && offset + L > c.cap) return PIE_EINVAL_SHAPE;
  count_launch();
  kvq_write_kernel<<<dim3(L, Hkv, B), D / 2, 0, (cudaStream_t)stream>>>(c, (const uint16_t*)k, (const uint16_t*)v, L, D,
                                                                       offset);
  return status_of(cudaGetLastError());
}

int pie_decode_attn_q8_pick_splits(int B, int Hq, int Hkv, int max_context) {
  if (B <= 0 || Hq <= 0 || Hkv <= 0 || Hq % Hkv) return 1;
  const int n_rep = Hq / Hkv, R = n_rep % 4 == 0 ? 4 : (n_rep % 2 == 0 ? 2 : 1);
  const long long units = (long long)B * Hkv * (n_rep <= 8 ? 1 : n_rep / R);   // <= 8: the tensor-core kernel, one CTA per kv head
  // two CTAs of four warps fit an SM; one wave of them, and at least 512 tokens per split: a CTA's fixed cost (q load,
  // ring fill, merge through the workspace) is worth about 200 tokens (measured: bs32 ctx2048 79 / 93 / 99 us at 1 / 2 /
  // 4 splits; bs1 ctx8192 31 / 37 / 60 us at 16 / 32 / 64)
  long long n = (2LL * num_sms()) / units;
  const long long by_ctx = max_context / 512 > 1 ? max_context / 512 : 1;
  if (n > by_ctx) n = by_ctx;
  if (n > 64) n = 64;
  return n < 1 ? 1 : (int)n;
}

int pie_decode_attn_q8(void* out, const void* q, const pie_kvq_cache* cache, const int32_t* positions, int B, int Hq,
                       int Hkv, int D, float scale, int nsplit, void* workspace, size_t workspace_bytes, void* stream) {
  if (B == 0) return PIE_OK;
  if (!out || !q || !positions) return PIE_ENULL;
  if (B < 0 || Hq <= 0 || Hkv <= 0 || Hq % Hkv) return PIE_EINVAL_SHAPE;
  DecAttnQArgs a;
  if (int s = kvq_check(cache, Hkv, D, &a.c)) return s;
  if ((uintptr_t)q & 15) return PIE_EALIGN;
  if (nsplit <= 0) nsplit = 1;
  if (nsplit > 1 && (!workspace || workspace_bytes < pie_decode_attn_workspace_bytes(B, Hq, D, nsplit)))
    return PIE_EWORKSPACE;
  a.out = (uint16_t*)out; a.q = (const uint16_t*)q; a.pos = positions;
  a.Hq = Hq; a.nsplit = nsplit; a.rg = 1;
  a.scale_log2e = scale * 1.4426950408889634f;
  const size_t cnt_bytes = (((size_t)B * Hq * sizeof(unsigned)) + 255) & ~(size_t)255;
  a.ws_cnt = reinterpret_cast<unsigned int*>(workspace);
  a.ws_part = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(workspace) + cnt_bytes);
  // tensor-core kernel for head_dim 128 with up to 8 query heads per kv head (PIE_KVQ_TC=0: the SIMT kernel)
  static const bool tc = [] { const char* e = getenv("PIE_KVQ_TC"); return !(e && e[0] == '0'); }();
  if (tc && D == 128 && Hq / Hkv <= 8) return status_of(launch_decode_attn_q8_tc(a, B, (cudaStream_t)stream));
  return status_of(D == 128 ? launch_decode_attn_q8_d<128>(a, B, Hq / Hkv, (cudaStream_t)stream)
                            : launch_decode_attn_q8_d<64>(a, B, Hq / Hkv, (cudaStream_t)stream));
}

}  // extern "C"
