// One whole decode step (L == 1) as ONE persistent cooperative kernel, sm_100a.
//
// Replaces, for a single new token per sequence, Model.__call__ (PIE/models/llama/language.py:199-210) with everything
// under it -- LlamaModel.__call__ :170-187, TransformerBlock.__call__ :144-154, Attention.__call__ :75-108,
// MLP.__call__ :126-127, QuantizedEmbedding.__call__ (MLX/python/mlx/nn/layers/quantized.py:100-107) -- and the
// temp == 0 sampler + `cache.offset += 1` of generate_step (PIE/engine/inference_engine.py:282-297).
//
// Why one kernel: at batch 1 a layer's matrices are 9..66 MB, i.e. 1.5..10 us of HBM time each, while a kernel
// boundary costs ~3-5 us of drain + launch + pipeline ramp.  Here one CTA per SM walks the step's matrices in order
// (4 per layer + the head); every warp's TMA ring simply goes on with the NEXT matrix's units when the current
// matrix's are all requested, so the weight stream crosses the grid barriers that separate the matrices.  Between
// q/k/v and o_proj the same CTAs run RoPE + KV append + split-K flash-decode attention (o_proj's first units are
// already in shared memory).  The GEMV arithmetic is that of pie_qmv2.cu (TMA-fed, fp16 staging, one row).
// Built for batch 1, 4-bit weights, K a multiple of 512 (the Llama decode step); other shapes use the
// per-operator launches.
#include "pie_internal.h"
#include "pie_ptx.cuh"
#include "pie_qmv2.h"
#include "pie_qmv2_dev.cuh"

#include <cstdlib>
#include <cstring>

namespace pie {

enum StepPre : int { PRE_BARRIER = 0, PRE_EMBED = 1, PRE_ATTN = 2 };

struct StepPhase {
  Qmv2Args q;   // holds the matrix's TMA tensor maps: the plan lives in global memory, 64-byte aligned
  int pre;
  int pad_;
  // PRE_ATTN: attention of this layer, between the q/k/v matrix and o_proj
  const uint16_t* qkv;   // bf16 [B, (Hq + 2 Hkv) D], un-rotated
  uint16_t* attn_out;    // bf16 [B, Hq D]
  uint16_t* kc;
  uint16_t* vc;
  KvLayout kv;
};

struct StepArgs {
  const StepPhase* phases;
  int nphases;
  int B, Hq, Hkv, D, R, rg, nsplit;
  float scale_log2e;
  const float* freqs;
  const float* rope_tab;   // optional [B][2][D / 2]: sin, cos of this step's rotary angles (pie_rope_table), else computed here
  int32_t* pos;
  float* attn_part;        // [B * Hq][nsplit][D + 2]
  unsigned int* attn_cnt;  // [B * Hkv * rg]
  // embedding gather
  const uint32_t* e_wq;
  const uint16_t* e_s;
  const uint16_t* e_b;
  int32_t* tokens;
  int table_rows, hidden;
  uint16_t* h;
  // sampler
  const uint16_t* logits;
  int vocab;
  int32_t* seq;
  int seq_ld;
  unsigned int* bar;  // [0] arrivals, [1] generation (grid_barrier); [2] running arrivals, [3] departures (flat form)
  int pf_cross;           // units of a later matrix a warp may request before the current matrix is done (0..2)
  unsigned long long* trace;
  unsigned long long* trace_stage;  // optional: 4 stamps per matrix from inside the x staging
};

struct alignas(64) StepHeader {  // head of the host/device plan buffer (the phases behind it hold CUtensorMaps)
  uint32_t magic, bits, grid, smem, onecol, nphases, plan_bytes, pad_;
  StepArgs args;  // args.phases is patched at launch
};
constexpr uint32_t kStepMagic = 0x50494553u;

// ---- grid-wide barrier: arrivals counter + generation word; leaves the counter at zero --------------------------
__device__ __forceinline__ void grid_barrier(unsigned int* bar) {
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned int gen = ld_acquire_gpu(bar + 1);
    __threadfence();
    if (atomicAdd(bar, 1u) == gridDim.x - 1) {
      atomicExch(bar, 0u);
      __threadfence();
      red_release_gpu_add(bar + 1, 1u);
    } else {
      while (ld_acquire_gpu(bar + 1) == gen) {}
    }
    __threadfence();
  }
  __syncthreads();
}

// Flat form used by the one-launch step: one fire-and-forget arrival and one polled word.  The word counts arrivals and
// is not reset between barriers: barrier number k (from 1) is passed when it reaches k * gridDim.x.  The last CTA to
// leave the kernel zeroes it.
__device__ __forceinline__ void grid_barrier_flat(unsigned int* cnt, unsigned int target) {
  __syncthreads();
  if (threadIdx.x == 0) {
    red_release_gpu_add(cnt, 1u);
    unsigned int spins = 0;
    while ((int)(ld_acquire_gpu(cnt) - target) < 0) {
      if (++spins > (1u << 26)) __trap();
    }
  }
  __syncthreads();
}

__device__ __forceinline__ unsigned long long globaltimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
__device__ __forceinline__ void cluster_sync() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// fp32 at shared-memory address `local_addr` of CTA `rank` of this cluster (distributed shared memory)
__device__ __forceinline__ float ld_dsmem_f32(uint32_t local_addr, int rank) {
  uint32_t remote;
  float v;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(local_addr), "r"(rank));
  asm volatile("ld.shared::cluster.f32 %0, [%1];" : "=f"(v) : "r"(remote) : "memory");
  return v;
}
// One 4-byte remote store into CTA `rank`'s shared memory that also counts 4 bytes on that CTA's mbarrier: the data is
// visible to whoever sees the barrier's phase complete -- no fence, no cluster-wide barrier (st.async, sm_90+).
__device__ __forceinline__ void st_async_f32(uint32_t local_addr, uint32_t local_bar, int rank, float v) {
  uint32_t ra, rb;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(ra) : "r"(local_addr), "r"(rank));
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(rb) : "r"(local_bar), "r"(rank));
  asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.b32 [%0], %1, [%2];" ::"r"(ra), "f"(v), "r"(rb)
               : "memory");
}
__device__ __forceinline__ size_t step_kv_addr(const KvLayout& kv, int b, int h, int t) {
  if (kv.page_table) {
    const int page = __ldg(kv.page_table + (size_t)b * kv.pt_stride + t / kv.page_size);
    return ((size_t)page * kv.page_size + t % kv.page_size) * kv.tok_stride + (size_t)h * kv.head_stride;
  }
  return (size_t)b * kv.batch_stride + (size_t)h * kv.head_stride + (size_t)t * kv.tok_stride;
}
__device__ __forceinline__ void unpack8(const uint4& v, float* f) {
  f[0] = bf16lo(v.x); f[1] = bf16hi(v.x); f[2] = bf16lo(v.y); f[3] = bf16hi(v.y);
  f[4] = bf16lo(v.z); f[5] = bf16hi(v.z); f[6] = bf16lo(v.w); f[7] = bf16hi(v.w);
}
__device__ __forceinline__ uint4 pack8(const float* f) {
  return make_uint4(pack_bf16(f[0], f[1]), pack_bf16(f[2], f[3]), pack_bf16(f[4], f[5]), pack_bf16(f[6], f[7]));
}

// Rotate one head row held 8-lanes-wide (lane `sub` owns elements 8 sub + 64 c + j, c < C, j < 8) exactly like
// rope_kv_append_kernel (pie_rope_kv.cu): non-traditional pairs (i, i + D/2), fp32, each product rounded, bf16 result.
// sn / cs: sin / cos of this lane's 8 pair indices.
template <int C>
__device__ __forceinline__ void rope_row(float* v, const float* sn, const float* cs, int sub) {
  if (C == 2) {
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float x1 = v[j], x2 = v[8 + j];
      v[j] = bfround(__fsub_rn(__fmul_rn(x1, cs[j]), __fmul_rn(x2, sn[j])));
      v[8 + j] = bfround(__fadd_rn(__fmul_rn(x1, sn[j]), __fmul_rn(x2, cs[j])));
    }
  } else {  // D == 64: the pair partner lives 4 lanes away
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const float other = __shfl_xor_sync(0xffffffffu, v[j], 4);
      const float x1 = sub < 4 ? v[j] : other, x2 = sub < 4 ? other : v[j];
      v[j] = sub < 4 ? bfround(__fsub_rn(__fmul_rn(x1, cs[j]), __fmul_rn(x2, sn[j])))
                     : bfround(__fadd_rn(__fmul_rn(x1, sn[j]), __fmul_rn(x2, cs[j])));
    }
  }
}

// RoPE (a7) + KV append (a8) + split-K flash-decode (a9) for one layer, spread over the whole grid.
// Unit = (batch row, kv head, query-head group of R, split of the context); a CTA's 16 warps take 4 tokens each per
// round.  The token being appended is taken from the q/k/v projection output (rotated in registers and written to
// the cache by the rgi == 0 unit); older tokens come from the cache.  Split partials are merged in split order by
// the last unit to arrive, as in decode_attn_kernel (pie_attn.cu).
// CLUSTER: the splits of one (row, kv head, head group) are the CTAs of one thread-block cluster (stand-alone launch,
// nsplit <= 8): their partials meet through distributed shared memory behind a cluster barrier instead of a global
// workspace + fence + atomic + re-read, and every CTA of the cluster finishes a slice of the output.
template <int D, int R, bool CLUSTER, int NW, bool PREWAIT = false>
__device__ __forceinline__ void attn_phase_impl(const StepArgs& s, const StepPhase& P, float* scratch, int* s_flag,
                                        unsigned long long* st = nullptr, uint64_t* cl_bar = nullptr) {
  constexpr int C = D / 64, E = 8 * C, PW = D + 2;
  auto stamp = [&](int i) { if (st && threadIdx.x == 0) st[i] = globaltimer_ns(); };
  if (CLUSTER && s.nsplit > 1) {
    // the cluster's first CTA collects the other splits' partials: R * (D + 2) floats each, counted by its mbarrier
    if (blockIdx.x % s.nsplit == 0 && threadIdx.x == 0) {
      mbar_init(cl_bar, 1);
      mbar_arrive_expect_tx(cl_bar, (uint32_t)(s.nsplit - 1) * R * PW * 4);
      fence_mbar_init();
    }
    asm volatile("barrier.cluster.arrive.relaxed.aligned;" ::: "memory");   // waited for just before the pushes
  }
  stamp(8);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, sub = lane & 7, tg = lane >> 3;
  const int units = s.B * s.Hkv * s.rg * s.nsplit;
  const int qkv_ld = (s.Hq + 2 * s.Hkv) * D;
  for (int u = blockIdx.x; u < units; u += gridDim.x) {
    const int split = u % s.nsplit;
    int rest = u / s.nsplit;
    const int rgi = rest % s.rg; rest /= s.rg;
    const int kvh = rest % s.Hkv;
    const int b = rest / s.Hkv;
    const int qh0 = kvh * (s.rg * R) + rgi * R;
    // Everything that depends only on the POSITION comes first -- with PREWAIT (stand-alone launch) before the wait on the
    // q/k/v projection: the rotation's sines, and the K/V rows of older tokens (written by earlier steps) of two rounds.
    const uint16_t* row = P.qkv + (size_t)b * qkv_ld;
    float4 s0, s1, c0, c1;   // this lane's 8 sines and cosines, requested together with the position
    s0 = s1 = c0 = c1 = make_float4(0.f, 0.f, 0.f, 0.f);
    if (s.rope_tab) {
      const float4* tb = reinterpret_cast<const float4*>(s.rope_tab + (size_t)b * D + (C == 2 ? 8 * sub : 8 * (sub & 3)));
      s0 = __ldg(tb); s1 = __ldg(tb + 1); c0 = __ldg(tb + D / 8); c1 = __ldg(tb + D / 8 + 1);
    }
    const int p = s.pos[b], ctx = p + 1;
    stamp(9 + (p < -5));
    const int per = (ctx + s.nsplit - 1) / s.nsplit;
    const int t0 = split * per, t1 = min(ctx, t0 + per);
    const size_t rec0 = ((size_t)b * s.Hq + qh0) * s.nsplit;
    // K/V rows are requested two rounds ahead: kn / vn hold the next round, kn2 / vn2 the one after
    uint4 kn[C], vn[C], kn2[C], vn2[C];
    auto request_to = [&](uint4 (&kd)[C], uint4 (&vd)[C], int tb_, bool fresh_ok) {
      const int t_ = tb_ + tg;
      if (t_ < t1) {
        const bool fresh_ = t_ == p;
        if (fresh_ && !fresh_ok) return;
        const size_t off = step_kv_addr(P.kv, b, kvh, t_);
        const uint4* kp = fresh_ ? reinterpret_cast<const uint4*>(row + (size_t)(s.Hq + kvh) * D)
                                 : reinterpret_cast<const uint4*>(P.kc + off);
        const uint4* vp = fresh_ ? reinterpret_cast<const uint4*>(row + (size_t)(s.Hq + s.Hkv + kvh) * D)
                                 : reinterpret_cast<const uint4*>(P.vc + off);
#pragma unroll
        for (int c = 0; c < C; ++c) { kd[c] = ldcg128(kp + sub + 8 * c); vd[c] = ldcg128(vp + sub + 8 * c); }
      }
    };
    const int tb0 = t0 + warp * 4;
    stamp(10);
    if (tb0 < t1) request_to(kn, vn, tb0, false);
    if (tb0 + NW * 4 < t1) request_to(kn2, vn2, tb0 + NW * 4, false);
    float sn[8], cs[8];
    stamp(1);
    if (s.rope_tab) {
      sn[0] = s0.x; sn[1] = s0.y; sn[2] = s0.z; sn[3] = s0.w; sn[4] = s1.x; sn[5] = s1.y; sn[6] = s1.z; sn[7] = s1.w;
      cs[0] = c0.x; cs[1] = c0.y; cs[2] = c0.z; cs[3] = c0.w; cs[4] = c1.x; cs[5] = c1.y; cs[6] = c1.z; cs[7] = c1.w;
    } else {
#pragma unroll
      for (int j = 0; j < 8; ++j)
        rope_sincos((float)p * (1.0f / __ldg(s.freqs + (C == 2 ? 8 * sub : 8 * (sub & 3)) + j)), &sn[j], &cs[j]);
    }
    stamp(3);
    if (PREWAIT) {
      pdl_wait();
      stamp(2);
    }
    uint4 qraw[R][C];
#pragma unroll
    for (int r = 0; r < R; ++r)
#pragma unroll
      for (int c = 0; c < C; ++c)
        qraw[r][c] = ldcg128(reinterpret_cast<const uint4*>(row + (size_t)(qh0 + r) * D) + sub + 8 * c);
    // the token of this step, if it falls into one of the two rounds requested before the projection was complete
    if (tb0 + tg == p && tb0 + tg < t1) request_to(kn, vn, tb0, true);
    if (tb0 + NW * 4 + tg == p && tb0 + NW * 4 + tg < t1) request_to(kn2, vn2, tb0 + NW * 4, true);
    if (t0 < t1) {
      float q[R][E];
#pragma unroll
      for (int r = 0; r < R; ++r) {
#pragma unroll
        for (int c = 0; c < C; ++c) unpack8(qraw[r][c], &q[r][8 * c]);
        rope_row<C>(q[r], sn, cs, sub);
#pragma unroll
        for (int e = 0; e < E; ++e) q[r][e] *= s.scale_log2e;
      }
      stamp(4 + (q[0][0] == 123456.f));
      float m[R], l[R], acc[R][E];
#pragma unroll
      for (int r = 0; r < R; ++r) {
        m[r] = -INFINITY; l[r] = 0.f;
#pragma unroll
        for (int e = 0; e < E; ++e) acc[r][e] = 0.f;
      }
      for (int tb = tb0; tb < t1; tb += NW * 4) {  // warp-uniform trip count; tail lanes idle
        const int t = tb + tg;
        const bool ok = t < t1;
        float kf[E], vf[E];
#pragma unroll
        for (int e = 0; e < E; ++e) { kf[e] = 0.f; vf[e] = 0.f; }
        const bool fresh = ok && t == p;  // the token of this step: not in the cache yet
        uint4 kr[C], vr[C];
#pragma unroll
        for (int c = 0; c < C; ++c) { kr[c] = kn[c]; vr[c] = vn[c]; kn[c] = kn2[c]; vn[c] = vn2[c]; }
        if (tb + 2 * NW * 4 < t1) request_to(kn2, vn2, tb + 2 * NW * 4, true);
        if (ok) {
#pragma unroll
          for (int c = 0; c < C; ++c) { unpack8(kr[c], &kf[8 * c]); unpack8(vr[c], &vf[8 * c]); }
        }
        // the rotation of the fresh key needs its 8-lane group converged (D == 64 exchanges lanes): do it warp-wide
        if (__any_sync(0xffffffffu, fresh)) {
          float kt[E];
#pragma unroll
          for (int e = 0; e < E; ++e) kt[e] = kf[e];
          rope_row<C>(kt, sn, cs, sub);
          if (fresh) {
#pragma unroll
            for (int e = 0; e < E; ++e) kf[e] = kt[e];
            if (rgi == 0) {
              const size_t off = step_kv_addr(P.kv, b, kvh, t);
              uint4* kd = reinterpret_cast<uint4*>(P.kc + off);
              uint4* vd = reinterpret_cast<uint4*>(P.vc + off);
#pragma unroll
              for (int c = 0; c < C; ++c) { kd[sub + 8 * c] = pack8(&kf[8 * c]); vd[sub + 8 * c] = vr[c]; }
            }
          }
        }
#pragma unroll
        for (int r = 0; r < R; ++r) {
          float sc = 0.f;
#pragma unroll
          for (int e = 0; e < E; ++e) sc = fmaf(q[r][e], kf[e], sc);
          sc += __shfl_xor_sync(0xffffffffu, sc, 1);
          sc += __shfl_xor_sync(0xffffffffu, sc, 2);
          sc += __shfl_xor_sync(0xffffffffu, sc, 4);
          if (ok) {
            const float mn = fmaxf(m[r], sc);
            const float corr = exp2f(m[r] - mn), pr = exp2f(sc - mn);
            l[r] = l[r] * corr + pr;
#pragma unroll
            for (int e = 0; e < E; ++e) acc[r][e] = fmaf(pr, vf[e], acc[r][e] * corr);
            m[r] = mn;
          }
        }
      }
      stamp(5 + (l[0] == 123456.f));
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
          float* rec = scratch + ((size_t)warp * R + r) * PW;
#pragma unroll
          for (int c = 0; c < C; ++c)
#pragma unroll
            for (int j = 0; j < 8; ++j) rec[(sub + 8 * c) * 8 + j] = acc[r][8 * c + j];
          if (sub == 0) { rec[D] = m[r]; rec[D + 1] = l[r]; }
        }
      }
    }
    __syncthreads();
    if (CLUSTER && s.nsplit > 1) asm volatile("barrier.cluster.wait.aligned;" ::: "memory");   // the first CTA's mbarrier is set up
    // merge the warps; thread -> (r, d).  An empty split leaves (m, l) = (-inf, 0).
    for (int i = threadIdx.x; i < R * D; i += blockDim.x) {
      const int r = i / D, d = i % D;
      float mn = -INFINITY, lt = 0.f, at = 0.f;
      if (t0 < t1) {
        for (int w = 0; w < NW; ++w) mn = fmaxf(mn, scratch[((size_t)w * R + r) * PW + D]);
        for (int w = 0; w < NW; ++w) {
          const float* rec = scratch + ((size_t)w * R + r) * PW;
          const float cw = rec[D] == -INFINITY ? 0.f : exp2f(rec[D] - mn);
          lt += rec[D + 1] * cw;
          at += rec[d] * cw;
        }
      }
      if (s.nsplit == 1) {
        P.attn_out[((size_t)b * s.Hq + qh0 + r) * D + d] = f2bf(at / lt);
      } else if (CLUSTER) {
        // Splits 1.. push their partial into the slot `split` of the cluster's first CTA (remote shared memory; every
        // store also counts on that CTA's mbarrier); the first CTA keeps its own in slot 0.
        float* rec = scratch + (size_t)NW * R * PW + ((size_t)split * R + r) * PW;
        if (split == 0) {
          rec[d] = at;
          if (d == 0) { rec[D] = mn; rec[D + 1] = lt; }
        } else {
          const uint32_t bar = smem_u32(cl_bar);
          st_async_f32(smem_u32(rec + d), bar, 0, at);
          if (d == 0) { st_async_f32(smem_u32(rec + D), bar, 0, mn); st_async_f32(smem_u32(rec + D + 1), bar, 0, lt); }
        }
      } else {
        float* rec = s.attn_part + (rec0 + (size_t)r * s.nsplit + split) * PW;
        rec[d] = at;
        if (d == 0) { rec[D] = mn; rec[D + 1] = lt; }
      }
    }
    stamp(6);
    if (CLUSTER && s.nsplit > 1) {
      if (split == 0) {
        __syncthreads();        // slot 0 is written
        mbar_wait(cl_bar, 0);   // every other split's R * (D + 2) values have landed
        const float* part = scratch + (size_t)NW * R * PW;
        for (int i = threadIdx.x; i < R * D; i += blockDim.x) {
          const int r = i / D, d = i % D;
          float mn = -INFINITY;
          for (int sp = 0; sp < s.nsplit; ++sp) mn = fmaxf(mn, part[((size_t)sp * R + r) * PW + D]);
          float lt = 0.f, at = 0.f;
          for (int sp = 0; sp < s.nsplit; ++sp) {  // split order: deterministic
            const float* rec = part + ((size_t)sp * R + r) * PW;
            const float cw = rec[D] == -INFINITY ? 0.f : exp2f(rec[D] - mn);
            lt += rec[D + 1] * cw;
            at += rec[d] * cw;
          }
          P.attn_out[((size_t)b * s.Hq + qh0 + r) * D + d] = f2bf(at / lt);
        }
      }
    }
    if (!CLUSTER && s.nsplit > 1) {
      __threadfence();
      __syncthreads();
      const int cidx = (b * s.Hkv + kvh) * s.rg + rgi;
      if (threadIdx.x == 0) *s_flag = (atomicAdd(&s.attn_cnt[cidx], 1u) == (unsigned)s.nsplit - 1);
      __syncthreads();
      if (*s_flag) {
        __threadfence();
        for (int i = threadIdx.x; i < R * D; i += blockDim.x) {
          const int r = i / D, d = i % D;
          const float* rec = s.attn_part + (rec0 + (size_t)r * s.nsplit) * PW;
          float mn = -INFINITY;
          for (int sp = 0; sp < s.nsplit; ++sp) mn = fmaxf(mn, ldcg_f32(rec + (size_t)sp * PW + D));
          float lt = 0.f, at = 0.f;
          for (int sp = 0; sp < s.nsplit; ++sp) {
            const float ms = ldcg_f32(rec + (size_t)sp * PW + D);
            const float cw = ms == -INFINITY ? 0.f : exp2f(ms - mn);
            lt += ldcg_f32(rec + (size_t)sp * PW + D + 1) * cw;
            at += ldcg_f32(rec + (size_t)sp * PW + d) * cw;
          }
          P.attn_out[((size_t)b * s.Hq + qh0 + r) * D + d] = f2bf(at / lt);
        }
        if (threadIdx.x == 0) s.attn_cnt[cidx] = 0;
      }
    }
    __syncthreads();  // scratch and s_flag are reused by the next unit
  }
}

// Out-of-line form for the one-launch step (four instantiations in one kernel).  The stand-alone launch inlines the body:
// out of line, the __grid_constant__ parameter blocks are reached through generic pointers, and the chain
// page table? -> strides -> cache pointers in front of the first K/V request cost ~0.8 us per launch (LD.E from the
// parameter bank, three dependent round trips) where the inlined code reads constant-bank operands.
template <int D, int R, bool CLUSTER, int NW>
__device__ __noinline__ void attn_phase(const StepArgs& s, const StepPhase& P, float* scratch, int* s_flag) {
  attn_phase_impl<D, R, CLUSTER, NW, false>(s, P, scratch, s_flag);
}

// a4: h[b, :] = dequantize(embed[tokens[b]]) -- same arithmetic as dequantize_kernel (pie_misc.cu): one packed word
// per thread, w_hat = bf16(bf16(q * scale) + bias).
__device__ __noinline__ void embed_phase(const StepArgs& s, int bits) {
  const int per = 32 / bits, wpr = s.hidden / per, gpr = s.hidden / kGroup;
  const uint32_t mask = (1u << bits) - 1u;
  for (int idx = blockIdx.x * blockDim.x + threadIdx.x; idx < s.B * wpr; idx += gridDim.x * blockDim.x) {
    const int b = idx / wpr, wi = idx % wpr;
    int r = s.tokens[b];
    r = r < 0 ? 0 : (r >= s.table_rows ? s.table_rows - 1 : r);
    const uint32_t w = __ldg(s.e_wq + (size_t)r * wpr + wi);
    const int grp = wi * per / kGroup;
    const float sc = bf16lo((uint32_t)__ldg(s.e_s + (size_t)r * gpr + grp));
    const float bi = bf16lo((uint32_t)__ldg(s.e_b + (size_t)r * gpr + grp));
    uint16_t o[8];
    for (int e = 0; e < per; ++e) o[e] = f2bf(bfround((float)((w >> (e * bits)) & mask) * sc) + bi);
    uint16_t* dst = s.h + (size_t)b * s.hidden + (size_t)wi * per;
    if (per == 8) *reinterpret_cast<uint4*>(dst) = *reinterpret_cast<const uint4*>(o);
    else *reinterpret_cast<uint2*>(dst) = *reinterpret_cast<const uint2*>(o);
  }
}

// greedy sampler + cache.offset += 1 (argmax_kernel, pie_misc.cu): first maximum; CTA b takes batch row b.
__device__ __noinline__ void argmax_phase(const StepArgs& s, float* sv, int* si) {
  for (int b = blockIdx.x; b < s.B; b += gridDim.x) {
    const uint16_t* row = s.logits + (size_t)b * s.vocab;
    float best = -INFINITY;
    int bi = 0x7fffffff;
    for (int i = threadIdx.x; i < s.vocab; i += blockDim.x) {
      const float v = bf16lo((uint32_t)ldcg16(row + i));
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
      const int nw = blockDim.x >> 5;
      best = threadIdx.x < nw ? sv[threadIdx.x] : -INFINITY;
      bi = threadIdx.x < nw ? si[threadIdx.x] : 0x7fffffff;
      for (int o = 16; o > 0; o >>= 1) {
        const float ov = __shfl_xor_sync(0xffffffffu, best, o);
        const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
        if (ov > best || (ov == best && oi < bi)) { best = ov; bi = oi; }
      }
      if (threadIdx.x == 0) {
        const int tok = bi == 0x7fffffff ? 0 : bi;
        s.tokens[b] = tok;
        const int np = s.pos[b] + 1;
        s.pos[b] = np;
        if (s.seq && np < s.seq_ld) s.seq[(size_t)b * s.seq_ld + np] = tok;
      }
    }
    __syncthreads();
  }
}

// Stand-alone form of the attention phase: one CTA per unit (pie_decode_attn_fused).
constexpr int kAttnWarps = 8;  // stand-alone launch: 8 warps and up to 255 registers (no spills, K/V double-buffered)
template <int D, int R, bool CLUSTER>
__global__ void __launch_bounds__(kAttnWarps * 32, 1) attn_fused_kernel(const __grid_constant__ StepArgs s,
                                                                       const __grid_constant__ StepPhase P) {
  __shared__ float scratch[(kAttnWarps + 8) * R * (D + 2)];   // per-warp partials + (first CTA of a cluster) per-split partials
  __shared__ uint64_t cl_bar;
  __shared__ int s_flag;
  if (s.trace && threadIdx.x == 0) s.trace[blockIdx.x * 16] = globaltimer_ns();
  pdl_launch_dependents();
  attn_phase_impl<D, R, CLUSTER, kAttnWarps, true>(s, P, scratch, &s_flag, s.trace ? s.trace + blockIdx.x * 16 : nullptr, &cl_bar);
  if (s.trace && threadIdx.x == 0) s.trace[blockIdx.x * 16 + 7] = globaltimer_ns();
}

// ---- the one-launch step: TMA-fed GEMV phases (the arithmetic of pie_qmv2.cu, one row, fp16 staging) ---------------
// Shared memory (base aligned to 1024 B): [weight stages 16 x 2 x 4096][scale/bias stages 16 x 2 x 512][mbarriers]
// [edge tile ids][edge values 16 x 2 x 16 f32][reduction scratch 1 KB][per-phase: start-value quads, sum_x pieces, x]
// The attention phase's scratch aliases the per-phase region (x is restaged by every matrix).
constexpr int kS2EdgeValOff = kQ2RedOff + 1024;
constexpr int kS2VarOff = kS2EdgeValOff + kQ2Warps * 2 * 16 * 4;
__host__ __device__ inline size_t s2_var_bytes(int K) {
  return (size_t)(q2_kpad(K) / kGroup) * 16 + (((size_t)3 * q2_gpitch(K) * 2 + 15) & ~(size_t)15) + (size_t)q2_xrow_bytes(K);
}

// this CTA's / warp's run of units of a matrix (same partition as qmv2_kernel)
struct S2Run { int ub, ue, upt; };
__device__ __forceinline__ S2Run s2_run_of(const Qmv2Args& a, int warp) {
  const int G = gridDim.x, c = blockIdx.x;
  const int ntiles = a.ntiles, upt = a.upt;
  // (32-bit arithmetic: c * ntiles < 2^31 for any matrix the plan accepts)
  const int tile_lo = (int)((unsigned)c * (unsigned)ntiles / (unsigned)G);
  const int tile_hi = (int)((unsigned)(c + 1) * (unsigned)ntiles / (unsigned)G);
  const int cu = (tile_hi - tile_lo) * upt;
  S2Run r;
  r.upt = upt;
  r.ub = tile_lo * upt + (int)((unsigned)warp * (unsigned)cu / kQ2Warps);
  r.ue = tile_lo * upt + (int)((unsigned)(warp + 1) * (unsigned)cu / kQ2Warps);
  return r;
}

__global__ void __launch_bounds__(kQ2Warps * 32, 1) decode_step2_kernel(const __grid_constant__ StepArgs s) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ int s_flag;
  // The descriptors of the current and the next matrix, copied from the plan one matrix ahead: every scalar the x
  // staging and the main loop read comes from shared memory (from global memory each is a dependent L2 round trip in
  // front of the matrix: measured 4.6 us of staging).  TMA itself takes the tensor maps from the plan in global memory.
  __shared__ __align__(64) StepPhase ph[2];
  static_assert(sizeof(StepPhase) % 16 == 0, "phase descriptors are copied 16 bytes at a time");
  constexpr int PHV = sizeof(StepPhase) / 16;
  auto fetch_phase = [&](int q) {
    if (q < s.nphases)
      for (int i = threadIdx.x; i < PHV; i += blockDim.x)
        reinterpret_cast<uint4*>(&ph[q & 1])[i] = __ldg(reinterpret_cast<const uint4*>(s.phases + q) + i);
  };
  fetch_phase(0);
  fetch_phase(1);
  __syncthreads();
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  const int tid = threadIdx.x, lane = tid & 31;
  const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
  const int g = lane >> 2, t = lane & 3;
  const uint32_t smem_s = smem_u32(smem);
  const uint32_t wst = smem_s + (uint32_t)(warp * kQ2Stages) * kQ2WBytes;
  const uint32_t sbst = smem_s + kQ2SbOff + (uint32_t)(warp * kQ2Stages) * kQ2SbBytes;
  const uint32_t bar0 = smem_s + kQ2BarOff + (uint32_t)(warp * kQ2Stages) * 8;
  int* edge_tile = reinterpret_cast<int*>(smem + kQ2EdgeOff);
  float* edge_val = reinterpret_cast<float*>(smem + kS2EdgeValOff);   // [warp][2][16]
  float* red = reinterpret_cast<float*>(smem + kQ2RedOff);
  uint8_t* var = smem + kS2VarOff;
  float* scratch = reinterpret_cast<float*>(var);
  const uint64_t pol = l2_policy_evict_first();
  const bool tracing = s.trace && blockIdx.x == 0 && tid == 0;
  if (tracing) s.trace[0] = globaltimer_ns();
  // grid barriers: arrivals accumulate in bar[2] (zero at launch: the last CTA to leave the kernel resets it)
  unsigned int bar_target = 0;
  auto gbar = [&]() { bar_target += gridDim.x; grid_barrier_flat(s.bar + 2, bar_target); };

  // (tile, k chunk) of matrix `a` -> TMA copies into stage st of this warp's ring
  // d: the matrix's scalars (shared copy, or the plan itself for a matrix further ahead); a: the plan entry (maps)
  auto issue = [&](const Qmv2Args& d, const Qmv2Args& a, int tile, int kc, int st) {
    const uint32_t bar = bar0 + st * 8, wd = wst + st * kQ2WBytes, sd = sbst + st * kQ2SbBytes;
    q2_issue_unit(d, a, tile, kc, wd, sd, bar, pol);
  };
  int cur_phase = 0;  // the matrix being consumed: ph[] holds cur_phase and cur_phase + 1
  auto desc_of = [&](int q) -> const Qmv2Args& {
    return (q == cur_phase || q == cur_phase + 1) ? ph[q & 1].q : s.phases[q].q;
  };
  // Producer cursor: the next unit to request, possibly of a LATER matrix than the one being consumed -- the ring keeps
  // streaming across the grid barriers.  pp = its phase, [p_u, p_ue) the rest of this warp's run there.
  int pp = 0, p_u, p_ue, p_tile, p_kc, p_upt;
  auto cursor_to_phase = [&](int q) {
    pp = q;
    if (q < s.nphases) {
      const S2Run r = s2_run_of(desc_of(q), warp);
      p_u = r.ub; p_ue = r.ue; p_upt = r.upt;
      p_tile = r.ub / r.upt; p_kc = r.ub - p_tile * r.upt;
    } else {
      p_u = p_ue = 0; p_tile = p_kc = 0; p_upt = 1;
    }
  };
  unsigned int n_issued = 0, n_used = 0;   // units requested / consumed by this warp since the kernel began
  // Keep the ring full (two units in flight or waiting), but let at most `s.pf_cross` units of a matrix LATER than the
  // one being consumed (`cur`) go out: requests that are in flight while the grid barrier and the x staging make their
  // round trips queue in front of them (measured: 21.8 MB in flight -> 6 us barriers, 4.7 us staging).
  auto top_up = [&](int cur) {   // warp-uniform; one elected lane issues
    while (n_issued - n_used < 2u) {
      while (pp < s.nphases && p_u >= p_ue) cursor_to_phase(pp + 1);
      if (pp >= s.nphases) return;
      if (pp > cur) {
        // units of later matrices already out = everything issued beyond the current matrix's own units
        if ((int)(n_issued - n_used) >= s.pf_cross) return;
      }
      if (elect_one()) issue(desc_of(pp), s.phases[pp].q, p_tile, p_kc, (int)(n_issued & 1u));
      ++n_issued;
      ++p_u;
      if (++p_kc == p_upt) { p_kc = 0; ++p_tile; }
    }
  };

  if (lane == 0) {
    mbar_init_s(bar0, 1);
    mbar_init_s(bar0 + 8, 1);
    fence_mbar_init();
    fence_proxy_async();
  }
  __syncwarp();
  cursor_to_phase(0);
  top_up(s.nphases);   // the very first units: nothing is waiting for a round trip yet

  for (int p = 0; p < s.nphases; ++p) {
    cur_phase = p;
    const StepPhase& P = ph[p & 1];
    const Qmv2Args& a = P.q;
    // ---- produce the matrix's input (embedding / attention), then wait for the whole grid ------------------------
    if (P.pre == PRE_EMBED) {
      embed_phase(s, 4);
    } else if (P.pre == PRE_ATTN) {
      gbar();
      if (s.D == 128) {
        if (s.R == 2) attn_phase<128, 2, false, kQ2Warps>(s, P, scratch, &s_flag);
        else attn_phase<128, 1, false, kQ2Warps>(s, P, scratch, &s_flag);
      } else {
        if (s.R == 2) attn_phase<64, 2, false, kQ2Warps>(s, P, scratch, &s_flag);
        else attn_phase<64, 1, false, kQ2Warps>(s, P, scratch, &s_flag);
      }
    }
    gbar();
    if (tracing) s.trace[2 + 3 * p] = globaltimer_ns();
    if (lane == 0) { edge_tile[warp * 2] = -1; edge_tile[warp * 2 + 1] = -1; }
    q2_stage_x<true, true>(a, var, red);   // (the staging's own stamps belong to the stand-alone launch's 16-stamp slots)
    __syncthreads();
    if (tracing) s.trace[3 + 3 * p] = globaltimer_ns();
    top_up(p);

    // ---- this warp's units of matrix p ---------------------------------------------------------------------------
    const S2Run run = s2_run_of(a, warp);
    const int ub = run.ub, ue = run.ue, upt = run.upt;
    const int dual = a.dual, epi = a.epi;
    const int gp = q2_gpitch(a.K);
    const uint32_t nsx_s = smem_u32(var);
    const uint32_t pcs_s = nsx_s + (uint32_t)(q2_kpad(a.K) / kGroup) * 16;
    const uint32_t xs_s = pcs_s + (uint32_t)((3 * gp * 2 + 15) & ~15);
    const uint32_t x_lane = xs_s + t * 32;
    const uint32_t pcs_lane = pcs_s + 2 * t * 2;
    const uint32_t hs = dual ? 2048u : 1024u, cs = dual ? 1024u : 2048u;
    const uint32_t w_lane = g * 128 + t * 8, gx5 = (g & 3) << 5;
    const uint32_t lmask = a.lift_mask, lmask2 = a.lift_mask2, lmagic = a.lift_magic;
    const float rs = red[208];

    auto seg_of = [&](int tile, int& si, int& nbase) {
      si = 0;
      if (dual) { nbase = tile * 8; return; }
      if (a.nseg > 1 && tile >= a.tile0[1]) si = 1;
      if (a.nseg > 2 && tile >= a.tile0[2]) si = 2;
      nbase = (tile - a.tile0[si]) * 16;
    };
    // finished tile: lanes t == 0 hold row g (v0) and row g + 8 (v1)
    auto epilogue = [&](int tile, float v0, float v1) {
      if (t != 0) return;
      int si, nbase;
      seg_of(tile, si, nbase);
      if (dual) {   // SwiGLU: row g = gate, row g + 8 = up of the same output column
        const int n = nbase + g;
        if (n < a.N[0]) {
          const float gt = bfround(v0), up = bfround(v1);
          const float sg = bfround(1.f / (1.f + expf(-gt)));
          a.out[n] = f2bf(bfround(gt * sg) * up);
        }
      } else {
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          const int n = nbase + g + 8 * h;
          if (n < a.N[si]) {
            float r = bfround(h ? v1 : v0);
            if (epi == EPI_RESADD) r = bf16lo((uint32_t)ldcg16(a.residual + a.out_off[si] + n)) + r;
            a.out[a.out_off[si] + n] = f2bf(r);
          }
        }
      }
    };

    float ys0 = 0.f, ys2 = 0.f, yb[4] = {0.f, 0.f, 0.f, 0.f};
    int n_edges = 0, run_start_kc = 0;
    int tile = ub / upt, kc = ub - tile * upt;
    for (int u = ub; u < ue; ++u) {
      const int st = n_used & 1;
      if (u == ub || kc == 0) { run_start_kc = kc; ys0 = ys2 = 0.f; yb[0] = yb[1] = yb[2] = yb[3] = 0.f; }
      const uint32_t tok = mbar_wait_s(bar0 + st * 8, (n_used >> 1) & 1);
      {
        const uint32_t wb = wst + st * kQ2WBytes + w_lane + tok, sp = sbst + st * kQ2SbBytes + tok;
        const uint32_t xp = x_lane + (uint32_t)kc * (8 * 128) + tok;
        const uint32_t np = nsx_s + (uint32_t)kc * 8 * 16 + tok;
#pragma unroll
        for (int cch = 0; cch < 2; ++cch) {
          const uint2 sA = lds64_t(sp + g * 16 + cch * 8, 0), sB = lds64_t(sp + 128 + g * 16 + cch * 8, 0);
          const uint32_t sAw[2] = {sA.x, sA.y}, sBw[2] = {sB.x, sB.y};
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const uint32_t wo = wb + cch * cs + ((uint32_t)(j << 5) ^ gx5);
            const uint2 w0 = lds64_t(wo, 0), w1 = lds64_t(wo + hs, 0);
            const uint4 x0 = lds128_t(xp + (cch * 4 + j) * 128, 0), x1 = lds128_t(xp + (cch * 4 + j) * 128 + 16, 0);
            const uint4 nq = lds128_t(np + (cch * 4 + j) * 16, 0);
            const uint32_t xb[8] = {x0.x, x0.y, x0.z, x0.w, x1.x, x1.y, x1.z, x1.w};
            float acc[4];
#pragma unroll
            for (int m = 0; m < 4; ++m) {
              const int sh = (m >> 1) * 8;
              const uint32_t mk = (m & 1) ? lmask2 : lmask;
              const uint32_t a0 = lift4b(w0.x, sh, mk, lmagic), a1 = lift4b(w1.x, sh, mk, lmagic),
                             a2 = lift4b(w0.y, sh, mk, lmagic), a3 = lift4b(w1.y, sh, mk, lmagic);
              if (m == 0) mma_16816<true>(acc, a0, a1, a2, a3, xb[0], xb[1], __uint_as_float(nq.x), __uint_as_float(nq.y),
                                          __uint_as_float(nq.z), __uint_as_float(nq.w));
              else mma_16816<true>(acc, a0, a1, a2, a3, xb[2 * m], xb[2 * m + 1], acc[0], acc[1], acc[2], acc[3]);
            }
            ys0 = fmaf(j & 1 ? bf16hi(sAw[j >> 1]) : bf16lo(sAw[j >> 1]), acc[0], ys0);
            ys2 = fmaf(j & 1 ? bf16hi(sBw[j >> 1]) : bf16lo(sBw[j >> 1]), acc[2], ys2);
          }
        }
        const uint32_t ba0 = lds32_t(sp + 256 + g * 16 + t * 4, 0), ba1 = lds32_t(sp + 384 + g * 16 + t * 4, 0);
        const uint32_t pb = pcs_lane + (uint32_t)kc * 16 + tok;
#pragma unroll
        for (int q = 0; q < 3; ++q) mma_bf16_1688(yb, ba0, ba1, lds32_t(pb + (uint32_t)q * gp * 2, 0), yb);
      }
      ++n_used;
      __syncwarp();
      top_up(p);
      const int kc_done = kc, tile_done = tile;
      if (++kc == upt) { kc = 0; ++tile; }
      if (kc_done != upt - 1 && u != ue - 1) continue;
      // ---- this warp's share of the tile is complete: lanes t == 0 hold rows g (v0) and g + 8 (v1) -----------------
      const float v0 = fmaf(ys0, rs, yb[0]), v1 = fmaf(ys2, rs, yb[2]);
      if (run_start_kc == 0 && kc_done == upt - 1) {
        epilogue(tile_done, v0, v1);
      } else {
        if (t == 0) { edge_val[(warp * 2 + n_edges) * 16 + g] = v0; edge_val[(warp * 2 + n_edges) * 16 + 8 + g] = v1; }
        if (lane == 0) edge_tile[warp * 2 + n_edges] = tile_done;
        ++n_edges;
      }
    }
    // the warp whose run holds a cut tile's LAST unit adds the tile's shares in warp (= k) order and finishes it
    int fin_tile = -1;
    {
      const int first_tile = ub / upt;
      if (ue > ub && ub > first_tile * upt && (first_tile + 1) * upt <= ue) fin_tile = first_tile;
    }
    __syncthreads();
    if (fin_tile >= 0) {
      float v0 = 0.f, v1 = 0.f;
      for (int w2 = 0; w2 <= warp; ++w2) {
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          if (edge_tile[w2 * 2 + e] != fin_tile) continue;
          v0 += edge_val[(w2 * 2 + e) * 16 + g];
          v1 += edge_val[(w2 * 2 + e) * 16 + 8 + g];
        }
      }
      epilogue(fin_tile, v0, v1);
    }
    if (tracing) s.trace[4 + 3 * p] = globaltimer_ns();
    // every warp of this CTA is past matrix p (the barrier above): its descriptor slot takes matrix p + 2.  A warp's
    // producer cursor may already sit in matrix p + 1 (its copy is in the other slot) but not beyond...
    __syncthreads();
    fetch_phase(p + 2);
  }
  gbar();
  if (tracing) s.trace[1] = globaltimer_ns();
  argmax_phase(s, scratch, reinterpret_cast<int*>(scratch + 32));
  // every CTA is past the last barrier once it gets here: the last one to leave zeroes the counters for the next launch
  if (tid == 0 && atomicAdd(s.bar + 3, 1u) == gridDim.x - 1) {
    s.bar[2] = 0;
    s.bar[3] = 0;
    __threadfence();
  }
}

static cudaError_t launch_step2(const StepArgs& s, int grid, size_t smem, cudaStream_t st) {
  cudaError_t e = cudaFuncSetAttribute(decode_step2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kQmvMaxSmem - 4096);
  if (e != cudaSuccess) return e;
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)grid, 1, 1);
  cfg.blockDim = dim3(kQ2Warps * 32, 1, 1);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeCooperative;  // every CTA must be resident: the kernel has grid barriers
  attr[0].val.cooperative = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  count_launch();
  return cudaLaunchKernelEx(&cfg, decode_step2_kernel, s);
}

}  // namespace pie

using namespace pie;

extern "C" {

size_t pie_decode_step_plan_bytes(int n_layers) {
  if (n_layers <= 0) return 0;
  return sizeof(StepHeader) + (size_t)(4 * n_layers + 1) * sizeof(StepPhase);
}

int pie_decode_step_pick_splits(int B, int Hq, int Hkv, int max_context) {
  if (B <= 0 || Hq <= 0 || Hkv <= 0 || Hq % Hkv) return 1;
  const int n_rep = Hq / Hkv;
  const int R = n_rep % 2 == 0 ? 2 : 1;  // query heads per unit (see pie_decode_step_plan)
  const int rows = B * Hkv * (n_rep / R);
  int ns = num_sms() / rows;                 // one unit per CTA when the context allows it
  const int cap = (max_context + 63) / 64;   // a unit covers >= 64 tokens (4 per warp) of the longest context
  if (ns > cap) ns = cap;
  if (ns > 64) ns = 64;
  return ns < 1 ? 1 : ns;
}

size_t pie_decode_step_attn_ws_bytes(int B, int Hq, int D, int nsplit) {
  return pie_decode_attn_workspace_bytes(B, Hq, D, nsplit);
}

int pie_decode_step_plan(const pie_step_model* m, const pie_step_layer* layers, const pie_step_io* io, void* plan_host,
                         size_t plan_bytes) {
  if (!m || !layers || !io || !plan_host) return PIE_ENULL;
  if (m->n_layers <= 0 || io->B != 1) return PIE_EUNSUPPORTED;           // batch 1 (the latency-bound case)
  if (m->group_size != 64 || m->bits != 4) return PIE_EUNSUPPORTED;       // 4-bit weights
  if (m->D != 64 && m->D != 128) return PIE_EUNSUPPORTED;
  if (m->Hq <= 0 || m->Hkv <= 0 || m->Hq % m->Hkv || m->hidden % 64 || m->ffn % 64) return PIE_EINVAL_SHAPE;
  if (plan_bytes < pie_decode_step_plan_bytes(m->n_layers)) return PIE_EWORKSPACE;
  if (!io->tokens || !io->positions || !io->h || !io->qkv || !io->attn || !io->act || !io->logits) return PIE_ENULL;
  if (!m->rope_freqs || !m->final_norm || !m->embed.wq || !m->embed.scales || !m->embed.biases) return PIE_ENULL;
  const int B = io->B, hid = m->hidden, qd = m->Hq * m->D, qkvd = (m->Hq + 2 * m->Hkv) * m->D;
  const int nsplit = io->nsplit < 1 ? 1 : io->nsplit;
  if (nsplit > 1 && (!io->attn_ws || io->attn_ws_bytes < pie_decode_attn_workspace_bytes(B, m->Hq, m->D, nsplit)))
    return PIE_EWORKSPACE;
  memset(plan_host, 0, plan_bytes);
  StepHeader* H = reinterpret_cast<StepHeader*>(plan_host);
  StepPhase* ph = reinterpret_cast<StepPhase*>(H + 1);
  size_t smem = 0;
  int np = 0;
  int max_k = 0;
  auto add = [&](void* out, int out_ld, const void* x, int x_ld, int K, const pie_qmv_seg* segs, int nseg,
                 const void* norm, int epi, const void* residual, int pre) -> int {
    StepPhase& P = ph[np];
    if (!qmv2_supported(1, K, m->bits, m->group_size, segs, nseg, epi == PIE_EPI_SILUMUL)) return PIE_EUNSUPPORTED;
    const int st = build_qmv2_args(P.q, out, out_ld, x, x_ld, 1, K, segs, nseg, norm, m->rms_eps, epi, residual);
    if (st != PIE_OK) return st;
    // fp16 staging throughout the step (one arithmetic path in the kernel)
    P.q.f16 = 1; P.q.lift_magic = 0x64006400u;
    if (K > max_k) max_k = K;
    P.pre = pre;
    ++np;
    return PIE_OK;
  };
  for (int i = 0; i < m->n_layers; ++i) {
    const pie_step_layer& L = layers[i];
    if (!L.kcache || !L.vcache || !L.input_norm || !L.post_norm) return PIE_ENULL;
    if (L.kv.tok_stride < m->D || L.kv.tok_stride % 8 || L.kv.head_stride % 8 || L.kv.batch_stride % 8)
      return PIE_EINVAL_SHAPE;
    if (L.kv.page_table && (L.kv.page_size <= 0 || L.kv.pt_stride <= 0)) return PIE_EINVAL_SHAPE;
    const pie_qmv_seg qkv[3] = {L.q, L.k, L.v};
    const pie_qmv_seg gu[2] = {L.gate, L.up};
    if (L.q.N != qd || L.k.N != m->Hkv * m->D || L.v.N != m->Hkv * m->D || L.o.N != hid || L.down.N != hid ||
        L.gate.N != m->ffn || L.up.N != m->ffn) return PIE_EINVAL_SHAPE;
    int st = add(io->qkv, qkvd, io->h, hid, hid, qkv, 3, L.input_norm, PIE_EPI_STORE, nullptr,
                 i == 0 ? PRE_EMBED : PRE_BARRIER);
    if (st != PIE_OK) return st;
    st = add(io->h, hid, io->attn, qd, qd, &L.o, 1, nullptr, PIE_EPI_RESADD, io->h, PRE_ATTN);
    if (st != PIE_OK) return st;
    StepPhase& A = ph[np - 1];
    A.qkv = (const uint16_t*)io->qkv; A.attn_out = (uint16_t*)io->attn;
    A.kc = (uint16_t*)L.kcache; A.vc = (uint16_t*)L.vcache;
    A.kv.batch_stride = L.kv.batch_stride; A.kv.head_stride = L.kv.head_stride; A.kv.tok_stride = L.kv.tok_stride;
    A.kv.page_table = L.kv.page_table; A.kv.pt_stride = L.kv.pt_stride; A.kv.page_size = L.kv.page_size;
    st = add(io->act, m->ffn, io->h, hid, hid, gu, 2, L.post_norm, PIE_EPI_SILUMUL, nullptr, PRE_BARRIER);
    if (st != PIE_OK) return st;
    st = add(io->h, hid, io->act, m->ffn, m->ffn, &L.down, 1, nullptr, PIE_EPI_RESADD, io->h, PRE_BARRIER);
    if (st != PIE_OK) return st;
  }
  if (m->head.N != m->vocab) return PIE_EINVAL_SHAPE;
  int st = add(io->logits, m->vocab, io->h, hid, hid, &m->head, 1, m->final_norm, PIE_EPI_STORE, nullptr, PRE_BARRIER);
  if (st != PIE_OK) return st;
  const int n_rep = m->Hq / m->Hkv;
  // two query heads per unit: with 512 threads per CTA (128 registers each) four heads' fp32 accumulators spill
  const int R = n_rep % 2 == 0 ? 2 : 1;
  // per-phase region: x, its start values and pieces for the widest K -- or the attention phase's scratch
  // ((16 warps + 1) x R x (D + 2) floats), whichever is larger
  size_t var = s2_var_bytes(max_k);
  const size_t attn_need = (size_t)(kQ2Warps + 1) * R * (m->D + 2) * sizeof(float);
  if (attn_need > var) var = attn_need;
  smem = 1024 + kS2VarOff + var;
  if (smem > (size_t)kQmvMaxSmem - 4096) return PIE_EUNSUPPORTED;  // the kernel also holds two descriptors statically
  H->magic = kStepMagic; H->bits = (uint32_t)m->bits; H->grid = (uint32_t)num_sms(); H->smem = (uint32_t)smem;
  H->onecol = B == 1; H->nphases = (uint32_t)np; H->plan_bytes = (uint32_t)plan_bytes;
  StepArgs& s = H->args;
  s.nphases = np; s.B = B; s.Hq = m->Hq; s.Hkv = m->Hkv; s.D = m->D; s.R = R; s.rg = n_rep / R; s.nsplit = nsplit;
  s.scale_log2e = m->attn_scale * 1.4426950408889634f;
  s.freqs = m->rope_freqs; s.pos = io->positions;
  const size_t cnt_bytes = (((size_t)B * m->Hq * sizeof(unsigned)) + 255) & ~(size_t)255;
  s.attn_cnt = reinterpret_cast<unsigned int*>(io->attn_ws);
  s.attn_part = io->attn_ws ? reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(io->attn_ws) + cnt_bytes) : nullptr;
  s.e_wq = m->embed.wq; s.e_s = (const uint16_t*)m->embed.scales; s.e_b = (const uint16_t*)m->embed.biases;
  s.tokens = io->tokens; s.table_rows = m->embed.N; s.hidden = hid; s.h = (uint16_t*)io->h;
  s.trace = io->trace;
  s.trace_stage = io->trace ? io->trace + 2 + 3 * (4 * m->n_layers + 1) + 2 : nullptr;  // behind the per-matrix stamps
  { const char* e = getenv("PIE_STEP_PF"); s.pf_cross = e && *e ? atoi(e) : 1; if (s.pf_cross < 0) s.pf_cross = 0; if (s.pf_cross > 2) s.pf_cross = 2; }
  s.logits = (const uint16_t*)io->logits; s.vocab = m->vocab; s.seq = io->seq; s.seq_ld = io->seq_ld;
  return PIE_OK;
}

}  // extern "C"
namespace pie {
// sin / cos of the rotary angles of positions[b], once per step instead of once per layer and lane
__global__ void rope_table_kernel(float* table, const int32_t* pos, const float* freqs, int half) {
  const int b = blockIdx.x;
  const float p = (float)pos[b];
  for (int j = threadIdx.x; j < half; j += blockDim.x) {
    float sn, cs;
    rope_sincos(p * (1.0f / freqs[j]), &sn, &cs);
    table[(size_t)b * 2 * half + j] = sn;
    table[(size_t)b * 2 * half + half + j] = cs;
  }
}

}  // namespace pie
extern "C" {

int pie_rope_table(float* table, const int32_t* positions, const float* freqs, int B, int half_dims, void* stream) {
  if (B == 0) return PIE_OK;
  if (!table || !positions || !freqs) return PIE_ENULL;
  if (B < 0 || half_dims <= 0) return PIE_EINVAL_SHAPE;
  count_launch();
  rope_table_kernel<<<B, 64, 0, (cudaStream_t)stream>>>(table, positions, freqs, half_dims);
  return status_of(cudaGetLastError());
}

int pie_decode_attn_fused(void* out, const void* qkv, void* kcache, void* vcache, const pie_kv_layout* kv,
                          const int32_t* positions, const float* freqs, const float* rope_table, int B, int Hq, int Hkv,
                          int D, float scale, int nsplit, void* workspace, size_t workspace_bytes, int pdl, void* stream) {
  if (B == 0) return PIE_OK;
  if (!out || !qkv || !kcache || !vcache || !kv || !positions || !freqs) return PIE_ENULL;
  if (B < 0 || Hq <= 0 || Hkv <= 0 || Hq % Hkv) return PIE_EINVAL_SHAPE;
  if (D != 64 && D != 128) return PIE_EUNSUPPORTED;
  if (kv->tok_stride < D || kv->tok_stride % 8 || kv->head_stride % 8 || kv->batch_stride % 8) return PIE_EINVAL_SHAPE;
  if (kv->page_table && (kv->page_size <= 0 || kv->pt_stride <= 0)) return PIE_EINVAL_SHAPE;
  if ((reinterpret_cast<uintptr_t>(qkv) | reinterpret_cast<uintptr_t>(kcache) | reinterpret_cast<uintptr_t>(vcache)) & 15)
    return PIE_EALIGN;
  if (nsplit <= 0) nsplit = 1;
  if (nsplit > 1 && (!workspace || workspace_bytes < pie_decode_attn_workspace_bytes(B, Hq, D, nsplit)))
    return PIE_EWORKSPACE;
  const int n_rep = Hq / Hkv, R = n_rep % 2 == 0 ? 2 : 1;
  StepArgs s;
  memset(&s, 0, sizeof(s));
  StepPhase P;
  memset(&P, 0, sizeof(P));
  s.B = B; s.Hq = Hq; s.Hkv = Hkv; s.D = D; s.R = R; s.rg = n_rep / R; s.nsplit = nsplit;
  s.scale_log2e = scale * 1.4426950408889634f;
  s.freqs = freqs; s.pos = const_cast<int32_t*>(positions); s.rope_tab = rope_table;
  const size_t cnt_bytes = (((size_t)B * Hq * sizeof(unsigned)) + 255) & ~(size_t)255;
  s.attn_cnt = reinterpret_cast<unsigned int*>(workspace);
  s.attn_part = workspace ? reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(workspace) + cnt_bytes) : nullptr;
  P.qkv = (const uint16_t*)qkv; P.attn_out = (uint16_t*)out; P.kc = (uint16_t*)kcache; P.vc = (uint16_t*)vcache;
  P.kv.batch_stride = kv->batch_stride; P.kv.head_stride = kv->head_stride; P.kv.tok_stride = kv->tok_stride;
  P.kv.page_table = kv->page_table; P.kv.pt_stride = kv->pt_stride; P.kv.page_size = kv->page_size;
  s.trace = (B * Hkv * s.rg * nsplit <= 256) ? next_stamp_slot() : nullptr;   // profiling aid (pie_debug_stamps)
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)(B * Hkv * s.rg * nsplit), 1, 1);
  cfg.blockDim = dim3(kAttnWarps * 32, 1, 1);
  cfg.stream = (cudaStream_t)stream;
  // 2..8 splits: the splits of a unit form one thread-block cluster and merge through distributed shared memory
  const bool cluster = nsplit > 1 && nsplit <= 8;
  cudaLaunchAttribute attr[2];
  int na = 0;
  if (pdl) {
    attr[na].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[na].val.programmaticStreamSerializationAllowed = 1;
    ++na;
  }
  if (cluster) {
    attr[na].id = cudaLaunchAttributeClusterDimension;
    attr[na].val.clusterDim.x = (unsigned)nsplit;
    attr[na].val.clusterDim.y = 1;
    attr[na].val.clusterDim.z = 1;
    ++na;
  }
  cfg.attrs = attr;
  cfg.numAttrs = na;
  count_launch();
  cudaError_t e;
#define PIE_ATTN_LAUNCH(DD, RR)                                                                  \
  (cluster ? cudaLaunchKernelEx(&cfg, attn_fused_kernel<DD, RR, true>, s, P)                     \
           : cudaLaunchKernelEx(&cfg, attn_fused_kernel<DD, RR, false>, s, P))
  if (D == 128) e = R == 2 ? PIE_ATTN_LAUNCH(128, 2) : PIE_ATTN_LAUNCH(128, 1);
  else e = R == 2 ? PIE_ATTN_LAUNCH(64, 2) : PIE_ATTN_LAUNCH(64, 1);
#undef PIE_ATTN_LAUNCH
  return status_of(e);
}

int pie_decode_step(const void* plan_dev, const void* plan_host, size_t plan_bytes, void* sync_ws, void* stream) {
  if (!plan_dev || !plan_host || !sync_ws) return PIE_ENULL;
  const StepHeader* H = reinterpret_cast<const StepHeader*>(plan_host);
  if (plan_bytes < sizeof(StepHeader) || H->magic != kStepMagic || H->plan_bytes > plan_bytes) return PIE_EINVAL_SHAPE;
  StepArgs s = H->args;
  s.phases = reinterpret_cast<const StepPhase*>(reinterpret_cast<const uint8_t*>(plan_dev) + sizeof(StepHeader));
  s.bar = reinterpret_cast<unsigned int*>(sync_ws);
  cudaStream_t st = (cudaStream_t)stream;
  return status_of(launch_step2(s, H->grid, H->smem, st));
}

}  // extern "C"
