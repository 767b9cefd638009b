// Decode engine for sm_100a: ONE persistent cooperative kernel runs a whole decode step (or several)
// of the quantized Llama forward pass (PIE/models/llama/language.py:75-210 over
// mlx.nn.QuantizedLinear, MLX/python/mlx/nn/layers/quantized.py:204-216).
//
// Structure (DESIGN.md has the full picture):
//   * warp 0 of every CTA is a TMA producer.  It walks the step's GEMV phases in program order and
//     streams this CTA's share of the packed weights (+ bf16 scales/biases) through a deep
//     shared-memory ring with cp.async.bulk.tensor + mbarrier complete_tx.  It never waits for
//     activations, so HBM keeps streaming across grid barriers and prologues.
//   * 16 consumer warps dequantise IN REGISTERS: 4-bit fields are lifted straight into bf16
//     mantissas ((w >> 4i) & 0x000f000f | 0x43004300 = 128+q) and fed to mma.sync.m16n8k16 as the A
//     operand (16 weight rows x 16 k), the activations (<= 8 batch columns) are the B operand.
//     Products x*(128+q) are exact in fp32; the per-group affine (scale, bias) is applied in fp32
//     once per 64-element group:  y += s * acc + (b - 128 s) * sum_x(group).
//     8-bit weights reuse the same datapath by splitting every byte into two nibbles and giving the
//     high nibble the activation 16*x (exact in bf16).
//   * work is cut stream-K style into (16-row tile, 512-element K chunk) units so that every SM
//     streams the same number of bytes; partial tiles are combined through shared-memory atomics
//     (inside a CTA) and a two-party global fix-up (across a CTA boundary).
//   * RMSNorm, embedding dequantisation, residual add, SwiGLU and arg-max are fused into the GEMV
//     prologue / epilogue; RoPE, KV append and split-K flash-decode attention form the one non-GEMV
//     phase per layer.
#include <cuda_bf16.h>
#include <math.h>

#include "pie_engine.h"
#include "pie_ptx.cuh"

namespace pie {

// ============================================================================================
// shared state
// ============================================================================================
struct Ctx {
  int bits, M, nstages;
  const CUtensorMap* maps;
  uint8_t* smem;        // dynamic smem base (1024-aligned)
  uint32_t ring_w, ring_s, ring_b, x_addr, sumx_addr;  // shared-space addresses
  float* slots;         // [kTileSlots][16 * M]
  float* sumx;          // [groups][8]
  uint8_t* xs;          // staged activations
  int x_ld;             // bf16 per staged row
  unsigned int* tile_cnt;
  float* tile_part;
  uint64_t* full;
  uint64_t* empty;
  unsigned int* slot_cnt;
  float* red;           // [64] block-reduce scratch
};

__device__ __forceinline__ int ctid() { return (int)threadIdx.x - 32; }  // consumer thread id
__device__ __forceinline__ void consumer_bar() {
  asm volatile("bar.sync 1, %0;" ::"n"(kConsumerWarps * 32) : "memory");
}

__device__ __forceinline__ void cta_range(const GemvDesc& d, int aligned, long long& u0, long long& u1) {
  const long long G = gridDim.x, c = blockIdx.x;
  if (aligned) {
    u0 = (c * d.ntiles / G) * d.chunks;
    u1 = ((c + 1) * d.ntiles / G) * d.chunks;
  } else {
    const long long U = (long long)d.ntiles * d.chunks;
    u0 = c * U / G;
    u1 = (c + 1) * U / G;
  }
}
// A phase is cut at unit granularity only when every CTA gets at least one whole tile's worth of
// units, so a tile is shared by at most two CTAs.
__device__ __forceinline__ int phase_aligned(const GemvDesc& d) {
  return ((long long)d.ntiles * d.chunks) / (long long)gridDim.x < d.chunks ? 1 : 0;
}

// ============================================================================================
// producer: stream one GEMV phase's weights for this CTA through the ring
// ============================================================================================
__device__ void produce_phase(const Ctx& c, const GemvDesc& d, uint32_t& cnt, uint64_t policy) {
  long long u0, u1;
  cta_range(d, phase_aligned(d), u0, u1);
  const int wbytes = stage_w_bytes(c.bits);
  const int nbox = stage_boxes(c.bits);
  const int stage_words = kStageK * c.bits / 32;
  for (long long u = u0; u < u1; ++u, ++cnt) {
    const int slot = cnt % c.nstages;
    const uint32_t par = (cnt / c.nstages) & 1;
    mbar_wait(&c.empty[slot], par ^ 1);
    const int tile = (int)(u / d.chunks), chunk = (int)(u % d.chunks);
    uint64_t* bar = &c.full[slot];
    mbar_arrive_expect_tx(bar, wbytes + 2 * stage_sb_bytes());
    const uint32_t wdst = c.ring_w + (uint32_t)slot * wbytes;
    const uint32_t sdst = c.ring_s + (uint32_t)slot * stage_sb_bytes();
    const uint32_t bdst = c.ring_b + (uint32_t)slot * stage_sb_bytes();
    const int kw0 = chunk * stage_words;
    const int kg0 = chunk * kStageGroups;
    if (!d.dual) {
      int si = 0;
      while (si + 1 < d.nseg && tile >= d.seg[si + 1].tile0) ++si;
      const GemvSeg& s = d.seg[si];
      const int row0 = (tile - s.tile0) * kTileRows;
      for (int b = 0; b < nbox; ++b)
        tma_load_2d(wdst + b * 2048, &c.maps[s.map_w], bar, kw0 + b * 32, row0, policy);
      tma_load_2d(sdst, &c.maps[s.map_s], bar, kg0, row0, policy);
      tma_load_2d(bdst, &c.maps[s.map_b], bar, kg0, row0, policy);
    } else {
      const int row0 = tile * 8;
      for (int b = 0; b < nbox; ++b) {
        tma_load_2d(wdst + b * 2048, &c.maps[d.seg[0].map_w], bar, kw0 + b * 32, row0, policy);
        tma_load_2d(wdst + b * 2048 + 1024, &c.maps[d.seg[1].map_w], bar, kw0 + b * 32, row0, policy);
      }
      tma_load_2d(sdst, &c.maps[d.seg[0].map_s], bar, kg0, row0, policy);
      tma_load_2d(sdst + 128, &c.maps[d.seg[1].map_s], bar, kg0, row0, policy);
      tma_load_2d(bdst, &c.maps[d.seg[0].map_b], bar, kg0, row0, policy);
      tma_load_2d(bdst + 128, &c.maps[d.seg[1].map_b], bar, kg0, row0, policy);
    }
  }
}

// ============================================================================================
// consumer: one (16-row tile, 512-k chunk) unit per warp
// ============================================================================================
__device__ __forceinline__ float sel_bf16(const uint4& v, int i) {
  const uint32_t w = (i >> 1) == 0 ? v.x : (i >> 1) == 1 ? v.y : (i >> 1) == 2 ? v.z : v.w;
  return (i & 1) ? bf16hi(w) : bf16lo(w);
}

template <int BITS>
__device__ __forceinline__ void consume_unit(const Ctx& c, int slot, int chunk, float (&fin)[4]) {
  const int lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
  constexpr int WB = stage_w_bytes(BITS);
  constexpr int SLABS_PER_GROUP = BITS / 4;         // 32-byte row slabs per 64-element group
  constexpr int STAGE_WORDS = kStageK * BITS / 32;  // words per row per stage
  constexpr float OFFS = (BITS == 4) ? 128.0f : 128.0f * 17.0f;
  const uint32_t wbase = c.ring_w + slot * WB;
  const uint32_t sbase = c.ring_s + slot * stage_sb_bytes();
  const uint32_t bbase = c.ring_b + slot * stage_sb_bytes();
  const uint4 sA = lds128(sbase + g * 16), sB = lds128(sbase + (g + 8) * 16);
  const uint4 bA = lds128(bbase + g * 16), bB = lds128(bbase + (g + 8) * 16);
  const bool xact = g < c.M;
  const uint32_t xrow = c.x_addr + (uint32_t)(g * c.x_ld) * 2u + (uint32_t)(chunk * STAGE_WORDS) * 16u;
  const uint32_t sxrow = c.sumx_addr + (uint32_t)(chunk * kStageGroups) * 32u + t * 8u;
  const uint32_t swz = (uint32_t)(g & 7);
#pragma unroll
  for (int gi = 0; gi < kStageGroups; ++gi) {
    float acc[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int h = 0; h < SLABS_PER_GROUP; ++h) {
      const int slab = gi * SLABS_PER_GROUP + h;
      const int box = slab >> 2, sl = slab & 3;
      const uint32_t off = ((((uint32_t)(2 * sl) + (uint32_t)(t >> 1)) ^ swz) << 4) + ((uint32_t)(t & 1) << 3);
      const uint2 wA = lds64(wbase + box * 2048 + g * 128 + off);
      const uint2 wB = lds64(wbase + box * 2048 + (g + 8) * 128 + off);
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        uint4 xf = make_uint4(0u, 0u, 0u, 0u);
        if (xact) xf = lds128(xrow + (uint32_t)(slab * 8 + 2 * t + j) * 16u);
        const uint32_t WA = j ? wA.y : wA.x, WB_ = j ? wB.y : wB.x;
        constexpr uint32_t MSK = 0x000f000fu, MAG = 0x43004300u;
        mma_bf16_16816(acc, (WA & MSK) | MAG, (WB_ & MSK) | MAG, ((WA >> 4) & MSK) | MAG,
                       ((WB_ >> 4) & MSK) | MAG, xf.x, xf.y);
        mma_bf16_16816(acc, ((WA >> 8) & MSK) | MAG, ((WB_ >> 8) & MSK) | MAG,
                       ((WA >> 12) & MSK) | MAG, ((WB_ >> 12) & MSK) | MAG, xf.z, xf.w);
      }
    }
    const float s0 = sel_bf16(sA, gi), s1 = sel_bf16(sB, gi);
    const float c0 = fmaf(-OFFS, s0, sel_bf16(bA, gi)), c1 = fmaf(-OFFS, s1, sel_bf16(bB, gi));
    const uint2 sx = lds64(sxrow + gi * 32);
    const float sx0 = __uint_as_float(sx.x), sx1 = __uint_as_float(sx.y);
    fin[0] = fmaf(s0, acc[0], fmaf(c0, sx0, fin[0]));
    fin[1] = fmaf(s0, acc[1], fmaf(c0, sx1, fin[1]));
    fin[2] = fmaf(s1, acc[2], fmaf(c1, sx0, fin[2]));
    fin[3] = fmaf(s1, acc[3], fmaf(c1, sx1, fin[3]));
  }
}

// ------------------------------------------------------------------ epilogues
__device__ __forceinline__ unsigned long long argmax_key(float v, int n) {
  uint32_t u = __float_as_uint(v);
  u = (u & 0x80000000u) ? ~u : (u | 0x80000000u);
  return ((unsigned long long)u << 32) | (unsigned long long)(0xFFFFFFFFu - (uint32_t)n);
}

// One warp finishes a tile whose complete fp32 sums sit in sp[16][M].
__device__ void tile_epilogue(const Ctx& c, const GemvDesc& d, int tile, float* sp, unsigned long long* argmax) {
  const int lane = threadIdx.x & 31, M = c.M;
  __nv_bfloat16* out = (__nv_bfloat16*)d.out;
  if (d.epi == EPI_SILUMUL) {
    const int n0 = tile * 8, N = d.seg[0].N;
    for (int e = lane; e < 8 * M; e += 32) {
      const int r = e / M, m = e % M, n = n0 + r;
      if (n < N) {
        const float gv = bfround(sp[r * M + m]), uv = bfround(sp[(r + 8) * M + m]);
        const float sg = bfround(1.0f / (1.0f + expf(-gv)));
        const float a = bfround(gv * sg);
        out[(size_t)m * d.out_ld + n] = __float2bfloat16_rn(a * uv);
      }
    }
    return;
  }
  int si = 0;
  while (si + 1 < d.nseg && tile >= d.seg[si + 1].tile0) ++si;
  const GemvSeg& s = d.seg[si];
  const int n0 = (tile - s.tile0) * kTileRows;
  if (d.epi == EPI_LOGITS) {
    for (int m = 0; m < M; ++m) {
      unsigned long long key = 0ull;
      const int n = n0 + lane;
      if (lane < kTileRows && n < s.N) {
        const __nv_bfloat16 y = __float2bfloat16_rn(sp[lane * M + m]);
        if (out) out[(size_t)m * d.out_ld + s.out_off + n] = y;
        key = argmax_key(__bfloat162float(y), s.out_off + n);
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        const unsigned long long other = __shfl_xor_sync(0xffffffffu, key, o);
        key = other > key ? other : key;
      }
      if (lane == 0) atomicMax(&argmax[m], key);
    }
    return;
  }
  for (int e = lane; e < kTileRows * M; e += 32) {
    const int r = e / M, m = e % M, n = n0 + r;
    if (n < s.N) {
      const size_t idx = (size_t)m * d.out_ld + s.out_off + n;
      const float y = sp[r * M + m];
      if (d.epi == EPI_RESADD) {
        const float hv = __uint_as_float(((uint32_t)ldcg16(out + idx)) << 16);
        out[idx] = __float2bfloat16_rn(hv + bfround(y));
      } else {
        out[idx] = __float2bfloat16_rn(y);
      }
    }
  }
}

// ------------------------------------------------------------------ consumer side of a GEMV phase
template <int BITS>
__device__ void consume_phase(const Ctx& c, const GemvDesc& d, uint32_t& cnt_base, unsigned long long* argmax) {
  const int warp = (threadIdx.x >> 5) - 1, lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
  const int M = c.M;
  const int aligned = phase_aligned(d);
  long long u0, u1;
  cta_range(d, aligned, u0, u1);
  const int nunits = (int)(u1 - u0);
  const int first_tile = (int)(u0 / d.chunks);
  for (int i = warp; i < nunits; i += kConsumerWarps) {
    const uint32_t cnt = cnt_base + i;
    const int slot = cnt % c.nstages;
    const long long u = u0 + i;
    const int tile = (int)(u / d.chunks), chunk = (int)(u % d.chunks);
    mbar_wait(&c.full[slot], (cnt / c.nstages) & 1);
    float fin[4] = {0.f, 0.f, 0.f, 0.f};
    consume_unit<BITS>(c, slot, chunk, fin);
    // accumulate this unit into the tile slot
    const int ts = (tile - first_tile) % kTileSlots;
    float* sp = c.slots + ts * (kTileRows * M);
    if (2 * t < M) {
      atomicAdd(&sp[g * M + 2 * t], fin[0]);
      atomicAdd(&sp[(g + 8) * M + 2 * t], fin[2]);
    }
    if (2 * t + 1 < M) {
      atomicAdd(&sp[g * M + 2 * t + 1], fin[1]);
      atomicAdd(&sp[(g + 8) * M + 2 * t + 1], fin[3]);
    }
    __threadfence_block();
    __syncwarp();
    const long long tb = (long long)tile * d.chunks, te = tb + d.chunks;
    const int expect = (int)((te < u1 ? te : u1) - (tb > u0 ? tb : u0));
    unsigned int old = 0;
    if (lane == 0) old = atomicAdd(&c.slot_cnt[ts], 1u);
    old = __shfl_sync(0xffffffffu, old, 0);
    if ((int)old == expect - 1) {
      // last unit of this tile inside this CTA
      __threadfence_block();
      bool do_final = true;
      if (expect != d.chunks) {
        // tile shared with a neighbouring CTA: two-party fix-up through global memory
        const int boundary = (tb < u0) ? (int)blockIdx.x : (int)blockIdx.x + 1;  // head or tail tile
        const int mypart = (tb < u0) ? 1 : 0;
        float* part = c.tile_part + ((size_t)boundary * 2) * (kTileRows * kMaxBatch);
        for (int e = lane; e < kTileRows * M; e += 32) part[mypart * (kTileRows * kMaxBatch) + e] = sp[e];
        __threadfence();
        __syncwarp();
        unsigned int prev = 0;
        if (lane == 0) prev = atomicAdd(&c.tile_cnt[boundary], 1u);
        prev = __shfl_sync(0xffffffffu, prev, 0);
        if (prev == 0) {
          do_final = false;  // the neighbour finishes the tile
        } else {
          __threadfence();
          const float* other = part + (1 - mypart) * (kTileRows * kMaxBatch);
          // fixed summation order (part 0 + part 1) keeps the result independent of arrival order
          for (int e = lane; e < kTileRows * M; e += 32) {
            const float o = ldcg_f32(other + e);
            sp[e] = mypart ? (o + sp[e]) : (sp[e] + o);
          }
          if (lane == 0) c.tile_cnt[boundary] = 0;
          __syncwarp();
        }
      }
      if (do_final) tile_epilogue(c, d, tile, sp, argmax);
      __syncwarp();
      for (int e = lane; e < kTileRows * M; e += 32) sp[e] = 0.f;
      if (lane == 0) c.slot_cnt[ts] = 0;
      __threadfence_block();
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(&c.empty[slot]);
  }
  cnt_base += nunits;
}

// ============================================================================================
// activation staging (GEMV prologue): x -> permuted bf16 B-fragments + per-group sums
// ============================================================================================
// 8 consecutive elements (bf16-exact floats) of row m starting at element 8*blk.
template <int BITS>
__device__ __forceinline__ void stage_block(const Ctx& c, int m, int blk, const float (&v)[8]) {
  if (BITS == 4) {
    uint4 o;
    o.x = pack_bf16(v[0], v[4]);
    o.y = pack_bf16(v[1], v[5]);
    o.z = pack_bf16(v[2], v[6]);
    o.w = pack_bf16(v[3], v[7]);
    *reinterpret_cast<uint4*>(c.xs + ((size_t)m * c.x_ld + (size_t)blk * 8) * 2) = o;
  } else {
#pragma unroll
    for (int w = 0; w < 2; ++w) {
      const float x0 = v[4 * w], x1 = v[4 * w + 1], x2 = v[4 * w + 2], x3 = v[4 * w + 3];
      uint4 o;
      o.x = pack_bf16(x0, x2);
      o.y = pack_bf16(16.f * x0, 16.f * x2);
      o.z = pack_bf16(x1, x3);
      o.w = pack_bf16(16.f * x1, 16.f * x3);
      *reinterpret_cast<uint4*>(c.xs + ((size_t)m * c.x_ld + (size_t)(2 * blk + w) * 8) * 2) = o;
    }
  }
}

__device__ __forceinline__ void unpack8(const uint4& r, float (&v)[8]) {
  v[0] = bf16lo(r.x); v[1] = bf16hi(r.x); v[2] = bf16lo(r.y); v[3] = bf16hi(r.y);
  v[4] = bf16lo(r.z); v[5] = bf16hi(r.z); v[6] = bf16lo(r.w); v[7] = bf16hi(r.w);
}

// block-wide sum over the 512 consumer threads (two consumer barriers)
__device__ __forceinline__ float consumer_block_sum(const Ctx& c, float v) {
  const int warp = (threadIdx.x >> 5) - 1, lane = threadIdx.x & 31;
  v = warp_sum(v);
  if (lane == 0) c.red[warp] = v;
  consumer_bar();
  float tot = 0.f;
#pragma unroll
  for (int w = 0; w < kConsumerWarps; ++w) tot += c.red[w];
  consumer_bar();
  return tot;
}

struct EmbedSrc {
  const uint32_t* w; const uint16_t* s; const uint16_t* b; const int* tokens; __nv_bfloat16* h_out;
};

// Stage x for one GEMV phase.  K = reduction length; rows m < M.
template <int BITS>
__device__ void stage_x(const Ctx& c, const GemvDesc& d, const EmbedSrc& emb) {
  const int tid = ctid(), lane = threadIdx.x & 31;
  const int K = d.K, nblk = K / 8, M = c.M;
  const int nblk_pad = d.chunks * (kStageK / 8);  // zero-padded to whole stages
  constexpr int NT = kConsumerWarps * 32;
  constexpr int MAXB = 4;  // register-cached blocks per thread (K <= 16384 for normed inputs)
  for (int m = 0; m < M; ++m) {
    if (d.xmode == X_PLAIN) {
      const uint4* src = reinterpret_cast<const uint4*>((const __nv_bfloat16*)d.x_in + (size_t)m * K);
      for (int b0 = 0; b0 < nblk_pad; b0 += NT) {
        const int blk = b0 + tid;
        float v[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        if (blk < nblk) unpack8(ldcg128(src + blk), v);
        float s = ((v[0] + v[1]) + (v[2] + v[3])) + ((v[4] + v[5]) + (v[6] + v[7]));
        s += __shfl_xor_sync(0xffffffffu, s, 1);
        s += __shfl_xor_sync(0xffffffffu, s, 2);
        s += __shfl_xor_sync(0xffffffffu, s, 4);
        if (blk < nblk_pad) {
          stage_block<BITS>(c, m, blk, v);
          if ((lane & 7) == 0) c.sumx[(blk >> 3) * 8 + m] = s;
        }
      }
    } else {
      // X_NORM / X_EMBED: rms_norm (MLX/mlx/fast.cpp:93-111): fp32 reduce, round to bf16, then * w in bf16
      float v[MAXB][8];
      float ss = 0.f;
#pragma unroll
      for (int i = 0; i < MAXB; ++i) {
        const int blk = i * NT + tid;
#pragma unroll
        for (int e = 0; e < 8; ++e) v[i][e] = 0.f;
        if (blk < nblk) {
          if (d.xmode == X_NORM) {
            unpack8(ldcg128(reinterpret_cast<const uint4*>((const __nv_bfloat16*)d.x_in + (size_t)m * K) + blk), v[i]);
          } else {
            // QuantizedEmbedding.__call__ (quantized.py:100-107) + dequantize fallback (fast.cpp:972-1017):
            // w_hat = bf16(bf16(q * scale) + bias)
            const size_t row = (size_t)emb.tokens[m];
            const int grp = blk >> 3;
            const float sc = __uint_as_float(((uint32_t)emb.s[row * (K / kGroup) + grp]) << 16);
            const float bi = __uint_as_float(((uint32_t)emb.b[row * (K / kGroup) + grp]) << 16);
            if (BITS == 4) {
              const uint32_t w = emb.w[row * (K / 8) + blk];
#pragma unroll
              for (int e = 0; e < 8; ++e) v[i][e] = bfround(bfround((float)((w >> (4 * e)) & 15u) * sc) + bi);
            } else {
              const uint2 w = *reinterpret_cast<const uint2*>(emb.w + row * (K / 4) + 2 * blk);
#pragma unroll
              for (int e = 0; e < 4; ++e) {
                v[i][e] = bfround(bfround((float)((w.x >> (8 * e)) & 255u) * sc) + bi);
                v[i][4 + e] = bfround(bfround((float)((w.y >> (8 * e)) & 255u) * sc) + bi);
              }
            }
            if (blockIdx.x == 0) {
              uint4 o;
              o.x = pack_bf16(v[i][0], v[i][1]); o.y = pack_bf16(v[i][2], v[i][3]);
              o.z = pack_bf16(v[i][4], v[i][5]); o.w = pack_bf16(v[i][6], v[i][7]);
              reinterpret_cast<uint4*>(emb.h_out + (size_t)m * K)[blk] = o;
            }
          }
#pragma unroll
          for (int e = 0; e < 8; ++e) ss = fmaf(v[i][e], v[i][e], ss);
        }
      }
      ss = consumer_block_sum(c, ss);
      const float inv = 1.0f / sqrtf(ss / (float)K + d.eps);
#pragma unroll
      for (int i = 0; i < MAXB; ++i) {
        const int blk = i * NT + tid;
        if (i * NT < nblk_pad) {  // warp-uniform
          float y[8] = {0, 0, 0, 0, 0, 0, 0, 0};
          if (blk < nblk) {
            float wv[8];
            unpack8(__ldg(reinterpret_cast<const uint4*>(d.norm_w) + blk), wv);
#pragma unroll
            for (int e = 0; e < 8; ++e) y[e] = bfround(bfround(v[i][e] * inv) * wv[e]);
          }
          float s = ((y[0] + y[1]) + (y[2] + y[3])) + ((y[4] + y[5]) + (y[6] + y[7]));
          s += __shfl_xor_sync(0xffffffffu, s, 1);
          s += __shfl_xor_sync(0xffffffffu, s, 2);
          s += __shfl_xor_sync(0xffffffffu, s, 4);
          if (blk < nblk_pad) {
            stage_block<BITS>(c, m, blk, y);
            if ((lane & 7) == 0) c.sumx[(blk >> 3) * 8 + m] = s;
          }
        }
      }
    }
  }
  consumer_bar();
}

// ============================================================================================
// grid barrier (consumer warps only; the producer keeps streaming)
// ============================================================================================
__device__ __forceinline__ void grid_sync(unsigned int* bar, unsigned int& target) {
  consumer_bar();
  target += gridDim.x;
  if (ctid() == 0) {
    __threadfence();
    red_release_gpu_add(bar, 1u);
    uint32_t spins = 0;
    while (ld_acquire_gpu(bar) < target) {
      if (++spins > PIE_SPIN_LIMIT) __trap();
    }
    __threadfence();
  }
  consumer_bar();
}

// ============================================================================================
// attention phase: RoPE (PIE/models/llama/utils.py:42-50 -> MLX/mlx/fast.cpp:397-471), KV append
// (PIE/cache/kv_cache/reusable.py:136-139) and split-K flash-decode (the sdpa_vector_2pass scheme,
// MLX/mlx/backend/metal/kernels/sdpa_vector.h:168-392) for L = 1.
// ============================================================================================
template <int D, int NREP>
__device__ void attn_phase(const EngineParams& p, const Ctx& c, int layer, const int* pos_s, float* scratch) {
  constexpr int EPL = D / 32;  // elements per lane
  const int warp = (threadIdx.x >> 5) - 1, lane = threadIdx.x & 31, tid = ctid();
  const int B = p.batch, Hkv = p.n_kv, nsplit = p.nsplit;
  const int qkv_ld = (p.n_heads + 2 * p.n_kv) * D;
  const __nv_bfloat16* qkv = (const __nv_bfloat16*)p.qkv;
  __nv_bfloat16* kc = (__nv_bfloat16*)p.attn[layer].kcache;
  __nv_bfloat16* vc = (__nv_bfloat16*)p.attn[layer].vcache;
  const int nitems = B * Hkv * nsplit;
  for (int item = blockIdx.x; item < nitems; item += gridDim.x) {
    const int split = item % nsplit, kvh = (item / nsplit) % Hkv, b = item / (nsplit * Hkv);
    const int pos = pos_s[b];
    const int T = pos + 1;
    int len = (T + nsplit - 1) / nsplit;
    len = (len + 7) & ~7;
    const int t0 = split * len;
    int t1 = t0 + len;
    if (t1 > T) t1 = T;
    const bool owner = (pos >= t0) && (pos < t0 + len);
    // ---- rotate q (all NREP heads of this kv head) and the new k ----
    float cs[EPL], sn[EPL];
#pragma unroll
    for (int e = 0; e < EPL; ++e) {
      const int i = (lane * EPL + e) % (D / 2);
      const float theta = (float)pos * (1.0f / p.rope_freqs[i]);
      sincosf(theta, &sn[e], &cs[e]);
    }
    auto load_row = [&](const __nv_bfloat16* src, float (&o)[EPL]) {
      if (EPL == 4) {
        const uint2 r = ldcg64(src + lane * 4);
        o[0] = bf16lo(r.x); o[1] = bf16hi(r.x); o[2] = bf16lo(r.y); o[3] = bf16hi(r.y);
      } else {
        const uint32_t r = ldcg32(src + lane * 2);
        o[0] = bf16lo(r); o[1] = bf16hi(r);
      }
    };
    auto rotate = [&](float (&x)[EPL]) {
#pragma unroll
      for (int e = 0; e < EPL; ++e) {
        const float other = __shfl_xor_sync(0xffffffffu, x[e], 16);
        // out1 = x1*cos - x2*sin ; out2 = x1*sin + x2*cos   (each product rounded, as the op graph)
        const float r = (lane < 16) ? __fsub_rn(__fmul_rn(x[e], cs[e]), __fmul_rn(other, sn[e]))
                                    : __fadd_rn(__fmul_rn(other, sn[e]), __fmul_rn(x[e], cs[e]));
        x[e] = bfround(r);  // op boundary: rope output is bf16
      }
    };
    float q[NREP][EPL];
#pragma unroll
    for (int r = 0; r < NREP; ++r) {
      load_row(qkv + (size_t)b * qkv_ld + (size_t)(kvh * NREP + r) * D, q[r]);
      rotate(q[r]);
#pragma unroll
      for (int e = 0; e < EPL; ++e) q[r][e] *= p.attn_scale;
    }
    float mx[NREP], l[NREP], acc[NREP][EPL];
#pragma unroll
    for (int r = 0; r < NREP; ++r) {
      mx[r] = -INFINITY;
      l[r] = 0.f;
#pragma unroll
      for (int e = 0; e < EPL; ++e) acc[r][e] = 0.f;
    }
    auto kv_off = [&](int t) -> size_t {
      if (p.page_table) {
        const int page = p.page_table[(size_t)b * p.pt_stride + t / p.page_size];
        return ((size_t)page * p.page_size + (size_t)(t % p.page_size)) * p.kv_tok_stride +
               (size_t)kvh * p.kv_head_stride;
      }
      return (size_t)b * p.kv_batch_stride + (size_t)kvh * p.kv_head_stride + (size_t)t * p.kv_tok_stride;
    };
    auto step = [&](const float (&kk)[EPL], const float (&vv)[EPL]) {
#pragma unroll
      for (int r = 0; r < NREP; ++r) {
        float s = 0.f;
#pragma unroll
        for (int e = 0; e < EPL; ++e) s = fmaf(q[r][e], kk[e], s);
        s = warp_sum(s);
        const float mn = fmaxf(mx[r], s);
        const float corr = __expf(mx[r] - mn), pr = __expf(s - mn);
        l[r] = l[r] * corr + pr;
#pragma unroll
        for (int e = 0; e < EPL; ++e) acc[r][e] = fmaf(pr, vv[e], acc[r][e] * corr);
        mx[r] = mn;
      }
    };
    const int tend = (owner && t1 > pos) ? pos : t1;  // the new token is handled from registers
    for (int t = t0 + warp; t < tend; t += kConsumerWarps) {
      float kk[EPL], vv[EPL];
      const size_t off = kv_off(t);
      load_row(kc + off, kk);
      load_row(vc + off, vv);
      step(kk, vv);
    }
    if (owner && warp == 0) {
      float kk[EPL], vv[EPL];
      load_row(qkv + (size_t)b * qkv_ld + (size_t)(p.n_heads + kvh) * D, kk);
      load_row(qkv + (size_t)b * qkv_ld + (size_t)(p.n_heads + p.n_kv + kvh) * D, vv);
      rotate(kk);
      const size_t off = kv_off(pos);
      if (EPL == 4) {
        *reinterpret_cast<uint2*>(kc + off + lane * 4) = make_uint2(pack_bf16(kk[0], kk[1]), pack_bf16(kk[2], kk[3]));
        *reinterpret_cast<uint2*>(vc + off + lane * 4) = make_uint2(pack_bf16(vv[0], vv[1]), pack_bf16(vv[2], vv[3]));
      } else {
        *reinterpret_cast<uint32_t*>(kc + off + lane * 2) = pack_bf16(kk[0], kk[1]);
        *reinterpret_cast<uint32_t*>(vc + off + lane * 2) = pack_bf16(vv[0], vv[1]);
      }
      step(kk, vv);
    }
    // ---- combine the 16 warps through shared memory ----
    constexpr int RS = D + 2;  // per (warp, head) record: acc[D], m, l
#pragma unroll
    for (int r = 0; r < NREP; ++r) {
      float* rec = scratch + ((size_t)warp * NREP + r) * RS;
#pragma unroll
      for (int e = 0; e < EPL; ++e) rec[lane * EPL + e] = acc[r][e];
      if (lane == 0) { rec[D] = mx[r]; rec[D + 1] = l[r]; }
    }
    consumer_bar();
    float* part = p.attn_part;
    for (int e = tid; e < NREP * D; e += kConsumerWarps * 32) {
      const int r = e / D, dd = e % D;
      float M_ = -INFINITY;
#pragma unroll
      for (int w = 0; w < kConsumerWarps; ++w) M_ = fmaxf(M_, scratch[((size_t)w * NREP + r) * RS + D]);
      float L_ = 0.f, A_ = 0.f;
#pragma unroll
      for (int w = 0; w < kConsumerWarps; ++w) {
        const float* rec = scratch + ((size_t)w * NREP + r) * RS;
        const float f = (rec[D] == -INFINITY) ? 0.f : __expf(rec[D] - M_);
        L_ = fmaf(rec[D + 1], f, L_);
        A_ = fmaf(rec[dd], f, A_);
      }
      const int head = kvh * NREP + r;
      if (nsplit == 1) {
        ((__nv_bfloat16*)p.attn_out)[(size_t)b * p.n_heads * D + (size_t)head * D + dd] = __float2bfloat16_rn(A_ / L_);
      } else {
        float* rec = part + (((size_t)b * p.n_heads + head) * nsplit + split) * RS;
        rec[dd] = A_;
        if (dd == 0) { rec[D] = M_; rec[D + 1] = L_; }
      }
    }
    if (nsplit > 1) {
      __threadfence();
      consumer_bar();
      if (tid == 0) {
        const unsigned int prev = atomicAdd(&p.attn_cnt[b * Hkv + kvh], 1u);
        c.red[32] = (prev == (unsigned)(nsplit - 1)) ? 1.f : 0.f;
        if (prev == (unsigned)(nsplit - 1)) p.attn_cnt[b * Hkv + kvh] = 0;
      }
      consumer_bar();
      if (c.red[32] != 0.f) {
        __threadfence();
        for (int e = tid; e < NREP * D; e += kConsumerWarps * 32) {
          const int r = e / D, dd = e % D, head = kvh * NREP + r;
          const float* base = part + ((size_t)b * p.n_heads + head) * nsplit * RS;
          float M_ = -INFINITY;
          for (int s = 0; s < nsplit; ++s) M_ = fmaxf(M_, ldcg_f32(base + (size_t)s * RS + D));
          float L_ = 0.f, A_ = 0.f;
          for (int s = 0; s < nsplit; ++s) {
            const float ms = ldcg_f32(base + (size_t)s * RS + D);
            const float f = (ms == -INFINITY) ? 0.f : __expf(ms - M_);
            L_ = fmaf(ldcg_f32(base + (size_t)s * RS + D + 1), f, L_);
            A_ = fmaf(ldcg_f32(base + (size_t)s * RS + dd), f, A_);
          }
          ((__nv_bfloat16*)p.attn_out)[(size_t)b * p.n_heads * D + (size_t)head * D + dd] = __float2bfloat16_rn(A_ / L_);
        }
      }
    }
    consumer_bar();  // scratch / red reuse
  }
}

// ============================================================================================
// common setup
// ============================================================================================
__device__ __forceinline__ void init_ring(Ctx& c, uint64_t* full, uint64_t* empty, unsigned int* slot_cnt, float* red) {
  c.full = full; c.empty = empty; c.slot_cnt = slot_cnt; c.red = red;
  if (threadIdx.x == 0) {
    for (int i = 0; i < c.nstages; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], 1); }
    fence_mbar_init();
  }
  for (int i = threadIdx.x; i < kTileSlots; i += blockDim.x) slot_cnt[i] = 0;
  for (int i = threadIdx.x; i < kTileSlots * kTileRows * c.M; i += blockDim.x) c.slots[i] = 0.f;
  __syncthreads();
}

// ============================================================================================
// kernels
// ============================================================================================
template <int BITS, int D, int NREP>
__global__ void __launch_bounds__(kThreads, 1) pie_decode_kernel(const __grid_constant__ EngineParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  __shared__ uint64_t full[kMaxStages], empty[kMaxStages];
  __shared__ unsigned int slot_cnt[kTileSlots];
  __shared__ float red[64];
  __shared__ int pos_s[kMaxBatch], tok_s[kMaxBatch];

  Ctx c;
  c.bits = BITS; c.M = p.batch; c.nstages = p.nstages; c.maps = p.maps;
  c.smem = smem_raw;
  const uint32_t sb = smem_u32(smem_raw);
  c.ring_w = sb + p.smem_ring_w; c.ring_s = sb + p.smem_ring_s; c.ring_b = sb + p.smem_ring_b;
  c.x_addr = sb + p.smem_x; c.sumx_addr = sb + p.smem_sumx;
  c.xs = smem_raw + p.smem_x; c.sumx = (float*)(smem_raw + p.smem_sumx);
  c.slots = (float*)(smem_raw + p.smem_slots);
  c.x_ld = p.x_ld;
  c.tile_cnt = p.tile_cnt; c.tile_part = p.tile_part;
  init_ring(c, full, empty, slot_cnt, red);

  const int L = p.n_layers;
  if (threadIdx.x < 32) {
    // ------------------------------ producer warp ------------------------------
    if (threadIdx.x == 0) {
      const uint64_t pol = l2_policy_evict_first();
      uint32_t cnt = 0;
      for (int step = 0; step < p.nsteps; ++step)
        for (int ph = 0; ph < 4 * L + 1; ++ph) produce_phase(c, p.gemv[ph], cnt, pol);
    }
    return;
  }
  // ------------------------------ consumer warps ------------------------------
  float* attn_scratch = (float*)(smem_raw + p.smem_attn);
  unsigned int bar_target = 0;
  uint32_t cnt = 0;
  EmbedSrc emb{p.embed_w, p.embed_s, p.embed_b, tok_s, (__nv_bfloat16*)p.h};
  if (ctid() < p.batch) { pos_s[ctid()] = p.positions[ctid()]; tok_s[ctid()] = p.tokens[ctid()]; }
  consumer_bar();
  for (int step = 0; step < p.nsteps; ++step) {
    unsigned long long* amax = p.argmax + (size_t)(step & 1) * kMaxBatch;
    for (int l = 0; l < L; ++l) {
      const GemvDesc* g = p.gemv + 4 * l;
      stage_x<BITS>(c, g[0], emb);                 // (embed ->) input_layernorm -> x
      consume_phase<BITS>(c, g[0], cnt, amax);     // q,k,v projections
      grid_sync(p.grid_bar, bar_target);
      if (l == 0 && step > 0 && blockIdx.x == 0 && ctid() < p.batch)
        p.argmax[(size_t)((step - 1) & 1) * kMaxBatch + ctid()] = 0ull;  // everyone has read it by now
      attn_phase<D, NREP>(p, c, l, pos_s, attn_scratch);
      grid_sync(p.grid_bar, bar_target);
      stage_x<BITS>(c, g[1], emb);
      consume_phase<BITS>(c, g[1], cnt, amax);     // o_proj + residual
      grid_sync(p.grid_bar, bar_target);
      stage_x<BITS>(c, g[2], emb);                 // post_attention_layernorm -> x
      consume_phase<BITS>(c, g[2], cnt, amax);     // gate/up + SwiGLU
      grid_sync(p.grid_bar, bar_target);
      stage_x<BITS>(c, g[3], emb);
      consume_phase<BITS>(c, g[3], cnt, amax);     // down_proj + residual
      grid_sync(p.grid_bar, bar_target);
    }
    stage_x<BITS>(c, p.gemv[4 * L], emb);          // final norm
    consume_phase<BITS>(c, p.gemv[4 * L], cnt, amax);  // lm_head + arg-max
    grid_sync(p.grid_bar, bar_target);
    if (ctid() < p.batch) {
      const unsigned long long key = *((volatile unsigned long long*)&amax[ctid()]);
      const int tok = (int)(0xFFFFFFFFu - (uint32_t)(key & 0xFFFFFFFFull));
      tok_s[ctid()] = tok;
      pos_s[ctid()] += 1;
      if (blockIdx.x == 0) {
        if (p.out_tokens) p.out_tokens[(size_t)step * p.batch + ctid()] = tok;
        if (step == p.nsteps - 1) { p.tokens[ctid()] = tok; p.positions[ctid()] = pos_s[ctid()]; }
      }
    }
    consumer_bar();
  }
}

// Standalone quantized matvec: one GEMV phase of the same machinery (mx.quantized_matmul for M <= 8).
template <int BITS>
__global__ void __launch_bounds__(kThreads, 1)
pie_qmv_kernel(const __grid_constant__ QmvParams p, const __grid_constant__ CUtensorMap map_w,
               const __grid_constant__ CUtensorMap map_s, const __grid_constant__ CUtensorMap map_b) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  __shared__ uint64_t full[kMaxStages], empty[kMaxStages];
  __shared__ unsigned int slot_cnt[kTileSlots];
  __shared__ float red[64];
  // The three tensor maps stay in param space (grid constants); TMA reads them from there.
  Ctx c;
  c.bits = BITS; c.M = p.batch; c.nstages = p.nstages;
  c.maps = nullptr;
  c.smem = smem_raw;
  const uint32_t sb = smem_u32(smem_raw);
  c.ring_w = sb + p.smem_ring_w; c.ring_s = sb + p.smem_ring_s; c.ring_b = sb + p.smem_ring_b;
  c.x_addr = sb + p.smem_x; c.sumx_addr = sb + p.smem_sumx;
  c.xs = smem_raw + p.smem_x; c.sumx = (float*)(smem_raw + p.smem_sumx);
  c.slots = (float*)(smem_raw + p.smem_slots);
  c.x_ld = p.x_ld;
  c.tile_cnt = p.tile_cnt; c.tile_part = p.tile_part;
  init_ring(c, full, empty, slot_cnt, red);
  if (threadIdx.x < 32) {
    if (threadIdx.x == 0) {
      // single segment: feed the three grid-constant maps directly
      const GemvDesc& d = p.d;
      const uint64_t pol = l2_policy_evict_first();
      long long u0, u1;
      cta_range(d, phase_aligned(d), u0, u1);
      const int wbytes = stage_w_bytes(BITS), nbox = stage_boxes(BITS), stage_words = kStageK * BITS / 32;
      uint32_t cnt = 0;
      for (long long u = u0; u < u1; ++u, ++cnt) {
        const int slot = cnt % c.nstages;
        mbar_wait(&c.empty[slot], ((cnt / c.nstages) & 1) ^ 1);
        const int tile = (int)(u / d.chunks), chunk = (int)(u % d.chunks);
        uint64_t* bar = &c.full[slot];
        mbar_arrive_expect_tx(bar, wbytes + 2 * stage_sb_bytes());
        const uint32_t wdst = c.ring_w + (uint32_t)slot * wbytes;
        for (int b = 0; b < nbox; ++b)
          tma_load_2d(wdst + b * 2048, &map_w, bar, chunk * stage_words + b * 32, tile * kTileRows, pol);
        tma_load_2d(c.ring_s + (uint32_t)slot * stage_sb_bytes(), &map_s, bar, chunk * kStageGroups,
                    tile * kTileRows, pol);
        tma_load_2d(c.ring_b + (uint32_t)slot * stage_sb_bytes(), &map_b, bar, chunk * kStageGroups,
                    tile * kTileRows, pol);
      }
    }
    return;
  }
  EmbedSrc emb{nullptr, nullptr, nullptr, nullptr, nullptr};
  uint32_t cnt = 0;
  stage_x<BITS>(c, p.d, emb);
  consume_phase<BITS>(c, p.d, cnt, nullptr);
}

// ============================================================================================
// host-side launchers (called from pie_api.cu)
// ============================================================================================
template <typename K>
static cudaError_t set_smem(K kernel, int bytes) {
  return cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
}

cudaError_t launch_qmv(const QmvParams& p, const CUtensorMap& mw, const CUtensorMap& ms, const CUtensorMap& mb,
                       int grid, cudaStream_t stream) {
  cudaError_t e;
  if (p.bits == 4) {
    if ((e = set_smem(pie_qmv_kernel<4>, p.smem_total)) != cudaSuccess) return e;
    pie_qmv_kernel<4><<<grid, kThreads, p.smem_total, stream>>>(p, mw, ms, mb);
  } else {
    if ((e = set_smem(pie_qmv_kernel<8>, p.smem_total)) != cudaSuccess) return e;
    pie_qmv_kernel<8><<<grid, kThreads, p.smem_total, stream>>>(p, mw, ms, mb);
  }
  return cudaGetLastError();
}

template <int BITS, int D, int NREP>
static cudaError_t launch_decode_t(const EngineParams& p, int grid, cudaStream_t stream) {
  cudaError_t e = set_smem(pie_decode_kernel<BITS, D, NREP>, p.smem_total);
  if (e != cudaSuccess) return e;
  void* args[] = {(void*)&p};
  return cudaLaunchCooperativeKernel((const void*)pie_decode_kernel<BITS, D, NREP>, dim3(grid), dim3(kThreads),
                                     args, (size_t)p.smem_total, stream);
}

cudaError_t launch_decode(const EngineParams& p, int grid, cudaStream_t stream) {
  const int nrep = p.n_heads / p.n_kv;
#define PIE_DISPATCH(B, DD, R) \
  if (p.bits == B && p.head_dim == DD && nrep == R) return launch_decode_t<B, DD, R>(p, grid, stream);
  PIE_DISPATCH(4, 128, 4) PIE_DISPATCH(4, 64, 4) PIE_DISPATCH(4, 128, 8)
  PIE_DISPATCH(8, 128, 4) PIE_DISPATCH(8, 64, 4) PIE_DISPATCH(8, 128, 8)
#undef PIE_DISPATCH
  return cudaErrorInvalidConfiguration;
}

bool decode_supported(int bits, int head_dim, int nrep) {
  return (bits == 4 || bits == 8) && ((head_dim == 128 && (nrep == 4 || nrep == 8)) || (head_dim == 64 && nrep == 4));
}

}  // namespace pie
