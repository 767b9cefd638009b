// Device-side pieces of the TMA-fed decode GEMV shared by the stand-alone kernel (pie_qmv2.cu) and the one-launch decode
// step (pie_step.cu): PTX wrappers, shared-memory geometry, x staging.  See pie_qmv2.cu for the description.
#pragma once
#include <cuda.h>
#include <cuda_fp16.h>

#include "pie_internal.h"
#include "pie_ptx.cuh"
#include "pie_qmv2.h"

namespace pie {

// ------------------------------------------------------------------------------------------- PTX helpers
__device__ __forceinline__ void tma_load_3d(uint32_t dst, const void* map, uint32_t bar, int c0, int c1, int c2,
                                            uint64_t pol) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes.L2::cache_hint"
      " [%0], [%1, {%3, %4, %5}], [%2], %6;" ::"r"(dst), "l"(map), "r"(bar), "r"(c0), "r"(c1), "r"(c2), "l"(pol)
      : "memory");
}
__device__ __forceinline__ void tma_load_2d_b(uint32_t dst, const void* map, uint32_t bar, int c0, int c1, uint64_t pol) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes.L2::cache_hint"
      " [%0], [%1, {%3, %4}], [%2], %5;" ::"r"(dst), "l"(map), "r"(bar), "r"(c0), "r"(c1), "l"(pol)
      : "memory");
}
__device__ __forceinline__ void tma_load_4d(uint32_t dst, const void* map, uint32_t bar, int c0, int c1, int c2, int c3,
                                            uint64_t pol) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.tile.mbarrier::complete_tx::bytes.L2::cache_hint"
      " [%0], [%1, {%3, %4, %5, %6}], [%2], %7;" ::"r"(dst), "l"(map), "r"(bar), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "l"(pol)
      : "memory");
}
__device__ __forceinline__ void mbar_expect_tx_s(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_init_s(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
// returns 0 through a value the compiler cannot see through (the loads of the stage are made to depend on it)
__device__ __forceinline__ uint32_t mbar_wait_s(uint32_t bar, uint32_t parity) {
  uint32_t spins = 0, ok = 0;
  while (true) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(bar), "r"(parity)
        : "memory");
    if (ok) break;
    if (++spins > PIE_SPIN_LIMIT) __trap();
  }
  return ok - 1;
}
// D(16x8) = A(16x8, bf16, row) * B(8x8, bf16, col) + C: a0 (g, 2t..2t+1), a1 (g+8, 2t..), b0 (k 2t..2t+1, n g).
// Not volatile, C separate from D: the compiler interleaves the independent accumulator chains of a unit.
__device__ __forceinline__ void mma_bf16_1688(float (&d)[4], uint32_t a0, uint32_t a1, uint32_t b0, const float (&c)[4]) {
  asm("mma.sync.aligned.m16n8k8.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5}, {%6}, {%7,%8,%9,%10};"
      : "=f"(d[0]), "=f"(d[1]), "=f"(d[2]), "=f"(d[3])
      : "r"(a0), "r"(a1), "r"(b0), "f"(c[0]), "f"(c[1]), "f"(c[2]), "f"(c[3]));
}
template <bool F16>
__device__ __forceinline__ void mma_16816(float (&d)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0,
                                          uint32_t b1, float c0, float c1, float c2, float c3) {
  if (F16)
    asm("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%10,%11,%12,%13};"
        : "=f"(d[0]), "=f"(d[1]), "=f"(d[2]), "=f"(d[3])
        : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1), "f"(c0), "f"(c1), "f"(c2), "f"(c3));
  else
    asm("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%10,%11,%12,%13};"
        : "=f"(d[0]), "=f"(d[1]), "=f"(d[2]), "=f"(d[3])
        : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1), "f"(c0), "f"(c1), "f"(c2), "f"(c3));
}
// Shared-memory loads that the compiler may schedule freely AFTER the value `tok` exists (tok = 0, produced by the
// mbarrier wait: nothing is hoisted above the wait; a later TMA issue is an asm volatile with a memory clobber).
__device__ __forceinline__ uint2 lds64_t(uint32_t addr, uint32_t tok) {
  uint2 v;
  asm("ld.shared.v2.u32 {%0,%1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(addr + tok));
  return v;
}
__device__ __forceinline__ uint4 lds128_t(uint32_t addr, uint32_t tok) {
  uint4 v;
  asm("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr + tok));
  return v;
}
__device__ __forceinline__ uint32_t lds32_t(uint32_t addr, uint32_t tok) {
  uint32_t v;
  asm("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(addr + tok));
  return v;
}
__device__ __forceinline__ uint32_t elect_one() {
  uint32_t pred;
  asm volatile("{\n\t.reg .pred P;\n\telect.sync _|P, 0xffffffff;\n\tselp.u32 %0, 1, 0, P;\n\t}" : "=r"(pred));
  return pred;
}
__device__ __forceinline__ uint32_t lift4b(uint32_t w, int sh, uint32_t mask, uint32_t magic) {
  uint32_t d;
  asm("lop3.b32 %0, %1, %2, %3, 0xEA;" : "=r"(d) : "r"(w >> sh), "r"(mask), "r"(magic));
  return d;
}

// ------------------------------------------------------------------------------------------- geometry
// Dynamic shared memory (base aligned to 1024 B):
//   [weight stages: warp w, stage s at (2 w + s) * 4096][scale/bias stages: (2 w + s) * 512: 256 B scales, 256 B biases]
//   [mbarriers 32 x 8 B][edge tile ids 32 x 4 B][reduction scratch 1 KB][-128 * group sums of x][bias-side pieces of the
//   group sums][x in fragment order]
constexpr int kQ2SbOff = kQ2Warps * kQ2Stages * kQ2WBytes;               // 131072
constexpr int kQ2BarOff = kQ2SbOff + kQ2Warps * kQ2Stages * kQ2SbBytes;  // + 16384
constexpr int kQ2EdgeOff = kQ2BarOff + kQ2Warps * kQ2Stages * 8;
constexpr int kQ2RedOff = kQ2EdgeOff + kQ2Warps * 2 * 4;
constexpr int kQ2VarOff = kQ2RedOff + 1024;

__host__ __device__ inline int q2_kpad(int K) { return (K + kQ2UnitK - 1) / kQ2UnitK * kQ2UnitK; }
__host__ __device__ inline int q2_xrow_bytes(int K) { return q2_kpad(K) * 2 + 16; }
__host__ __device__ inline int q2_gpitch(int K) { return q2_kpad(K) / kGroup + 8; }  // bf16 elements per pieces row

// ------------------------------------------------------------------------------------------- unit copies
// One unit (16 rows x 512 k: 4096 B of packed weights, 256 B of scales, 256 B of biases) into a stage: 2 copies when the
// arrays are co-located (see Qmv2Args), up to 6 otherwise.  d: the matrix's scalars, a: where its maps live.
__device__ __forceinline__ void q2_issue_unit(const Qmv2Args& d, const Qmv2Args& a, int tile, int kc, uint32_t wd, uint32_t sd,
                                              uint32_t bar, uint64_t pol) {
  mbar_expect_tx_s(bar, kQ2WBytes + kQ2SbBytes);
  if (d.dual) {
    const int r0 = tile * 8;
    if (d.gu_stacked) {
      tma_load_4d(wd, &a.wmap[0], bar, 0, r0, 2 * kc, 0, pol);
      tma_load_4d(sd, &a.smap[0], bar, 8 * kc, r0, 0, 0, pol);
      if (!d.sb_stacked) tma_load_4d(sd + 256, &a.bmap[0], bar, 8 * kc, r0, 0, 0, pol);
    } else {
      tma_load_4d(wd, &a.wmap[0], bar, 0, r0, 2 * kc, 0, pol);
      tma_load_4d(wd + 2048, &a.wmap[1], bar, 0, r0, 2 * kc, 0, pol);
      tma_load_4d(sd, &a.smap[0], bar, 8 * kc, r0, 0, 0, pol);
      tma_load_4d(sd + 128, &a.smap[1], bar, 8 * kc, r0, 0, 0, pol);
      tma_load_4d(sd + 256, &a.bmap[0], bar, 8 * kc, r0, 0, 0, pol);
      tma_load_4d(sd + 384, &a.bmap[1], bar, 8 * kc, r0, 0, 0, pol);
    }
  } else {
    int si = 0;
    if (d.nseg > 1 && tile >= d.tile0[1]) si = 1;
    if (d.nseg > 2 && tile >= d.tile0[2]) si = 2;
    const int r0 = (tile - d.tile0[si]) * 16;
    tma_load_4d(wd, &a.wmap[si], bar, 0, r0, 2 * kc, 0, pol);
    tma_load_4d(sd, &a.smap[si], bar, 8 * kc, r0, 0, 0, pol);
    if (!d.sb_stacked) tma_load_4d(sd + 256, &a.bmap[si], bar, 8 * kc, r0, 0, 0, pol);
  }
}

// ------------------------------------------------------------------------------------------- x staging
// x (optionally RMS-normalised: mx.fast.rms_norm, MLX/mlx/fast.cpp:93-111) -> shared memory in B-fragment order:
// within every block of 16 k, position 4 m + 2 h + l holds k = m + 8 h + 4 l (what lane t of the MMA needs for step m).
// Also per 64-element group: the accumulator's start value -OFF * sum(staged x) and sum(x) as three bf16 pieces.
//
// F16 = false: x is staged as the bf16 values themselves, weights lift to bf16 128 + q (OFF = 128).
// F16 = true : x is staged as fp16 after an exact power-of-two scaling of the row (max |x| lands in [2^14, 2^15), so
//   every element down to 2^-28 of the row maximum converts exactly; smaller ones keep subnormal precision and move the
//   dot product by less than an fp32 rounding), weights lift to fp16 1024 + q or 1024 + 16 q (two nibble positions
//   per 16-bit half without a shift; x at odd k is pre-divided by 16), OFF = 1024.  The row scale is undone when a tile
//   is finished.  Everything the MMA sees is exact in fp32 either way.
// MAXU: 16-element blocks a thread keeps in registers (rounds of the passes are unrolled MAXU-fold).  The code of this
// function is straight-line and large; q2_stage_x below picks the smallest MAXU that serves the shape so that a launch
// only pulls the instructions it needs through the instruction cache (K = 4096, one row: one block per thread).
template <bool ONECOL, bool F16, int MAXU>
__device__ __forceinline__ void q2_stage_x_impl(const Qmv2Args& a, uint8_t* var, float* red,
                                                unsigned long long* tr = nullptr) {
  // tr (profiling aid, one thread): %globaltimer after each of the four CTA barriers inside
  auto stamp = [&](int i) {
    if (tr) { unsigned long long t_; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_)); tr[i] = t_; }
  };
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  constexpr int NT = kQ2Warps * 32;
  constexpr int COLS = ONECOL ? 1 : 8;
  const int kpad = q2_kpad(a.K), G = kpad / kGroup, gp = q2_gpitch(a.K);
  constexpr int NSXB = ONECOL ? 16 : 64;  // bytes of start values per group: quads {n[2t], n[2t+1], n[2t], n[2t+1]}, t = 0..3
  float* nsx = reinterpret_cast<float*>(var);
  uint16_t* pieces = reinterpret_cast<uint16_t*>(var + (size_t)G * NSXB);
  uint8_t* xs = var + (size_t)G * NSXB + (((size_t)3 * COLS * gp * 2 + 15) & ~(size_t)15);
  auto put_nsx = [&](int grp, int m, float v) {
    if (ONECOL) {
      *reinterpret_cast<float4*>(nsx + grp * 4) = make_float4(v, v, v, v);
    } else {
      float* q = nsx + grp * 16 + (m >> 1) * 4 + (m & 1);
      q[0] = v; q[2] = v;
    }
  };
  const int xrow = q2_xrow_bytes(a.K);
  const int bpr = kpad / 16;           // 16-element blocks per row (multiple of 32)
  const int total = a.M * bpr;
  const int nits = (total + NT - 1) / NT;
  const bool keep = nits <= MAXU;         // every block of this thread stays in registers between the passes
  uint4 blk[MAXU][2];
  auto load_round = [&](int it0) {
#pragma unroll
    for (int u = 0; u < MAXU; ++u) {
      const int idx = (it0 + u) * NT + tid;
      blk[u][0] = blk[u][1] = make_uint4(0, 0, 0, 0);
      if (idx < total) {
        const int m = idx / bpr, b = idx - m * bpr;
        if (b * 16 < a.K) {
          const uint4* src = reinterpret_cast<const uint4*>(a.x + (size_t)m * a.x_ld + b * 16);
          blk[u][0] = ldcg128(src); blk[u][1] = ldcg128(src + 1);
        }
      }
    }
  };
  // Per-row reductions: entry e = (round, warp) holds the value of blocks [32 e, 32 e + 32), which lie in one row
  // (bpr % 32 == 0), so row m owns the entries [m * epr, (m + 1) * epr).  Warp m reduces row m: fixed order.
  const int epr = bpr / 32;
  if (a.norm_w) {
    for (int it0 = 0; it0 < nits; it0 += MAXU) {
      load_round(it0);
#pragma unroll
      for (int u = 0; u < MAXU; ++u) {
        if (it0 + u >= nits) break;
        const uint32_t w8[8] = {blk[u][0].x, blk[u][0].y, blk[u][0].z, blk[u][0].w,
                                blk[u][1].x, blk[u][1].y, blk[u][1].z, blk[u][1].w};
        float ss = 0.f;
#pragma unroll
        for (int j = 0; j < 8; ++j) { const float lo = bf16lo(w8[j]), hi = bf16hi(w8[j]); ss += lo * lo + hi * hi; }
        ss = warp_sum(ss);
        if (lane == 0) red[(it0 + u) * kQ2Warps + warp] = ss;
      }
    }
    __syncthreads();
    stamp(0);
    if (warp < a.M) {
      float tot = 0.f;
      for (int e = warp * epr + lane; e < (warp + 1) * epr; e += 32) tot += red[e];
      tot = warp_sum(tot);
      if (lane == 0) red[192 + warp] = 1.0f / sqrtf(tot / (float)a.K + a.eps);
    }
    __syncthreads();
    stamp(1);
  }
  // block u of this thread -> its 16 values (normalised when asked), as floats holding bf16 values
  auto values = [&](int u, int m, int b, bool live, float (&v)[16]) {
    const uint32_t w8[8] = {blk[u][0].x, blk[u][0].y, blk[u][0].z, blk[u][0].w,
                            blk[u][1].x, blk[u][1].y, blk[u][1].z, blk[u][1].w};
#pragma unroll
    for (int j = 0; j < 8; ++j) { v[2 * j] = bf16lo(w8[j]); v[2 * j + 1] = bf16hi(w8[j]); }
    if (a.norm_w && live) {
      const uint4* wp = reinterpret_cast<const uint4*>(a.norm_w + b * 16);
      const uint4 w0 = __ldg(wp), w1 = __ldg(wp + 1);
      const uint32_t wu[8] = {w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w};
      const float inv = red[192 + m];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        v[2 * j] = bfround(bfround(v[2 * j] * inv) * bf16lo(wu[j]));
        v[2 * j + 1] = bfround(bfround(v[2 * j + 1] * inv) * bf16hi(wu[j]));
      }
    }
  };
  if (F16) {
    // (nits <= MAXU here: the normalised blocks stay in registers, packed, while the row maxima are reduced)
    if (!a.norm_w) load_round(0);
#pragma unroll
    for (int u = 0; u < MAXU; ++u) {
      if (u >= nits) break;
      const int idx = u * NT + tid;
      const bool ok = idx < total;
      const int m = ok ? idx / bpr : 0, b = ok ? idx - m * bpr : 0;
      float v[16];
      values(u, m, b, ok && b * 16 < a.K, v);
      float amax = 0.f;
#pragma unroll
      for (int j = 0; j < 16; ++j) amax = fmaxf(amax, fabsf(v[j]));
      blk[u][0] = make_uint4(pack_bf16(v[0], v[1]), pack_bf16(v[2], v[3]), pack_bf16(v[4], v[5]), pack_bf16(v[6], v[7]));
      blk[u][1] = make_uint4(pack_bf16(v[8], v[9]), pack_bf16(v[10], v[11]), pack_bf16(v[12], v[13]), pack_bf16(v[14], v[15]));
      amax = warp_max(amax);
      if (lane == 0) red[u * kQ2Warps + warp] = amax;
    }
    __syncthreads();
    stamp(2);
    if (warp < a.M) {
      float mx = 0.f;
      for (int e = warp * epr + lane; e < (warp + 1) * epr; e += 32) mx = fmaxf(mx, red[e]);
      mx = warp_max(mx);
      // scale = 2^(14 - floor(log2 mx)); 1 for a zero (or denormal / non-finite) row
      const uint32_t ex = (__float_as_uint(mx) >> 23) & 0xffu;
      const bool sane = ex >= 16 && ex < 255;
      if (lane == 0) {
        red[200 + warp] = sane ? __uint_as_float((uint32_t)(127 + 14 + 127 - (int)ex) << 23) : 1.f;   // scale
        red[208 + warp] = sane ? __uint_as_float((uint32_t)((int)ex - 14) << 23) : 1.f;               // 1 / scale
      }
    }
    __syncthreads();
    stamp(3);
#pragma unroll
    for (int u = 0; u < MAXU; ++u) {
      if (u >= nits) break;
      const int idx = u * NT + tid;
      const bool ok = idx < total;
      const int m = ok ? idx / bpr : 0, b = ok ? idx - m * bpr : 0;
      const uint32_t w8[8] = {blk[u][0].x, blk[u][0].y, blk[u][0].z, blk[u][0].w,
                              blk[u][1].x, blk[u][1].y, blk[u][1].z, blk[u][1].w};
      float v[16];
#pragma unroll
      for (int j = 0; j < 8; ++j) { v[2 * j] = bf16lo(w8[j]); v[2 * j + 1] = bf16hi(w8[j]); }
      float s = 0.f;
#pragma unroll
      for (int j = 0; j < 16; ++j) s += v[j];
      const float sc = red[200 + m], sc16 = sc * 0.0625f;
      // what the MMA will see: fp16(v * scale), odd k also / 16; its exact sum starts the accumulator
      uint32_t h[8];  // h[j] = fp16 pair (k = 2j, 2j + 1)
      float sh = 0.f;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        __half2 hv = __floats2half2_rn(v[2 * j] * sc, v[2 * j + 1] * sc16);
        h[j] = *reinterpret_cast<uint32_t*>(&hv);
        const float2 back = __half22float2(hv);
        sh += back.x + back.y;
      }
      s += __shfl_xor_sync(0xffffffffu, s, 1);
      s += __shfl_xor_sync(0xffffffffu, s, 2);
      sh += __shfl_xor_sync(0xffffffffu, sh, 1);
      sh += __shfl_xor_sync(0xffffffffu, sh, 2);
      if (ok) {
        // position pair (4 mm + 2 hh, +1) holds k = (mm + 8 hh, mm + 8 hh + 4): halves of h[] regrouped by PRMT
        uint32_t o[8];
#pragma unroll
        for (int mm = 0; mm < 4; ++mm) {
#pragma unroll
          for (int hh = 0; hh < 2; ++hh) {
            const int k0 = mm + 8 * hh, k1 = k0 + 4;
            const uint32_t lo16 = (k0 & 1) ? (h[k0 >> 1] >> 16) : (h[k0 >> 1] & 0xffffu);
            const uint32_t hi16 = (k1 & 1) ? (h[k1 >> 1] >> 16) : (h[k1 >> 1] & 0xffffu);
            o[2 * mm + hh] = lo16 | (hi16 << 16);
          }
        }
        uint4* dst = reinterpret_cast<uint4*>(xs + (size_t)m * xrow + b * 32);
        dst[0] = make_uint4(o[0], o[1], o[2], o[3]);
        dst[1] = make_uint4(o[4], o[5], o[6], o[7]);
        if ((b & 3) == 0) {
          const int grp = b >> 2;
          put_nsx(grp, m, -1024.f * sh);
          const float hi = bfround(s), r1 = s - hi, mid = bfround(r1), lo = bfround(r1 - mid);
          pieces[(0 * COLS + m) * gp + grp] = (uint16_t)(__float_as_uint(hi) >> 16);
          pieces[(1 * COLS + m) * gp + grp] = (uint16_t)(__float_as_uint(mid) >> 16);
          pieces[(2 * COLS + m) * gp + grp] = (uint16_t)(__float_as_uint(lo) >> 16);
        }
      }
    }
  } else {
    for (int it0 = 0; it0 < nits; it0 += MAXU) {
      if (!(keep && a.norm_w)) load_round(it0);
#pragma unroll
      for (int u = 0; u < MAXU; ++u) {
        if (it0 + u >= nits) break;
        const int idx = (it0 + u) * NT + tid;
        const bool ok = idx < total;
        const int m = ok ? idx / bpr : 0, b = ok ? idx - m * bpr : 0;
        float v[16];
        values(u, m, b, ok && b * 16 < a.K, v);
        float s = 0.f;
#pragma unroll
        for (int j = 0; j < 16; ++j) s += v[j];
        s += __shfl_xor_sync(0xffffffffu, s, 1);
        s += __shfl_xor_sync(0xffffffffu, s, 2);
        if (ok) {
          uint32_t o[8];
#pragma unroll
          for (int mm = 0; mm < 4; ++mm) {
            o[2 * mm] = pack_bf16(v[mm], v[mm + 4]);
            o[2 * mm + 1] = pack_bf16(v[mm + 8], v[mm + 12]);
          }
          uint4* dst = reinterpret_cast<uint4*>(xs + (size_t)m * xrow + b * 32);
          dst[0] = make_uint4(o[0], o[1], o[2], o[3]);
          dst[1] = make_uint4(o[4], o[5], o[6], o[7]);
          if ((b & 3) == 0) {
            const int grp = b >> 2;
            put_nsx(grp, m, -128.f * s);
            const float hi = bfround(s), r1 = s - hi, mid = bfround(r1), lo = bfround(r1 - mid);
            pieces[(0 * COLS + m) * gp + grp] = (uint16_t)(__float_as_uint(hi) >> 16);
            pieces[(1 * COLS + m) * gp + grp] = (uint16_t)(__float_as_uint(mid) >> 16);
            pieces[(2 * COLS + m) * gp + grp] = (uint16_t)(__float_as_uint(lo) >> 16);
          }
        }
      }
    }
  }
  if (!ONECOL) {  // columns >= M of the MMA are never stored, but their start values must be finite
    for (int i = tid; i < G * 8; i += NT)
      if ((i & 7) >= a.M) put_nsx(i >> 3, i & 7, 0.f);
    for (int i = tid; i < 3 * COLS * gp; i += NT)
      if ((i / gp) % COLS >= a.M) pieces[i] = 0;
  }
}

// One activation row (the decode step), all 16 warps busy: a thread owns chunks of 8 consecutive k (16 bytes of x), the
// two threads of a 16-k block swap halves so that each stores 16 contiguous bytes of the fragment order, the 8 threads of
// a 64-k group add up its sums.  With the RMS norm and / or fp16 staging there is ONE CTA barrier inside: the first pass
// yields both reductions at once -- sum of squares and max |x * w| -- and every thread then adds the 16 per-warp partials
// itself (same order everywhere: identical, deterministic).  The row scale comes from the bound
//   max |v| <= max |x w| * inv_rms * (1 + 2^-8)^2   on the twice-rounded normalised values v
// so the scaled maximum lands in [2^13, 2^15), still exact in fp16.
// MAXC > 0: a thread's chunks stay in registers between the passes (K <= 4096 * MAXC); MAXC == 0: any K, x is read twice.
// The code is deliberately small (a launch runs it once, cold).
template <bool F16, int MAXC, class AfterLoads>
__device__ __forceinline__ void q2_stage_row(const Qmv2Args& a, uint8_t* var, float* red, AfterLoads after_loads,
                                             unsigned long long* tr = nullptr) {
  auto stamp = [&](int i) {   // profiling aid (one thread)
    if (tr) { unsigned long long t_; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_)); tr[i] = t_; }
  };
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  constexpr int NT = kQ2Warps * 32;
  constexpr int NC = MAXC > 0 ? MAXC : 1;
  const int kpad = q2_kpad(a.K), G = kpad / kGroup, gp = q2_gpitch(a.K);
  float* nsx = reinterpret_cast<float*>(var);
  uint16_t* pieces = reinterpret_cast<uint16_t*>(var + (size_t)G * 16);
  uint8_t* xs = var + (size_t)G * 16 + (((size_t)3 * gp * 2 + 15) & ~(size_t)15);
  const int nchunks = a.K / 8;   // K is a multiple of 512
  const bool norm = a.norm_w != nullptr;
  // Consumer side of a fused tensor-parallel exchange: every rank's o / down projection pushed its partial row into
  // slot r of this rank's buffer (pie_comm.cu has the layout); wait for the `world` flags of this exchange, then
  // x = residual + bf16(sum of the slots in rank order) -- the all_sum + `h + r` of the reference, in the x load.
  const bool tp = a.tp_mode == 2;
  unsigned int tp_set = 0;
  if (tp) {
    const unsigned int e = *reinterpret_cast<volatile unsigned int*>(a.tp_epoch);
    tp_set = e & 1u;
    if (tid < a.tp_world)
      wait_flag(reinterpret_cast<unsigned int*>(a.tp_peer[a.tp_rank] + tp_flags_off(a.tp_world, a.tp_max_elems)) +
                    ((size_t)tp_set * a.tp_world + tid) * kTpMaxCtas, e);
    __syncthreads();
  }
  auto load_x = [&](int c) -> uint4 {
    uint4 xv = ldcg128(a.x + (size_t)c * 8);
    if (tp) {
      float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
      const uint16_t* mine = reinterpret_cast<const uint16_t*>(a.tp_peer[a.tp_rank]) +
                             (size_t)tp_set * a.tp_world * a.tp_max_elems + (size_t)c * 8;
      for (int r = 0; r < a.tp_world; ++r) {
        const uint4 pv = ldcg128(mine + (size_t)r * a.tp_max_elems);
        const uint32_t pw[4] = {pv.x, pv.y, pv.z, pv.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) { acc[2 * j] += bf16lo(pw[j]); acc[2 * j + 1] += bf16hi(pw[j]); }
      }
      const uint32_t hw[4] = {xv.x, xv.y, xv.z, xv.w};
      uint32_t o[4];
#pragma unroll
      for (int j = 0; j < 4; ++j)
        o[j] = pack_bf16(bf16lo(hw[j]) + bfround(acc[2 * j]), bf16hi(hw[j]) + bfround(acc[2 * j + 1]));
      xv = make_uint4(o[0], o[1], o[2], o[3]);
      if (blockIdx.x == 0 && a.tp_h_out)   // the new residual stream, for the next exchange's consumer
        *reinterpret_cast<uint4*>(a.tp_h_out + (size_t)c * 8) = xv;
    }
    return xv;
  };
  uint4 xb[NC], wb[NC];
  float inv = 1.f, sc = 1.f;
  {
    float ss = 0.f, amax = 0.f;
    auto reduce_chunk = [&](const uint4& xv, const uint4& wv) {
      const uint32_t x4[4] = {xv.x, xv.y, xv.z, xv.w}, w4[4] = {wv.x, wv.y, wv.z, wv.w};
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float lo = bf16lo(x4[j]), hi = bf16hi(x4[j]);
        ss = fmaf(lo, lo, ss); ss = fmaf(hi, hi, ss);
        if (norm) amax = fmaxf(amax, fmaxf(fabsf(lo * bf16lo(w4[j])), fabsf(hi * bf16hi(w4[j]))));
        else amax = fmaxf(amax, fmaxf(fabsf(lo), fabsf(hi)));
      }
    };
    if (MAXC > 0) {
#pragma unroll
      for (int u = 0; u < NC; ++u) {
        const int c = u * NT + tid;
        xb[u] = wb[u] = make_uint4(0, 0, 0, 0);
        if (c < nchunks) {
          xb[u] = load_x(c);
          if (norm) wb[u] = __ldg(reinterpret_cast<const uint4*>(a.norm_w) + c);
        }
      }
      after_loads();
      if (norm || F16) {
#pragma unroll
        for (int u = 0; u < NC; ++u) reduce_chunk(xb[u], wb[u]);
      }
      stamp(9 + (ss == 123456.f));
    } else {
      after_loads();
      if (norm || F16) {
        for (int c = tid; c < nchunks; c += NT) {
          const uint4 xv = load_x(c);
          uint4 wv = make_uint4(0, 0, 0, 0);
          if (norm) wv = __ldg(reinterpret_cast<const uint4*>(a.norm_w) + c);
          reduce_chunk(xv, wv);
        }
      }
    }
    if (norm || F16) {
      ss = warp_sum(ss);
      amax = warp_max(amax);
      if (lane == 0) { red[warp] = ss; red[32 + warp] = amax; }
      __syncthreads();
      float tot = 0.f, mx = 0.f;
      stamp(10 + (red[0] == 123456.f));
#pragma unroll
      for (int w4 = 0; w4 < kQ2Warps / 4; ++w4) {
        const float4 sv = *reinterpret_cast<const float4*>(red + 4 * w4), mv = *reinterpret_cast<const float4*>(red + 32 + 4 * w4);
        tot += sv.x; tot += sv.y; tot += sv.z; tot += sv.w;
        mx = fmaxf(fmaxf(mx, mv.x), fmaxf(fmaxf(mv.y, mv.z), mv.w));
      }
      if (norm) { inv = 1.0f / sqrtf(tot / (float)a.K + a.eps); mx = mx * inv * 1.0079f; }
      if (F16) {
        const uint32_t ex = (__float_as_uint(mx) >> 23) & 0xffu;
        const bool sane = ex >= 16 && ex < 255;
        sc = sane ? __uint_as_float((uint32_t)(127 + 14 + 127 - (int)ex) << 23) : 1.f;
        if (tid == 0) red[208] = sane ? __uint_as_float((uint32_t)((int)ex - 14) << 23) : 1.f;   // 1 / scale, read after the caller's barrier
      }
    }
  }
  const float sc16 = sc * 0.0625f;
  stamp(11 + (sc16 == 123456.f));
  // second pass: normalise, convert, store in fragment order, group sums
  auto emit_chunk = [&](int c, const uint4& xv, const uint4& wv) {   // every lane of the warp calls this (c may be out of range)
    const bool ok = c < nchunks;
    const uint32_t x4[4] = {xv.x, xv.y, xv.z, xv.w}, w4[4] = {wv.x, wv.y, wv.z, wv.w};
    float v[8];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      float lo = bf16lo(x4[j]), hi = bf16hi(x4[j]);
      if (norm) {
        const uint32_t t1 = pack_bf16(lo * inv, hi * inv);
        const uint32_t t2 = pack_bf16(bf16lo(t1) * bf16lo(w4[j]), bf16hi(t1) * bf16hi(w4[j]));
        lo = bf16lo(t2); hi = bf16hi(t2);
      }
      v[2 * j] = lo; v[2 * j + 1] = hi;
    }
    float s = 0.f, sh = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) s += v[j];
    // word j of this thread = the pair (k = 8 h + j, 8 h + j + 4), j < 4
    uint32_t w[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      if (F16) {
        const float scj = (j & 1) ? sc16 : sc;   // odd k are pre-divided by 16 (their nibbles lift to 1024 + 16 q)
        __half2 hv = __floats2half2_rn(v[j] * scj, v[j + 4] * scj);
        w[j] = *reinterpret_cast<uint32_t*>(&hv);
        const float2 back = __half22float2(hv);
        sh += back.x + back.y;
      } else {
        w[j] = pack_bf16(v[j], v[j + 4]);
      }
    }
    const int h = c & 1;
    const uint32_t ra = __shfl_xor_sync(0xffffffffu, h ? w[0] : w[2], 1), rb = __shfl_xor_sync(0xffffffffu, h ? w[1] : w[3], 1);
    float gs = F16 ? sh : s;   // what the accumulator's start value is built from
    s += __shfl_xor_sync(0xffffffffu, s, 1);
    s += __shfl_xor_sync(0xffffffffu, s, 2);
    s += __shfl_xor_sync(0xffffffffu, s, 4);
    if (F16) {
      gs += __shfl_xor_sync(0xffffffffu, gs, 1);
      gs += __shfl_xor_sync(0xffffffffu, gs, 2);
      gs += __shfl_xor_sync(0xffffffffu, gs, 4);
    } else {
      gs = s;
    }
    if (ok) {
      *reinterpret_cast<uint4*>(xs + (size_t)c * 16) = h ? make_uint4(ra, w[2], rb, w[3]) : make_uint4(w[0], ra, w[1], rb);
      if ((c & 7) == 0) {
        const int grp = c >> 3;
        const float n = (F16 ? -1024.f : -128.f) * gs;
        *reinterpret_cast<float4*>(nsx + grp * 4) = make_float4(n, n, n, n);
        const float hi = bfround(s), r1 = s - hi, mid = bfround(r1), lo = bfround(r1 - mid);
        pieces[0 * gp + grp] = (uint16_t)(__float_as_uint(hi) >> 16);
        pieces[1 * gp + grp] = (uint16_t)(__float_as_uint(mid) >> 16);
        pieces[2 * gp + grp] = (uint16_t)(__float_as_uint(lo) >> 16);
      }
    }
  };
  if (MAXC > 0) {
#pragma unroll
    for (int u = 0; u < NC; ++u) {
      if (u * NT >= nchunks) break;   // CTA-uniform
      emit_chunk(u * NT + tid, xb[u], wb[u]);
    }
  } else {
    for (int c0 = 0; c0 < nchunks; c0 += NT) {   // nchunks is a multiple of 64: whole warps are in or out of range
      const int c = c0 + tid;
      uint4 xv = make_uint4(0, 0, 0, 0), wv = make_uint4(0, 0, 0, 0);
      if (c < nchunks) {
        xv = load_x(c);
        if (norm) wv = __ldg(reinterpret_cast<const uint4*>(a.norm_w) + c);
      }
      emit_chunk(c, xv, wv);
    }
  }
}

struct Q2Nothing { __device__ void operator()() const {} };
template <bool ONECOL, bool F16, class AfterLoads = Q2Nothing>
__device__ __forceinline__ void q2_stage_x(const Qmv2Args& a, uint8_t* var, float* red, AfterLoads after_loads = AfterLoads(),
                                           unsigned long long* tr = nullptr) {
  if (ONECOL) {   // the decode step's case
    if (a.K <= 4096) q2_stage_row<F16, 1>(a, var, red, after_loads, tr);
    else if (a.K <= 16384) q2_stage_row<F16, 4>(a, var, red, after_loads, tr);   // (all loads of x in flight at once)
    else q2_stage_row<F16, 0>(a, var, red, after_loads, tr);
    return;
  }
  after_loads();
  const int nits = (a.M * (q2_kpad(a.K) / 16) + kQ2Warps * 32 - 1) / (kQ2Warps * 32);
  if (nits <= 1) q2_stage_x_impl<ONECOL, F16, 1>(a, var, red, tr);
  else if (nits <= 2) q2_stage_x_impl<ONECOL, F16, 2>(a, var, red, tr);
  else q2_stage_x_impl<ONECOL, F16, 4>(a, var, red, tr);
}

}  // namespace pie
