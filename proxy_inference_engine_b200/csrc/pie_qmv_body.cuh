// Device code of the decode GEMV shared by the stand-alone kernel (pie_qmv.cu) and the fused decode step
// (pie_step.cu).  See pie_qmv.cu for the description of the kernel.
#pragma once
#include "pie_internal.h"
#include "pie_ptx.cuh"

#include <type_traits>

namespace pie {

template <int BITS>
struct QmvGeom {
  static constexpr int CK = kRowBytes * 8 / BITS;  // K elements per stage: 256 (q4) / 128 (q8)
  static constexpr int GPC = CK / kGroup;          // groups per stage: 4 / 2
  static constexpr int SB_BYTES = kTileRows * GPC * 2;  // scales (or biases) bytes per stage
  static constexpr int STAGE_BYTES = kTileRows * kRowBytes + 2 * SB_BYTES;
};

__host__ __device__ inline int qmv_xrow_bytes(int kslice) { return kslice * 2 + 16; }

// ---- swizzled position (bytes within the 128-byte stage row) of 16-byte chunk c of tile row r ----------
template <int BITS>
__device__ __forceinline__ int swz_chunk(int r, int c) {
  if (BITS == 4) return ((((c >> 1) ^ (r & 3)) << 1) | (c & 1)) * 16;  // 32-byte groups, 4 rows conflict-free
  return ((((c >> 2) ^ (r & 1)) << 2) | (c & 3)) * 16;                 // 64-byte groups, 2 rows conflict-free
}

// two bf16 values 128 + nibble: ((w >> sh) & 0x000f000f) | 0x43004300.  mask and magic arrive in registers (kernel
// arguments) so that the and-or is ONE LOP3; as immediates ptxas needs two.
__device__ __forceinline__ uint32_t lift4(uint32_t w, int sh, uint32_t mask, uint32_t magic) {
  uint32_t d;
  asm("lop3.b32 %0, %1, %2, %3, 0xEA;" : "=r"(d) : "r"(w >> sh), "r"(mask), "r"(magic));
  return d;
}

// ---- stage x (optionally RMS-normalised) into shared memory in B-fragment order + per-group sums ----
// Within every block of 16 k: q4 position 4m+i <-> k = m + 4i ; q8 position 4m+{0,1,2,3} <-> k = 4m+{0,2,1,3}.
template <int BITS, int MT>
__device__ void stage_x(const QmvArgs& a, int k0, int klen, uint8_t* xs, float* sumx, float* red) {
  const int tid = threadIdx.x, nthr = blockDim.x, lane = tid & 31;
  const int xrow = qmv_xrow_bytes(a.kslice);
  // fused RMSNorm (mx.fast.rms_norm, MLX/mlx/fast.cpp:93-111): inverse rms over the FULL row
  // M <= 2: every thread takes part in each row (one round trip per row).  More rows: a warp per row with 8 loads in
  // flight per lane, all rows concurrently, ONE barrier.  The inverse rms values end up in red[0 .. M).
  // Decode fast path (M <= 2, the whole row staged by this CTA, one 16-element block per thread): the block each thread
  // will stage is loaded ONCE, serves the sum of squares, and stays in registers for the second pass -- one L2 round
  // trip instead of two in front of every q/k/v and gate/up launch.
  const int bpr0 = klen / 16;
  const bool one_pass = a.norm_w && a.M <= 2 && klen == a.K && a.M * bpr0 <= nthr && (bpr0 & 31) == 0;
  uint4 keep0 = make_uint4(0, 0, 0, 0), keep1 = make_uint4(0, 0, 0, 0);
  if (one_pass) {
    float ss = 0.f;
    if (tid < a.M * bpr0) {
      const uint4* src = reinterpret_cast<const uint4*>(a.x + (size_t)(tid / bpr0) * a.x_ld + k0 + (tid % bpr0) * 16);
      keep0 = ldcg128(src); keep1 = ldcg128(src + 1);
      const uint32_t u[8] = {keep0.x, keep0.y, keep0.z, keep0.w, keep1.x, keep1.y, keep1.z, keep1.w};
#pragma unroll
      for (int j = 0; j < 8; ++j) { float lo = bf16lo(u[j]), hi = bf16hi(u[j]); ss += lo * lo + hi * hi; }
    }
    ss = warp_sum(ss);
    if (lane == 0) red[32 + (tid >> 5)] = ss;
    __syncthreads();
    if (tid < a.M) {
      const int wpr = bpr0 >> 5;  // warps per row
      float tot = 0.f;
      for (int w = 0; w < wpr; ++w) tot += red[32 + tid * wpr + w];
      red[tid] = 1.0f / sqrtf(tot / (float)a.K + a.eps);
    }
    __syncthreads();
  } else if (a.norm_w) {
    if (a.M <= 2) {
      float inv2[2] = {0.f, 0.f};
      for (int m = 0; m < a.M; ++m) {
        float ss = 0.f;
        const uint4* xr = reinterpret_cast<const uint4*>(a.x + (size_t)m * a.x_ld);
        for (int i = tid; i < a.K / 8; i += nthr) {
          uint4 v = ldcg128(xr + i);
          uint32_t u[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
          for (int j = 0; j < 4; ++j) { float lo = bf16lo(u[j]), hi = bf16hi(u[j]); ss += lo * lo + hi * hi; }
        }
        ss = warp_sum(ss);
        if (lane == 0) red[tid >> 5] = ss;
        __syncthreads();
        float tot = 0.f;
        for (int w = 0; w < (nthr >> 5); ++w) tot += red[w];
        inv2[m] = 1.0f / sqrtf(tot / (float)a.K + a.eps);
        __syncthreads();
      }
      if (tid < a.M) red[tid] = inv2[tid];
    } else {
      const int nch = a.K / 8;
      for (int m = tid >> 5; m < a.M; m += nthr >> 5) {
        float ss = 0.f;
        const uint4* xr = reinterpret_cast<const uint4*>(a.x + (size_t)m * a.x_ld);
        for (int i0 = lane; i0 < nch; i0 += 256) {
          uint4 v[8];
#pragma unroll
          for (int u = 0; u < 8; ++u) v[u] = i0 + 32 * u < nch ? ldcg128(xr + i0 + 32 * u) : make_uint4(0, 0, 0, 0);
#pragma unroll
          for (int u = 0; u < 8; ++u) {
            const uint32_t w4[4] = {v[u].x, v[u].y, v[u].z, v[u].w};
#pragma unroll
            for (int j = 0; j < 4; ++j) { float lo = bf16lo(w4[j]), hi = bf16hi(w4[j]); ss += lo * lo + hi * hi; }
          }
        }
        ss = warp_sum(ss);
        if (lane == 0) red[m] = 1.0f / sqrtf(ss / (float)a.K + a.eps);
      }
    }
    __syncthreads();
  }
  const int bpr = klen / 16;             // 16-element blocks per row in this slice (multiple of 4)
  const int total = a.M * bpr;
  const int rounds = (total + nthr - 1) / nthr;
  for (int it0 = 0; it0 < rounds; it0 += 4) {
   uint4 pre[4][2];  // the loads of four rounds are requested before the first is used
#pragma unroll
   for (int u = 0; u < 4; ++u) {
     const int idx = (it0 + u) * nthr + tid;
     if (one_pass && it0 + u == 0) {
       pre[u][0] = keep0; pre[u][1] = keep1;
     } else if (idx < total) {
       const uint4* src = reinterpret_cast<const uint4*>(a.x + (size_t)(idx / bpr) * a.x_ld + k0 + (idx % bpr) * 16);
       pre[u][0] = ldcg128(src); pre[u][1] = ldcg128(src + 1);
     }
   }
#pragma unroll
   for (int u = 0; u < 4; ++u) {
    const int it = it0 + u;
    if (it >= rounds) break;
    const int idx = it * nthr + tid;
    const bool ok = idx < total;
    const int m = ok ? idx / bpr : 0, blk = ok ? idx % bpr : 0;
    float v[16];
    if (ok) {
      const uint4 p0 = pre[u][0], p1 = pre[u][1];
      uint32_t u8[8] = {p0.x, p0.y, p0.z, p0.w, p1.x, p1.y, p1.z, p1.w};
#pragma unroll
      for (int j = 0; j < 8; ++j) { v[2 * j] = bf16lo(u8[j]); v[2 * j + 1] = bf16hi(u8[j]); }
      if (a.norm_w) {
        const uint4* wp = reinterpret_cast<const uint4*>(a.norm_w + k0 + blk * 16);
        uint4 w0 = __ldg(wp), w1 = __ldg(wp + 1);
        uint32_t wu[8] = {w0.x, w0.y, w0.z, w0.w, w1.x, w1.y, w1.z, w1.w};
        const float inv_m = red[m];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          v[2 * j] = bfround(bfround(v[2 * j] * inv_m) * bf16lo(wu[j]));
          v[2 * j + 1] = bfround(bfround(v[2 * j + 1] * inv_m) * bf16hi(wu[j]));
        }
      }
    } else {
#pragma unroll
      for (int j = 0; j < 16; ++j) v[j] = 0.f;
    }
    float s = 0.f;
#pragma unroll
    for (int j = 0; j < 16; ++j) s += v[j];
    s += __shfl_xor_sync(0xffffffffu, s, 1);
    s += __shfl_xor_sync(0xffffffffu, s, 2);
    if (ok) {
      uint32_t o[8];
#pragma unroll
      for (int mm = 0; mm < 4; ++mm) {
        if (BITS == 4) {
          o[2 * mm] = pack_bf16(v[mm], v[mm + 4]);
          o[2 * mm + 1] = pack_bf16(v[mm + 8], v[mm + 12]);
        } else {
          o[2 * mm] = pack_bf16(v[4 * mm], v[4 * mm + 2]);
          o[2 * mm + 1] = pack_bf16(v[4 * mm + 1], v[4 * mm + 3]);
        }
      }
      uint4* dst = reinterpret_cast<uint4*>(xs + (size_t)m * xrow + blk * 32);
      dst[0] = make_uint4(o[0], o[1], o[2], o[3]);
      dst[1] = make_uint4(o[4], o[5], o[6], o[7]);
      if ((blk & 3) == 0) sumx[((blk >> 2) * MT + (m >> 3)) * 8 + (m & 7)] = s;
    }
   }
  }
}

}  // namespace pie


// Main loop and epilogues of the persistent decode GEMV (included by pie_qmv.cu).
//
// One CTA per SM, W warps.  The CTA owns one slice of K (gridDim.x % ksplit slices, normally 1) and stages that
// slice of x ONCE.  Its warps form W / ks_w teams of ks_w warps; a team walks the tile sequence
//     tile = (team * ctas_per_slice + cta_lane) + i * (ctas_per_slice * nteams),  i = 0, 1, ...
// and each warp of the team streams its 1/ks_w share of the tile's K slice.  A warp's (tile, chunk) stages form
// one flattened sequence fed through a private cp.async ring that never drains at a tile boundary.
//
// End of a tile: no CTA or team barrier.  The team's warps take turns as the tile's FINISHER (tile i -> warp
// i % ks_w).  The others drop their partial sums (in MMA accumulator layout, one float4 per lane) into a
// double-buffered exchange slot and arrive on an mbarrier; the finisher waits for it, adds the partials in warp
// order (deterministic), applies the epilogue straight from registers and releases the slot.
namespace pie {

// `a` may live in constant (kernel parameter), shared or global memory.  `hook(phase)` runs after this warp's
// weight prefetch has been issued and before x is read: the stand-alone kernel waits there for the previous kernel
// (programmatic dependent launch); the fused decode step (pie_step.cu) runs its grid barriers and the attention
// phase there.  `reinit`: the mbarriers at the head of `smem` hold state of an earlier call and are invalidated first.
//
// MODE: QMV_FULL = the whole matrix (stand-alone kernel); QMV_PREFETCH = only issue this warp's first S - 1 ring
// stages and return; QMV_MAIN = the rest, for a caller that ran QMV_PREFETCH on the same descriptor (the producer
// state is re-derived, the cp.async groups are already committed).  The fused decode step splits the matrix this way
// so that the function holding the main loop contains no calls: anything that spills there goes to local memory
// that does not fit beside 200+ KB of shared memory in L1.
enum { QMV_FULL = 0, QMV_PREFETCH = 1, QMV_MAIN = 2 };
template <int BITS, int MT, bool ONECOL, bool COHERENT, int MODE, class Hook>
__device__ __forceinline__ void qmv_body(const QmvArgs& a_in, uint8_t* smem, bool reinit, Hook&& hook) {
  const QmvArgs& a = a_in;
  using G = QmvGeom<BITS>;
  constexpr int S = kQmvStages;
  constexpr int NS = kQmvSlots;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, W = blockDim.x >> 5;
  const int g = lane >> 2, t = lane & 3;
  const int ks_w = a.ks_w, nteams = W >> a.ks_w_log2, team = warp >> a.ks_w_log2, wt = warp & (ks_w - 1);
  int ksg = 0, cta_lane = blockIdx.x, ncta = gridDim.x;
  if (a.ksplit > 1) { ksg = blockIdx.x % a.ksplit; cta_lane = blockIdx.x / a.ksplit; ncta = gridDim.x / a.ksplit; }
  const int k0 = ksg * a.kslice;
  const int nch = a.nch;                      // stages per (warp, tile) = kslice / (CK * ks_w)
  const int kw0 = k0 + wt * nch * G::CK;      // first K element of this warp's share
  const int tstride = ncta * nteams;
  const int tile_first = team * ncta + cta_lane;  // CTAs first: a short tile list still spreads over every SM
  const int my_tiles = (tile_first < a.ntiles && cta_lane < ncta) ? (a.ntiles - 1 - tile_first) / tstride + 1 : 0;
  const int total = my_tiles * nch;

  // ---- shared-memory carve-up (must match qmv_smem_bytes) ----------------------------------------
  // [mbarriers 512 B][per-warp rings: W * S stages][x slice][group sums of x][exchange slots]; the rings sit at
  // offsets that do not depend on the matrix, so a fused caller can prefetch the next matrix while this one drains
  const int xrow = qmv_xrow_bytes(a.kslice);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem);                               // [nteams][2][NS]
  uint8_t* ring = smem + kQmvBarBytes + (size_t)warp * S * G::STAGE_BYTES;
  uint8_t* xs = smem + kQmvBarBytes + (size_t)W * S * G::STAGE_BYTES;
  float* sumx = reinterpret_cast<float*>(xs + (((size_t)a.M * xrow + 127) & ~(size_t)127));
  float4* exch = reinterpret_cast<float4*>(sumx + (a.kslice / kGroup) * 8 * MT);   // [W][NS][MT][32]
  const uint32_t ring_s = smem_u32(ring);
  uint64_t* full_bar = bars + team * 2 * NS;   // finisher waits: ks_w - 1 arrivals
  uint64_t* empty_bar = full_bar + NS;         // writers wait: 1 arrival (the finisher has read the slot)

  const int gpr = a.K / kGroup;  // groups per weight row
  // cp.async pieces of a stage: piece p = lane + 32 i -> tile row (lane >> 3) + 4 i, 16-byte chunk lane & 7.  The
  // swizzle only looks at row & 3, which does not depend on i: piece i lands 4 rows (512 B) after piece i - 1.
  const uint32_t wdst0 = (lane >> 3) * kRowBytes + swz_chunk<BITS>(lane >> 3, lane & 7);
  const uint32_t sbdst = kTileRows * kRowBytes + (lane >> 4) * G::SB_BYTES + (lane & 15) * G::GPC * 2;

  // tile -> (segment, first row); a dual tile is 8 gate rows + 8 up rows
  auto tile_seg = [&](int tile, int& sidx, int& nbase) {
    sidx = 0;
    if (a.dual) { nbase = tile * 8; return; }
    if (a.nseg > 1 && tile >= a.seg[1].tile0) sidx = 1;
    if (a.nseg > 2 && tile >= a.seg[2].tile0) sidx = 2;
    nbase = (tile - a.seg[sidx].tile0) * kTileRows;
  };

  // ---- producer state: running global pointers of the next stage to issue ---------------------------
  // pieces 0 and 2 carry 64-bit pointers; pieces 1 and 3 sit a 32-bit distance (4 rows, unless clamped) after them
  const uint8_t* wsrc[4];  // [1] and [3] unused
  uint32_t wdelta[2];
  const uint8_t* sbsrc;
  auto set_tile_ptrs = [&](int tile) {
    int sidx, nbase;
    tile_seg(tile, sidx, nbase);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int p = lane + 32 * i, r = p >> 3, c = p & 7;
      const int si = a.dual ? (r >> 3) : sidx;
      const int n = min(a.dual ? nbase + (r & 7) : nbase + r, a.seg[si].N - 1);
      const uint8_t* ptr = reinterpret_cast<const uint8_t*>(a.seg[si].wq) + ((size_t)n * a.K + kw0) * BITS / 8 + c * 16;
      if (i & 1) wdelta[i >> 1] = (uint32_t)(ptr - wsrc[i - 1]); else wsrc[i] = ptr;
    }
    const int r = lane & 15;
    const int si = a.dual ? (r >> 3) : sidx;
    const int n = min(a.dual ? nbase + (r & 7) : nbase + r, a.seg[si].N - 1);
    const uint16_t* base = (lane >> 4) ? a.seg[si].biases : a.seg[si].scales;
    sbsrc = reinterpret_cast<const uint8_t*>(base + (size_t)n * gpr + kw0 / kGroup);
  };
  int p_tile = tile_first, p_left = nch, p_todo = total;
  uint32_t p_soff = 0;
  if (total > 0) set_tile_ptrs(p_tile);
  auto produce = [&](bool copy) {
    if (p_todo > 0) {
      const uint32_t sbase = ring_s + p_soff;
      if (copy) {
#pragma unroll
        for (int i = 0; i < 4; ++i) cp_async16(sbase + wdst0 + i * 4 * kRowBytes, i & 1 ? wsrc[i - 1] + wdelta[i >> 1] : wsrc[i]);
        if (BITS == 4) cp_async8(sbase + sbdst, sbsrc); else cp_async4(sbase + sbdst, sbsrc);
      }
      wsrc[0] += kRowBytes;
      wsrc[2] += kRowBytes;
      sbsrc += G::GPC * 2;
      p_soff = p_soff == (S - 1) * G::STAGE_BYTES ? 0u : p_soff + G::STAGE_BYTES;
      --p_todo;
      if (--p_left == 0) {
        p_left = nch;
        p_tile += tstride;
        if (p_todo > 0) set_tile_ptrs(p_tile);
      }
    }
    if (copy) cp_async_commit();
  };
  // Fused caller (COHERENT): between two stages the five address words of the producer live in a per-thread shared
  // memory slot.  The main loop is a few registers over budget there (the descriptor is not in the constant bank),
  // and what ptxas would spill is exactly this state -- to local memory, whose write-through traffic to L2 costs
  // far more than two conflict-free LDS/STS bursts.
  const uint32_t prod_s = smem_u32(exch) + (uint32_t)W * NS * MT * 512 + tid * 4;
  auto prod_save = [&]() {
    if (!COHERENT) return;
    const uint64_t p0 = reinterpret_cast<uint64_t>(wsrc[0]), p2 = reinterpret_cast<uint64_t>(wsrc[2]),
                   p4 = reinterpret_cast<uint64_t>(sbsrc);
    const uint32_t st = blockDim.x * 4;
    sts32(prod_s, (uint32_t)p0); sts32(prod_s + st, (uint32_t)(p0 >> 32));
    sts32(prod_s + 2 * st, (uint32_t)p2); sts32(prod_s + 3 * st, (uint32_t)(p2 >> 32));
    sts32(prod_s + 4 * st, (uint32_t)p4); sts32(prod_s + 5 * st, (uint32_t)(p4 >> 32));
    sts32(prod_s + 6 * st, wdelta[0]); sts32(prod_s + 7 * st, wdelta[1]);
  };
  auto prod_load = [&]() {
    if (!COHERENT) return;
    const uint32_t st = blockDim.x * 4;
    wsrc[0] = reinterpret_cast<const uint8_t*>((uint64_t)lds32(prod_s) | ((uint64_t)lds32(prod_s + st) << 32));
    wsrc[2] = reinterpret_cast<const uint8_t*>((uint64_t)lds32(prod_s + 2 * st) | ((uint64_t)lds32(prod_s + 3 * st) << 32));
    sbsrc = reinterpret_cast<const uint8_t*>((uint64_t)lds32(prod_s + 4 * st) | ((uint64_t)lds32(prod_s + 5 * st) << 32));
    wdelta[0] = lds32(prod_s + 6 * st); wdelta[1] = lds32(prod_s + 7 * st);
  };
  auto issue_next = [&]() {
    prod_load();  // unconditional: a conditional reload would make the words loop-carried registers again
    produce(true);
    prod_save();
  };

  // ---- prologue: weights first (they never depend on the previous kernel), then x --------------------
#pragma unroll
  for (int s = 0; s < S - 1; ++s) produce(MODE != QMV_MAIN);
  if (MODE == QMV_PREFETCH) return;
  if (MODE == QMV_FULL) hook();
  prod_save();
  uint32_t* live_bars = reinterpret_cast<uint32_t*>(smem + kQmvBarBytes - 8);  // how many mbarriers the last call set up
  if (reinit) {  // every warp is past the earlier call here (the hook of a fused caller ends with a CTA barrier)
    if (tid < (int)*live_bars) mbar_inval(bars + tid);
    __syncthreads();
  }
  if (ks_w > 1 && wt == 0 && lane == 0) {
    for (int i = 0; i < NS; ++i) { mbar_init(full_bar + i, ks_w - 1); mbar_init(empty_bar + i, 1); }
    fence_mbar_init();
  }
  if (tid == 0) *live_bars = ks_w > 1 ? nteams * 2 * NS : 0;
  stage_x<BITS, MT>(a_in, k0, a.kslice, xs, sumx, reinterpret_cast<float*>(exch));
  __syncthreads();
#include "pie_qmv_loop.inl"
}

}  // namespace pie
