// TMA-fed decode GEMV for 4-bit weights, 1 <= M <= 8 activation rows, sm_100a (the decode step's dominant kernel).
//
// Replaces mx.quantized_matmul(transpose=True) at decode sizes (QuantizedLinear.__call__,
// MLX/python/mlx/nn/layers/quantized.py:204-216 -> Metal qmv_fast, MLX/mlx/backend/metal/kernels/quantized.h:623-689):
//     y[m, n] = bf16( sum_k x[m, k] * (scales[n, k/64] * q[n, k] + biases[n, k/64]) ),  fp32 accumulation.
//
// Shape of the kernel (one CTA per SM, 16 warps)
//   * work unit = 16 weight rows x 512 k: 4 KB of packed weights + 256 B of bf16 scales + 256 B of bf16 biases.  The
//     CTA owns a contiguous range of whole 16-row tiles; the tiles' units, flattened tile-major, are cut into 16
//     contiguous runs, one per warp (a tile's K is split over warps whenever the CTA has fewer tiles than warps);
//   * every warp owns a private two-stage ring fed by TMA: one elected lane issues three cp.async.bulk.tensor copies
//     per unit (weights as a 3-D box 128 B x 16 rows x 2 chunks in the 32-byte-atom 128-byte swizzle, so that the
//     fragment loads below are bank-conflict free; scales and biases as 16 B x 16 row boxes) that complete on the
//     stage's mbarrier (complete_tx).  No cross-warp synchronisation in the main loop, no per-lane address state;
//   * dequantisation in registers: ((w >> 4i) & 0x000f000f) | 0x43004300 = two bf16 values 128 + q, fed to
//     mma.sync.m16n8k16 as the A operand (16 weight rows x 16 k); the activations are the B operand (batch rows as
//     columns), staged once per CTA in fragment order.  The accumulator of a 64-element group starts at
//     -128 * sum_x(group), so that it ends as sum_k q x exactly as in fp32; y += scale * acc;
//   * the bias term sum_g biases[n, g] * sum_x(g) is itself a small dense product: one mma.sync.m16n8k8 per unit on
//     the bf16 biases as they lie in shared memory against sum_x split exactly into three bf16 pieces
//     (the factoring s * sum(x q) + b * sum(x) of qmv_fast, kernels/quantized.h:225);
//   * a tile whose K was cut over several warps is summed in warp (= k) order through shared memory at the end of the
//     kernel: deterministic.  Epilogues from registers: store / residual add / SwiGLU (tiles of 8 gate + 8 up rows);
//   * programmatic dependent launch: the first two units of every warp are requested before griddepcontrol.wait, x is
//     read after it.  (Asking L2 for the first units of the NEXT matrix of the step with cp.async.bulk.prefetch.tensor
//     once a warp's own loads were out was tried and measured slower by 0.5 - 1.8 us per launch: removed.)
#include <cuda.h>
#include <cuda_fp16.h>

#include <cstdlib>
#include <cstring>
#include <mutex>
#include <unordered_map>

#include "pie_internal.h"
#include "pie_ptx.cuh"
#include "pie_qmv2.h"
#include "pie_qmv2_dev.cuh"

namespace pie {

size_t qmv2_smem_bytes(int M, int K) {
  const int cols = M == 1 ? 1 : 8;
  const size_t nsx = (size_t)(q2_kpad(K) / kGroup) * (cols == 1 ? 16 : 64);  // start values as register quads
  const size_t pieces = ((size_t)3 * cols * q2_gpitch(K) * 2 + 15) & ~(size_t)15;
  const size_t xs = (size_t)M * q2_xrow_bytes(K);
  return 1024 /* alignment slack */ + kQ2VarOff + nsx + pieces + xs;
}

// ------------------------------------------------------------------------------------------- the kernel
template <bool ONECOL, bool F16, bool DUAL>
__global__ void __launch_bounds__(kQ2Warps * 32, 1) qmv2_kernel(const __grid_constant__ Qmv2Args a) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  const int tid = threadIdx.x, lane = tid & 31;
  // `warp` is the warp's SLOT in the CTA's run of units, not its hardware number: slots 4q .. 4q+3 are the four warps of
  // scheduler q (hardware warp w issues on scheduler w % 4).  The run is cut into 16 consecutive shares of n or n + 1
  // units; with slot == hardware warp the n + 1 shares of a x.5 split (q/k/v: 24 units, down: 56, gate/up: 104) all land on
  // schedulers 1 and 3, and the main loop is issue-bound per scheduler: 8 / 16 / 28 units there against 6 / 14 / 26 when
  // every scheduler gets two of each kind.  (PIE_QMV2_DBG & 64: slot = hardware warp.)
  const int hw_warp = __shfl_sync(0xffffffffu, tid >> 5, 0);  // warp-uniform for the compiler: TMA operands stay in uniform registers
  const int warp = (a.dbg & 64) ? hw_warp : (((hw_warp & 3) << 2) | (hw_warp >> 2));
  const int g = lane >> 2, t = lane & 3;
  constexpr int COLS = ONECOL ? 1 : 8;
#define Q2_STAMP(i)                                                                        \
  if (a.stamps && tid == 0) {                                                              \
    unsigned long long t_;                                                                 \
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_));                                 \
    a.stamps[blockIdx.x * 16 + (i)] = t_;                                                  \
  }
#define Q2_STAMP_ORDERED(i) /* not to be moved across the CTA barrier next to it */       \
  if (a.stamps && tid == 0) {                                                              \
    unsigned long long t_;                                                                 \
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_)::"memory");                       \
    a.stamps[blockIdx.x * 16 + (i)] = t_;                                                  \
  }
  Q2_STAMP(0)

  // ---- this CTA's tiles and this warp's run of units -------------------------------------------------
  // (no divisions here: this prologue is on the launch's critical path)
  const int c = blockIdx.x;
  const int tile_lo = c * a.tiles_q + min(c, a.tiles_r), tile_hi = tile_lo + a.tiles_q + (c < a.tiles_r ? 1 : 0);
  const int upt = a.upt;
  const int cu = (tile_hi - tile_lo) * upt;                 // units of this CTA
  const int ub = tile_lo * upt + ((warp * cu) >> 4);
  const int ue = tile_lo * upt + (((warp + 1) * cu) >> 4);
  static_assert(kQ2Warps == 16, "the warp split above shifts by 4");

  const uint32_t smem_s = smem_u32(smem);
  const uint32_t wst = smem_s + (uint32_t)(warp * kQ2Stages) * kQ2WBytes;
  const uint32_t sbst = smem_s + kQ2SbOff + (uint32_t)(warp * kQ2Stages) * kQ2SbBytes;
  const uint32_t bar0 = smem_s + kQ2BarOff + (uint32_t)(warp * kQ2Stages) * 8;
  int* edge_tile = reinterpret_cast<int*>(smem + kQ2EdgeOff);
  float* red = reinterpret_cast<float*>(smem + kQ2RedOff);
  uint8_t* var = smem + kQ2VarOff;

  const uint64_t pol = l2_policy_evict_first();
  // (tile, k chunk) -> the unit's TMA copies into stage st
  auto issue = [&](int tile, int kc, int st) {
    q2_issue_unit(a, a, tile, kc, wst + st * kQ2WBytes, sbst + st * kQ2SbBytes, bar0 + st * 8, pol);
  };
  // producer cursor (lane 0 only): the next unit to request
  const int p_tile0 = upt == 1 ? ub : (int)__umulhi((unsigned)ub, a.upt_inv);
  int p_u = ub, p_tile = p_tile0, p_kc = ub - p_tile * upt;
  auto advance = [&]() {
    ++p_u;
    if (++p_kc == upt) { p_kc = 0; ++p_tile; }
  };

  // the TMA unit fetches a descriptor on first use (cold: every launch has its own parameter buffer): ask for all of
  // them at once, from different lanes, before anything else
  if (warp == 0 && lane < 9) {
    const int i = lane / 3, kind = lane - 3 * i;
    if (i < (DUAL ? (a.gu_stacked ? 1 : 2) : a.nseg) && !(kind == 2 && a.sb_stacked))
      tma_prefetch_desc(kind == 0 ? &a.wmap[i] : kind == 1 ? &a.smap[i] : &a.bmap[i]);
  }
  if (lane == 0) {
    mbar_init_s(bar0, 1);
    mbar_init_s(bar0 + 8, 1);
    fence_mbar_init();
    fence_proxy_async();
    edge_tile[warp * 2] = -1;
    edge_tile[warp * 2 + 1] = -1;
  }
  // First loads: stage 0 before the dependency wait, stage 1 right behind the loads of x.  Measured on the 8B step
  // (tools/timeline_step.py, us per layer): this order 47.8-49.3; both stages before the wait 49.7 (when all CTAs of a
  // launch start together -- gate/up behind o -- 148 x 16 x 9 KB of weight requests queue in front of the loads of x every
  // CTA is about to wait for); both behind the loads of x 52.5 (the elected lane's copies -- cold descriptors -- then sit
  // in every warp's staging path).  PIE_QMV2_DBG: 8 = both before the wait, 16 = both behind x, 4 = both after the staging.
  Q2_STAMP(8)
  const int pf_mode = a.dbg & (4 | 8 | 16);
  if (pf_mode == 0 || pf_mode == 8) {
    if (p_u < ue) { if (elect_one()) issue(p_tile, p_kc, 0); advance(); }
    if (pf_mode == 8 && p_u < ue) { if (elect_one()) issue(p_tile, p_kc, 1); advance(); }
  }
  auto after_x_loads = [&]() {
    if (pf_mode == 16 && p_u < ue) { if (elect_one()) issue(p_tile, p_kc, 0); advance(); }
    if ((pf_mode == 0 || pf_mode == 16) && p_u < ue) { if (elect_one()) issue(p_tile, p_kc, 1); advance(); }
  };
  Q2_STAMP(1)
  pdl_launch_dependents();
  pdl_wait();
  Q2_STAMP(2)
  unsigned int tp_e = 0, tp_set = 0;
  if (ONECOL && a.tp_mode == 1) { tp_e = *reinterpret_cast<volatile unsigned int*>(a.tp_epoch); tp_set = tp_e & 1u; }
  q2_stage_x<ONECOL, F16>(a, var, red, after_x_loads, a.stamps && tid == 0 ? a.stamps + blockIdx.x * 16 : nullptr);
  __syncthreads();
  Q2_STAMP(3)
  if (a.dbg & 4) {  // experiment: first loads only after the activations are staged
    if (p_u < ue) { if (elect_one()) issue(p_tile, p_kc, 0); advance(); }
    if (p_u < ue) { if (elect_one()) issue(p_tile, p_kc, 1); advance(); }
  }

  // ---- per-lane constants -------------------------------------------------------------------------
  const int kpad = q2_kpad(a.K), Gp = kpad / kGroup, gp = q2_gpitch(a.K);
  const uint32_t nsx_s = smem_u32(var);
  const uint32_t pcs_s = nsx_s + (uint32_t)Gp * (ONECOL ? 16 : 64);
  const uint32_t xs_s = pcs_s + (uint32_t)((3 * COLS * gp * 2 + 15) & ~15);
  const int xrow = q2_xrow_bytes(a.K);
  const int col = ONECOL ? 0 : (g < a.M ? g : 0);
  const uint32_t x_lane = xs_s + (uint32_t)col * xrow + t * 32;
  const uint32_t nsx_lane = ONECOL ? nsx_s : nsx_s + t * 16;
  const uint32_t pcs_lane = pcs_s + (uint32_t)((ONECOL ? 0 : g) * gp + 2 * t) * 2;
  constexpr uint32_t hs = DUAL ? 2048u : 1024u, cs = DUAL ? 1024u : 2048u;  // strides: row half, 128-byte chunk
  uint32_t woff[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) woff[j] = g * 128 + ((j ^ (g & 3)) << 5) + t * 8;
  const uint32_t lmask = a.lift_mask, lmask2 = a.lift_mask2, lmagic = a.lift_magic;

  float rs[2] = {1.f, 1.f};  // F16: 1 / row scale of activation rows 2t, 2t + 1
  if (F16) { rs[0] = red[208 + (ONECOL ? 0 : 2 * t)]; rs[1] = red[208 + (ONECOL ? 0 : 2 * t + 1)]; }
  float ys[4] = {0.f, 0.f, 0.f, 0.f}, yb[4] = {0.f, 0.f, 0.f, 0.f};
  float e0v[4] = {0.f, 0.f, 0.f, 0.f};
  int e0_tile = -1;
  uint32_t res_raw[4] = {0, 0, 0, 0};

  auto seg_of = [&](int tile, int& si, int& nbase) {
    si = 0;
    if (DUAL) { nbase = tile * 8; return; }
    if (a.nseg > 1 && tile >= a.tile0[1]) si = 1;
    if (a.nseg > 2 && tile >= a.tile0[2]) si = 2;
    nbase = (tile - a.tile0[si]) * 16;
  };
  auto load_residual = [&](int tile) {
    int si, nbase;
    seg_of(tile, si, nbase);
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int m = 2 * t + (q & 1), n = nbase + g + (q >> 1) * 8;
      res_raw[q] = 0;
      if (m < a.M && n < a.N[si] && !(ONECOL && ((q & 1) || t)))
        res_raw[q] = (uint32_t)a.residual[(size_t)m * a.out_ld + a.out_off[si] + n];
    }
  };
  // epilogue from registers: lane (g, t) holds rows g, g + 8 x activation rows 2t, 2t + 1
  auto epilogue = [&](int tile, const float (&v)[4]) {
    int si, nbase;
    seg_of(tile, si, nbase);
    if (ONECOL && a.tp_mode == 1) {
      // fused exchange, producer: the tile's 16 partial sums (rows g in v[0], rows g + 8 in v[2] of the lanes t == 0) are
      // gathered into lane 0 and go to every peer's slot as two 16-byte stores (2-byte stores over NVLink are slow)
      const uint32_t lo = f2bf(v[0]), hi = f2bf(v[2]);
      uint32_t w[8];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        w[i] = __shfl_sync(0xffffffffu, lo, 8 * i) | (__shfl_sync(0xffffffffu, lo, 8 * i + 4) << 16);
        w[4 + i] = __shfl_sync(0xffffffffu, hi, 8 * i) | (__shfl_sync(0xffffffffu, hi, 8 * i + 4) << 16);
      }
      if (lane == 0) {
        for (int p = 0; p < a.tp_world; ++p) {
          uint4* dst = reinterpret_cast<uint4*>(reinterpret_cast<uint16_t*>(a.tp_peer[p]) +
                                                ((size_t)tp_set * a.tp_world + a.tp_rank) * a.tp_max_elems + a.out_off[si] + nbase);
          dst[0] = make_uint4(w[0], w[1], w[2], w[3]);
          dst[1] = make_uint4(w[4], w[5], w[6], w[7]);
        }
      }
      return;
    }
#pragma unroll
    for (int cc = 0; cc < (ONECOL ? 1 : 2); ++cc) {
      const int m = 2 * t + cc;
      if (m >= a.M) continue;
      if (DUAL) {  // SwiGLU: row g = gate, row g + 8 = up of the same output column
        const int n = nbase + g;
        if (n < a.N[0]) {
          const float gt = bfround(v[cc]), up = bfround(v[2 + cc]);
          const float sg = bfround(1.f / (1.f + expf(-gt)));  // nn.silu = x * sigmoid(x), each op rounded
          a.out[(size_t)m * a.out_ld + n] = f2bf(bfround(gt * sg) * up);
        }
      } else {
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          const int n = nbase + g + 8 * h;
          if (n < a.N[si]) {
            float r = bfround(v[2 * h + cc]);
            if (a.epi == EPI_RESADD) r = bf16lo(res_raw[2 * h + cc]) + r;
            a.out[(size_t)m * a.out_ld + a.out_off[si] + n] = f2bf(r);
          }
        }
      }
    }
  };

  // ---- main loop ---------------------------------------------------------------------------------------
  int run_start_kc = 0;  // k chunk at which this warp's share of the current tile began
  int tile = p_tile0, kc = ub - tile * upt;
  for (int u = ub; u < ue; ++u) {
    const int i = u - ub, st = i & 1;
    if (i == 0 || kc == 0) {
      run_start_kc = kc;
#pragma unroll
      for (int q = 0; q < 4; ++q) ys[q] = yb[q] = 0.f;
    }
    const bool tile_ends = kc == upt - 1 || u == ue - 1;
    const uint32_t tok = mbar_wait_s(bar0 + st * 8, (i >> 1) & 1);
    if (i == 0) { Q2_STAMP(4) }
    const uint32_t wb = wst + st * kQ2WBytes, sb = sbst + st * kQ2SbBytes;
    if (!(a.dbg & 1)) {
      const uint32_t sp = sb + tok;
      // two 128-byte chunks of every row, four groups each
#pragma unroll
      for (int cch = 0; cch < 2; ++cch) {
      const uint32_t wp = wb + tok + cch * cs;
      const uint2 sA = lds64_t(sp + g * 16 + cch * 8, 0), sB = lds64_t(sp + 128 + g * 16 + cch * 8, 0);
      const uint32_t sAw[2] = {sA.x, sA.y}, sBw[2] = {sB.x, sB.y};
      const uint32_t xp = x_lane + (uint32_t)(kc * 8 + cch * 4) * 128 + tok;
      const uint32_t np = nsx_lane + (uint32_t)(kc * 8 + cch * 4) * (ONECOL ? 16 : 64) + tok;
#pragma unroll
      for (int jj = 0; jj < 4; ++jj) {
        const int j = jj;
        const uint2 w0 = lds64_t(wp + woff[j], 0);
        const uint2 w1 = lds64_t(wp + woff[j] + hs, 0);
        // the accumulator's start value arrives as a register quad {n0, n1, n0, n1}: no moves in front of the first MMA
        uint4 x0, x1, nq;
        {  // (loading x only in the 4 lanes that feed column 0 -- ptxas turns the predicated loads into divergent branches --
           //  and the start value as a 4-byte broadcast: measured 561 vs 629 tok/s on the 8B step; the loop runs at HBM rate as it is)
          x0 = lds128_t(xp + jj * 128, 0); x1 = lds128_t(xp + jj * 128 + 16, 0);
          nq = lds128_t(np + jj * (ONECOL ? 16 : 64), 0);
        }
        const uint32_t xb[8] = {x0.x, x0.y, x0.z, x0.w, x1.x, x1.y, x1.z, x1.w};
        float acc[4];
#pragma unroll
        for (int m = 0; m < 4; ++m) {
          const int sh = F16 ? (m >> 1) * 8 : 4 * m;
          const uint32_t mk = (F16 && (m & 1)) ? lmask2 : lmask;
          uint32_t a0, a1, a2, a3;
          {
            a0 = lift4b(w0.x, sh, mk, lmagic); a1 = lift4b(w1.x, sh, mk, lmagic);
            a2 = lift4b(w0.y, sh, mk, lmagic); a3 = lift4b(w1.y, sh, mk, lmagic);
          }
          if (m == 0) mma_16816<F16>(acc, a0, a1, a2, a3, xb[0], xb[1], __uint_as_float(nq.x), __uint_as_float(nq.y),
                                     __uint_as_float(nq.z), __uint_as_float(nq.w));
          else mma_16816<F16>(acc, a0, a1, a2, a3, xb[2 * m], xb[2 * m + 1], acc[0], acc[1], acc[2], acc[3]);
        }
        const float s0 = jj & 1 ? bf16hi(sAw[jj >> 1]) : bf16lo(sAw[jj >> 1]);
        const float s1 = jj & 1 ? bf16hi(sBw[jj >> 1]) : bf16lo(sBw[jj >> 1]);
        ys[0] = fmaf(s0, acc[0], ys[0]);
        ys[2] = fmaf(s1, acc[2], ys[2]);
        if (!ONECOL) {
          ys[1] = fmaf(s0, acc[1], ys[1]);
          ys[3] = fmaf(s1, acc[3], ys[3]);
        }
      }
      }
      {  // bias term: biases (16 rows x 8 groups, as they lie in the stage) x the three pieces of sum_x
        const uint32_t ba0 = lds32_t(sp + 256 + g * 16 + t * 4, 0), ba1 = lds32_t(sp + 384 + g * 16 + t * 4, 0);
        const uint32_t pp = pcs_lane + (uint32_t)kc * 16 + tok;
#pragma unroll
        for (int p = 0; p < 3; ++p) {
          const uint32_t b0 = lds32_t(pp + (uint32_t)p * COLS * gp * 2, 0);
          mma_bf16_1688(yb, ba0, ba1, b0, yb);
        }
      }
    }
    __syncwarp();
    if (p_u < ue) {  // warp-uniform; one elected lane issues, every lane advances the cursor
      if (elect_one()) issue(p_tile, p_kc, st);
      ++p_u;
      if (++p_kc == upt) { p_kc = 0; ++p_tile; }
    }
    const int kc_done = kc, tile_done = tile;
    if (++kc == upt) { kc = 0; ++tile; }
    if (!tile_ends) continue;
    // ---- this warp's share of the tile is complete ---------------------------------------------------
    float v[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) v[q] = F16 ? fmaf(ys[q], rs[q & 1], yb[q]) : ys[q] + yb[q];
    if (run_start_kc == 0 && kc_done == upt - 1) {
      // (a whole tile inside one warp's run only happens for short-K, many-row matrices: the residual's L2 latency
      // is paid here; the Llama o/down projections always finish through the cut-tile path below, which hides it)
      if (a.epi == EPI_RESADD) load_residual(tile_done);
      epilogue(tile_done, v);
    } else if (e0_tile < 0) {
      // first partial share of this warp: parked in registers until the ring is idle
      e0_tile = tile_done;
#pragma unroll
      for (int q = 0; q < 4; ++q) e0v[q] = v[q];
    } else {
      // second partial share (the run's last tile): straight to this warp's stage 1 (its loads have all landed)
      __syncwarp();
      asm volatile("st.shared.v4.f32 [%0], {%1,%2,%3,%4};" ::"r"(wst + kQ2WBytes + lane * 16), "f"(v[0]), "f"(v[1]),
                   "f"(v[2]), "f"(v[3]) : "memory");
      if (lane == 0) edge_tile[warp * 2 + 1] = tile_done;
    }
  }
  Q2_STAMP(5)
  // park the first edge in this warp's stage 0 (every stage of this warp has been consumed by now)
  if (e0_tile >= 0) {
    asm volatile("st.shared.v4.f32 [%0], {%1,%2,%3,%4};" ::"r"(wst + lane * 16), "f"(e0v[0]), "f"(e0v[1]), "f"(e0v[2]),
                 "f"(e0v[3]) : "memory");
    if (lane == 0) edge_tile[warp * 2] = e0_tile;
  }
  // the warp whose run holds a cut tile's LAST unit adds the tile's shares in warp (= k) order and finishes it
  int fin_tile = -1;
  if (e0_tile >= 0 && (e0_tile + 1) * upt <= ue) fin_tile = e0_tile;   // edge 0 ends inside this run
  if (fin_tile >= 0 && a.epi == EPI_RESADD) load_residual(fin_tile);
  __syncthreads();
  Q2_STAMP_ORDERED(6)
  if (fin_tile >= 0) {
    // slot (2 w + e) of the 32 parked shares lives in stage buffer (2 w + e); lane i looks at slot i, the ballot gives the
    // slots of this tile (all of them at or before this warp's own), taken in slot (= k) order, four loads in flight
    static_assert(kQ2Stages == 2 && kQ2Warps == 16, "one parked share per stage buffer, 32 slots");
    unsigned m = __ballot_sync(0xffffffffu, edge_tile[lane] == fin_tile);
    float v[4] = {0.f, 0.f, 0.f, 0.f};
    while (m) {
      int sl[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) { sl[i] = m ? __ffs(m) - 1 : -1; m &= m - 1; }
      uint4 p[4];
#pragma unroll
      for (int i = 0; i < 4; ++i)
        p[i] = sl[i] >= 0 ? lds128(smem_s + (uint32_t)sl[i] * kQ2WBytes + lane * 16) : make_uint4(0, 0, 0, 0);
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        if (sl[i] < 0) break;   // (a parked -0.0f must not meet +0.0f padding: keep the sum exactly as the shares give it)
        v[0] += __uint_as_float(p[i].x); v[1] += __uint_as_float(p[i].y);
        v[2] += __uint_as_float(p[i].z); v[3] += __uint_as_float(p[i].w);
      }
    }
    epilogue(fin_tile, v);
  }
  Q2_STAMP_ORDERED(7)
  if (ONECOL && a.tp_mode == 1) {
    // producer: when this rank's LAST CTA is done, every peer may read slot `tp_rank` of exchange tp_e
    // (one fence by one thread behind the CTA barrier -- cumulative over the CTA's pushes -- not one per thread)
    __syncthreads();
    if (tid == 0) __threadfence_system();
    if (tid == 0 && atomicAdd(&a.tp_epoch[2], 1u) == gridDim.x - 1) {
      a.tp_epoch[2] = 0;   // (the release of the flag stores below is the only other fence: every CTA fenced its own pushes)
      const size_t foff = tp_flags_off(a.tp_world, a.tp_max_elems);
      for (int p = 0; p < a.tp_world; ++p)
        st_release_sys(reinterpret_cast<unsigned int*>(a.tp_peer[p] + foff) + ((size_t)tp_set * a.tp_world + a.tp_rank) * kTpMaxCtas, tp_e);
    }
  } else if (ONECOL && a.tp_mode == 2) {
    // consumer: the last CTA out advances the exchange number (every CTA read it in the x staging)
    __syncthreads();
    if (tid == 0) {
      __threadfence();
      if (atomicAdd(&a.tp_epoch[1], 1u) == gridDim.x - 1) {
        a.tp_epoch[1] = 0;
        __threadfence();
        a.tp_epoch[0] += 1;
      }
    }
  }
}

// ------------------------------------------------------------------------------------------- host side
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  });
  return fn;
}

struct MapKey {
  const void* ptr; int N, K, rows, kind, T, S; long long tstride, sstride;
  bool operator==(const MapKey& o) const {
    return ptr == o.ptr && N == o.N && K == o.K && rows == o.rows && kind == o.kind && T == o.T && S == o.S &&
           tstride == o.tstride && sstride == o.sstride;
  }
};
struct MapKeyHash {
  size_t operator()(const MapKey& k) const {
    size_t h = reinterpret_cast<size_t>(k.ptr);
    h ^= (size_t)k.N * 0x9E3779B97F4A7C15ull + (size_t)k.K * 0xC2B2AE3D27D4EB4Full + (size_t)k.rows * 131 + k.kind +
         (size_t)k.T * 7 + (size_t)k.S * 13 + (size_t)k.tstride * 31 + (size_t)k.sstride * 17;
    return h;
  }
};
// kind 0: packed 4-bit weights [N, K/2 bytes] as (128 B, row, 256-k chunk, matrix t < T at byte stride tstride);
// kind 1: bf16 scales / biases [N, K/64] as (group, row, matrix t < T, array s < S at byte stride sstride)
static bool get_map(CUtensorMap* out, const void* ptr, int N, int K, int rows, int kind, int T, long long tstride, int S,
                    long long sstride) {
  static std::unordered_map<MapKey, CUtensorMap, MapKeyHash> cache;
  static std::mutex mu;
  const MapKey key{ptr, N, K, rows, kind, T, S, tstride, sstride};
  std::lock_guard<std::mutex> lk(mu);
  auto it = cache.find(key);
  if (it != cache.end()) { *out = it->second; return true; }
  EncodeTiledFn enc = encode_fn();
  if (!enc) return false;
  CUtensorMap m;
  CUresult r;
  const cuuint32_t es[4] = {1, 1, 1, 1};
  if (kind == 0) {
    const cuuint64_t dims[4] = {128, (cuuint64_t)N, (cuuint64_t)(K / 256), (cuuint64_t)T};
    const cuuint64_t strides[3] = {(cuuint64_t)(K / 2), 128, (cuuint64_t)(T > 1 ? tstride : 16)};
    const cuuint32_t box[4] = {128, (cuuint32_t)rows, 2, (cuuint32_t)T};
    r = enc(&m, CU_TENSOR_MAP_DATA_TYPE_UINT8, 4, const_cast<void*>(ptr), dims, strides, box, es,
            CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  } else {
    const cuuint64_t dims[4] = {(cuuint64_t)(K / 64), (cuuint64_t)N, (cuuint64_t)T, (cuuint64_t)S};
    const cuuint64_t strides[3] = {(cuuint64_t)(K / 64) * 2, (cuuint64_t)(T > 1 ? tstride : 16),
                                   (cuuint64_t)(S > 1 ? sstride : 16)};
    const cuuint32_t box[4] = {8, (cuuint32_t)rows, (cuuint32_t)T, (cuuint32_t)S};
    r = enc(&m, CU_TENSOR_MAP_DATA_TYPE_UINT16, 4, const_cast<void*>(ptr), dims, strides, box, es,
            CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  }
  if (r != CUDA_SUCCESS) return false;
  if (cache.size() > 65536) cache.clear();
  cache.emplace(key, m);
  *out = m;
  return true;
}

bool qmv2_supported(int M, int K, int bits, int group_size, const pie_qmv_seg* segs, int nseg, int dual) {
  if (bits != 4 || group_size != 64 || M < 1 || M > 8 || K % kQ2UnitK != 0) return false;
  if (qmv2_smem_bytes(M, K) > (size_t)kQmvMaxSmem) return false;
  for (int i = 0; i < nseg; ++i) {
    if ((reinterpret_cast<uintptr_t>(segs[i].wq) & 15) || (reinterpret_cast<uintptr_t>(segs[i].scales) & 15) ||
        (reinterpret_cast<uintptr_t>(segs[i].biases) & 15))
      return false;
  }
  (void)dual;
  return encode_fn() != nullptr;
}

// a usable stride between two arrays: positive, a multiple of 16 bytes, below 2^40
static bool stride_ok(const void* lo, const void* hi, long long* out) {
  const long long d = (const char*)hi - (const char*)lo;
  *out = d;
  return d > 0 && d % 16 == 0 && d < (1ll << 40);
}
static const bool kNoStack = getenv("PIE_QMV2_NOSTACK") && atoi(getenv("PIE_QMV2_NOSTACK"));

// Fill the matrix part of a descriptor (maps, tiling).  Returns false when a tensor map cannot be encoded.
static bool fill_matrix(Qmv2Args& a, const pie_qmv_seg* segs, int nseg, int K, int dual) {
  int tile = 0, off = 0;
  const int rows = dual ? 8 : 16;
  // biases a fixed stride behind the scales (the same stride for both matrices of a dual pair)?
  long long sbs = 0, ws = 0, ss = 0, t = 0;
  bool sb = !kNoStack, gu = dual && !kNoStack;
  for (int i = 0; i < nseg && sb; ++i) {
    sb = stride_ok(segs[i].scales, segs[i].biases, &t) && (i == 0 || !dual || t == sbs);
    if (i == 0) sbs = t;
  }
  if (dual) {
    gu = gu && sb && nseg == 2 && segs[0].N == segs[1].N && stride_ok(segs[0].wq, segs[1].wq, &ws) &&
         stride_ok(segs[0].scales, segs[1].scales, &ss);
    if (!gu) sb = false;   // (the stage layout of a pair needs both or neither)
  }
  a.sb_stacked = sb; a.gu_stacked = gu;
  for (int i = 0; i < nseg; ++i) {
    if (!(gu && i == 1)) {
      long long sbi = 0;
      if (sb) stride_ok(segs[i].scales, segs[i].biases, &sbi);
      if (!get_map(&a.wmap[i], segs[i].wq, segs[i].N, K, rows, 0, gu ? 2 : 1, ws, 1, 0)) return false;
      if (!get_map(&a.smap[i], segs[i].scales, segs[i].N, K, rows, 1, gu ? 2 : 1, ss, sb ? 2 : 1, sbi)) return false;
      if (!sb && !get_map(&a.bmap[i], segs[i].biases, segs[i].N, K, rows, 1, 1, 0, 1, 0)) return false;
    }
    a.N[i] = segs[i].N; a.tile0[i] = tile; a.out_off[i] = off;
    tile += (segs[i].N + 15) / 16;
    off += segs[i].N;
  }
  a.ntiles = dual ? (segs[0].N + 7) / 8 : tile;
  return true;
}

int build_qmv2_args(Qmv2Args& a, void* out, int out_ld, const void* x, int x_ld, int M, int K, const pie_qmv_seg* segs,
                    int nseg, const void* norm_weight, float eps, int epilogue, const void* residual,
                    const pie_tp_fused* tp) {
  memset(&a, 0, sizeof(a));
  const int dual = epilogue == EPI_SILUMUL;
  if (!fill_matrix(a, segs, nseg, K, dual)) return PIE_EUNSUPPORTED;
  a.nseg = nseg; a.dual = dual; a.M = M; a.K = K; a.upt = q2_kpad(K) / kQ2UnitK;
  a.x = (const uint16_t*)x; a.x_ld = x_ld;
  a.norm_w = (const uint16_t*)norm_weight; a.eps = eps;
  a.out = (uint16_t*)out; a.out_ld = out_ld; a.residual = (const uint16_t*)residual; a.epi = epilogue;
  // fp16 staging of x (see q2_stage_x) needs every block of a thread in registers: M * kpad / 16 <= 4 * 512 blocks
  static const int f16_env = getenv("PIE_QMV2_F16") ? atoi(getenv("PIE_QMV2_F16")) : -1;
  static const int dbg_env = getenv("PIE_QMV2_DBG") ? atoi(getenv("PIE_QMV2_DBG")) : 0;  // experiments, see the kernel
  a.dbg = dbg_env;
  a.upt_inv = a.upt > 1 ? (uint32_t)(((1ull << 32) + a.upt - 1) / a.upt) : 0u;   // (upt == 1 is special-cased)
  // fp16 staging (see q2_stage_x) needs every block of a thread in registers.  Measured on B200, one row, us bf16 / fp16
  // staging: q/k/v 8.0 / 7.7, gate+up 17.4 / 16.5, head 52.2 / 49.7 (inputs that are RMS-normalised on the way in: the scale
  // comes out of the same reduction pass), o 5.3 / 5.7, down 11.5 / 11.6 (plain inputs: the row maximum is an extra pass).
  // Default: fp16 exactly when the RMS norm is fused.  PIE_QMV2_F16 = 0 / 1 forces.
  const bool f16_fits = (long long)M * q2_kpad(K) / 16 <= 4 * kQ2Warps * 32;
  a.f16 = f16_fits && (f16_env < 0 ? norm_weight != nullptr : f16_env != 0);
  if (tp && tp->mode) {
    const pie_tp_peers& pr = tp->peers;
    if (M != 1 || pr.world < 1 || pr.world > kTpMaxWorld || !pr.epoch) return PIE_EUNSUPPORTED;
    if (tp->mode == 1)   // producer tiles are pushed as whole 16-row, 32-byte pieces
      for (int i = 0; i < nseg; ++i)
        if (segs[i].N % 16 != 0) return PIE_EUNSUPPORTED;
    if (tp->mode == 2 && q2_kpad(K) / 16 > 2 * kQ2Warps * 32) return PIE_EUNSUPPORTED;  // the one-barrier staging form
    if (tp->mode == 2) a.f16 = 1;
    a.tp_mode = tp->mode; a.tp_rank = pr.rank; a.tp_world = pr.world; a.tp_max_elems = pr.max_elems;
    for (int i = 0; i < pr.world; ++i) a.tp_peer[i] = (uint8_t*)pr.base[i];
    a.tp_epoch = (unsigned int*)pr.epoch;
    a.tp_h_out = (uint16_t*)tp->h_out;
    if (a.f16) a.lift_magic = 0x64006400u;
  }
  a.lift_mask = 0x000f000fu; a.lift_mask2 = 0x00f000f0u; a.lift_magic = a.f16 ? 0x64006400u : 0x43004300u;
  return PIE_OK;
}

// profiling aid: every launch takes the next slot of 256 CTAs x 16 stamps from the buffer set by pie_debug_stamps
static unsigned long long* g_stamp_buf = nullptr;
static size_t g_stamp_slots = 0, g_stamp_next = 0;
void debug_stamps(void* buf, size_t bytes) {
  g_stamp_buf = (unsigned long long*)buf; g_stamp_slots = buf ? bytes / kStampSlotBytes : 0; g_stamp_next = 0;
}
unsigned long long* next_stamp_slot() {
  if (!g_stamp_buf || g_stamp_next >= g_stamp_slots) return nullptr;
  return g_stamp_buf + (g_stamp_next++) * (kStampSlotBytes / 8);
}

template <bool ONECOL, bool F16, bool DUAL>
static cudaError_t launch_qmv2_t(const Qmv2Args& a, bool pdl, cudaStream_t st) {
  cudaError_t e = cudaFuncSetAttribute(qmv2_kernel<ONECOL, F16, DUAL>, cudaFuncAttributeMaxDynamicSharedMemorySize, kQmvMaxSmem);
  if (e != cudaSuccess) return e;
  const int grid = a.tiles_q > 0 ? num_sms() : a.ntiles;   // (launch_qmv2 set the partition for this grid)
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)grid, 1, 1);
  cfg.blockDim = dim3(kQ2Warps * 32, 1, 1);
  cfg.dynamicSmemBytes = qmv2_smem_bytes(a.M, a.K);
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl ? 1 : 0;
  count_launch();
  return cudaLaunchKernelEx(&cfg, qmv2_kernel<ONECOL, F16, DUAL>, a);
}
template <bool ONECOL, bool F16>
static cudaError_t launch_qmv2_d(const Qmv2Args& a, bool pdl, cudaStream_t st) {
  return a.dual ? launch_qmv2_t<ONECOL, F16, true>(a, pdl, st) : launch_qmv2_t<ONECOL, F16, false>(a, pdl, st);
}

cudaError_t launch_qmv2(const Qmv2Args& a0, bool pdl, cudaStream_t st) {
  Qmv2Args a = a0;
  a.stamps = next_stamp_slot();
  {
    int grid = num_sms();
    if (grid > a.ntiles) grid = a.ntiles;
    a.tiles_q = a.ntiles / grid; a.tiles_r = a.ntiles % grid;
    if (grid == a.ntiles) { a.tiles_q = 0; a.tiles_r = a.ntiles; }   // one tile per CTA, grid = ntiles (see launch_qmv2_t)
  }
  if (a.f16) return a.M == 1 ? launch_qmv2_d<true, true>(a, pdl, st) : launch_qmv2_d<false, true>(a, pdl, st);
  return a.M == 1 ? launch_qmv2_d<true, false>(a, pdl, st) : launch_qmv2_d<false, false>(a, pdl, st);
}

}  // namespace pie
