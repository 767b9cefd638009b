// Batched-decode quantized GEMM (9 .. 32 activation rows per pass) on tcgen05 / TMEM, sm_100a.
//
// Replaces mx.quantized_matmul(transpose=True) where MLX itself leaves the vector kernel for qmm_t (M >= vector_limit,
// MLX/mlx/backend/metal/quantized.cpp; kernels/quantized.h:954-1078): w_hat = bf16(scale * q + bias) is formed once in
// shared memory and the contraction accumulates in fp32.  The work is HBM-bound like the GEMV (every weight byte is
// read once per step), but at 32 rows mma.sync leaves the 4x larger accumulator / affine work on the FMA pipe; here the
// operands are swapped -- the WEIGHT tile is the 128-row A operand, the batch is the N = 32 side -- so the whole
// contraction is 8 tcgen05.mma per 128-element K stage and the SM's issue slots go to dequantisation only.
//
// Persistent CTA (one per SM), 21 warps:
//   * warps 0-15, thread t <-> weight row t / 4, quarter t % 4 of the stage's 128 k: a private cp.async ring (R stages
//     deep) of its packed words + its share of the scales / biases and of the stage's activation tile (copied straight
//     into the 128B-swizzled K-major layout the MMA reads); dequantise its 32 values -> A stage (2 deep).  4-bit: two
//     values per instruction (bf16x2 lift, HFMA2 for (128 + q) - 128 and q * s + b); 8-bit: fp32 lift + FFMA + packed cvt.
//   * warp 16, one lane: tcgen05.mma M128 N32 K16 x 8 per stage into one of two TMEM accumulators; commit frees the A
//     stage and the ring slot's activation tile; after the last stage of a unit, commit hands the accumulator over.
//   * warps 17-20: epilogue.  TMEM lane quadrant -> registers; cross-CTA split-K partials (small N) are summed in slice
//     order by the last CTA to arrive; then STORE / RESADD / SILUMUL exactly as the GEMV epilogues (pie_qmv_loop.inl).
// A unit = (128-row tile of one matrix, K slice); SwiGLU units run gate rows then up rows into two accumulators.
#include "pie_internal.h"
#include "pie_ptx.cuh"
#include "pie_umma.cuh"

namespace pie {

constexpr int kQdRows = 128, kQdMb = 32, kQdBK = 128;
constexpr int kQdAStage = kQdRows * kQdBK * 2;         // 32 KB: two swizzle atoms of [128 rows][64 k]
constexpr int kQdXBytes = kQdMb * kQdBK * 2;           // 8 KB: two atoms of [32 rows][64 k]
constexpr int kQdSbBytes = kQdRows * 8;                // per row: scales of the 2 groups, biases of the 2 groups
constexpr int kQdProducers = 512, kQdLoaderWarp0 = kQdProducers / 32 + 5, kQdThreads = kQdProducers + 32 + 128 + 128;
constexpr int kQdNA = 3;                               // A / B stages: hides fence + MMA + commit latency behind two dequant periods
constexpr int kQdTmemCols = 128;                       // 2 accumulator buffers x (gate | up) x 32 columns
constexpr uint32_t kQdIdesc = umma_idesc_bf16_m128(kQdMb);

template <int BITS> struct QdGeom {
  // A ring slot holds 128 BYTES of every weight row (one L2 line per row and request: 64-byte requests ran at half the
  // DRAM request rate): two stages of 4-bit weights, one stage of 8-bit weights, + their scales / biases / activations.
  static constexpr int SPS = BITS == 4 ? 2 : 1;                   // stages per slot
  static constexpr int RAW = kQdRows * 128;                       // 16 KB
  static constexpr int XB = SPS * kQdXBytes;                      // 16 KB / 8 KB, row-major [32][SPS * 128 k]
  static constexpr int SB = SPS * kQdSbBytes;                     // per row: SPS*2 scales, then SPS*2 biases
  static constexpr int SLOT = XB + RAW + SB;                      // 34 KB / 25 KB
  static constexpr int R = BITS == 4 ? 3 : 4;                     // ring depth (slots)
  static constexpr int SMEM = R * SLOT + kQdNA * (kQdAStage + kQdXBytes) + 1024;  // + A / B stages + alignment slack
};

struct QmdArgs {
  QmvSeg seg[kMaxSeg];   // tile0 counts 128-row tiles here
  int nseg, dual;
  int M, K;
  int ntiles;            // output tiles (dual: tiles of seg[0])
  int ksplit, kstages;   // K slices per tile; 128-element stages per slice
  const uint16_t* x;     // bf16 [M, x_ld], already normalised
  int x_ld;
  uint16_t* out;
  int out_ld;
  const uint16_t* residual;
  int epi;
  float* ws_part;        // [ntiles][ksplit][32][128] fp32
  unsigned int* ws_cnt;  // [ntiles][4]
  int dbg;               // PIE_QMD_DBG timing experiments (results are then wrong): 1 no dequant, 2 no a_empty wait, 4 no fence
};

__device__ __forceinline__ void qd_tile_seg(const QmdArgs& a, int tile, int pass, int& si, int& nb) {
  if (a.dual) { si = pass; nb = tile * kQdRows; return; }
  si = 0;
  for (int i = 1; i < a.nseg; ++i) if (tile >= a.seg[i].tile0) si = i;
  nb = (tile - a.seg[si].tile0) * kQdRows;
}

template <int BITS>
__global__ void __launch_bounds__(kQdThreads, 1) qmd_kernel(const __grid_constant__ QmdArgs a) {
  using G = QdGeom<BITS>;
  constexpr int R = G::R;
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t a_full[kQdNA], a_empty[kQdNA], acc_full[2], acc_empty[2], raw_full[8], raw_empty[8];
  __shared__ uint32_t tmem_slot;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t ring = base, smA = base + R * G::SLOT, smB = smA + kQdNA * kQdAStage;
  const int Rn = (a.dbg >> 8) ? min(a.dbg >> 8, R) : R;  // ring depth actually used (timing experiment)
  const int nunits = a.ntiles * a.ksplit;
  const int passes = a.dual ? 2 : 1;
  const int spu = passes * a.kstages;  // stages per unit
  const int my_units = blockIdx.x < nunits ? (nunits - 1 - blockIdx.x) / gridDim.x + 1 : 0;
  const int total = my_units * spu;

  if (tid == 0) {
    for (int i = 0; i < kQdNA; ++i) { mbar_init(&a_full[i], kQdProducers / 32); mbar_init(&a_empty[i], 1); }
    for (int i = 0; i < 2; ++i) { mbar_init(&acc_full[i], 1); mbar_init(&acc_empty[i], 4); }
    for (int i = 0; i < R; ++i) { mbar_init(&raw_full[i], 128); mbar_init(&raw_empty[i], kQdProducers / 32); }
    fence_mbar_init();
  }
  if (warp == kQdProducers / 32) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_slot)),
                 "n"(kQdTmemCols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = tmem_slot;

  if (warp < kQdProducers / 32) {
    // ================= dequantisers: thread <-> (weight row r, quarter qtr of the stage's 128 k) ===========
    // No cp.async is ever outstanding in these threads: fence.proxy.async waits for the thread's in-flight copies, and
    // with the ring in the same threads it serialised every stage on an HBM round trip (ncu: 28 % of the stalls).
    const int r = tid >> 2, qtr = tid & 3, h = qtr >> 1, cbase = (qtr & 1) * 4;
    const uint32_t swz = (uint32_t)(r & 7);
    constexpr int WB = 32 * BITS / 8;  // packed bytes of this thread's 32 values: 16 / 32
    uint32_t lmask, lmagic;            // in registers, so that and-or is ONE LOP3 (as immediates ptxas needs two)
    asm volatile("mov.b32 %0, 0x000f000f;" : "=r"(lmask));
    asm volatile("mov.b32 %0, 0x43004300;" : "=r"(lmagic));
    const int xm = tid >> 4, xc = tid & 15;  // this thread's 16-byte chunk of the stage's activation tile
    constexpr int SPS = G::SPS;
    const uint32_t xsrc = xm * (SPS * 256) + xc * 16;
    const uint32_t xdst = (xc >> 3) * (kQdMb * 128) + xm * 128 + (((xc & 7) ^ (xm & 7)) << 4);
    uint32_t slot = ring;
    int sl = 0, ring_use = 0, as = 0, use = 0, half = 0;
    for (int f = 0; f < total; ++f) {
      if (half == 0) mbar_wait(&raw_full[sl], (uint32_t)ring_use & 1);
      if (use > 0 && !(a.dbg & 2)) mbar_wait(&a_empty[as], (uint32_t)(use - 1) & 1);  // MMA of stage f - NA has read this A / B stage
      const uint4 xv = lds128(slot + xsrc + half * 256);
      const uint32_t wsm = slot + G::XB + (SPS == 2 ? r * 128 + half * 64 + qtr * 16 : tid * WB);
      uint2 sb;
      if (SPS == 2) {
        const uint4 sb4 = lds128(slot + G::XB + G::RAW + r * 16);  // s0 s1 | s2 s3 | b0 b1 | b2 b3
        sb = half ? make_uint2(sb4.y, sb4.w) : make_uint2(sb4.x, sb4.z);
      } else {
        sb = lds64(slot + G::XB + G::RAW + r * 8);
      }
      const uint32_t dA = smA + as * kQdAStage + h * (kQdRows * 128) + r * 128;
      asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(smB + as * kQdXBytes + xdst), "r"(xv.x), "r"(xv.y),
                   "r"(xv.z), "r"(xv.w) : "memory");
      if (a.dbg & 1) {
      } else if (BITS == 4) {
        // two values per instruction: (128 + q_j, 128 + q_j+4) lifted with one shift + one LOP3, minus 128 (exact),
        // then bf16(q * s + b) with one rounding on the FMA pipe; byte permutes restore the k order
        const uint32_t s2 = __byte_perm(sb.x, sb.x, h ? 0x3232 : 0x1010), b2 = __byte_perm(sb.y, sb.y, h ? 0x3232 : 0x1010);
        const uint4 wv = lds128(wsm);
        const uint32_t w[4] = {wv.x, wv.y, wv.z, wv.w};
#pragma unroll
        for (int i0 = 0; i0 < 4; ++i0) {
          const int i = (i0 + 2 * h) & 3;  // the two groups of a row start two chunks apart: conflict-free stores
          const uint32_t wi = h ? (i0 < 2 ? w[i0 + 2] : w[i0 - 2]) : w[i0];
          uint32_t v[4];
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            uint32_t p;
            asm("lop3.b32 %0, %1, %2, %3, 0xEA;" : "=r"(p) : "r"(wi >> (4 * j)), "r"(lmask), "r"(lmagic));
            v[j] = fma_bf16x2(fma_bf16x2(p, 0x3f803f80u, 0xc300c300u), s2, b2);
          }
          const uint32_t o0 = __byte_perm(v[0], v[1], 0x5410), o1 = __byte_perm(v[2], v[3], 0x5410);
          const uint32_t o2 = __byte_perm(v[0], v[1], 0x7632), o3 = __byte_perm(v[2], v[3], 0x7632);
          asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(dA + (((uint32_t)(cbase + i) ^ swz) << 4)), "r"(o0),
                       "r"(o1), "r"(o2), "r"(o3) : "memory");
        }
      } else {
        const float sc = h ? bf16hi(sb.x) : bf16lo(sb.x);
        const float bb = fmaf(-256.f, sc, h ? bf16hi(sb.y) : bf16lo(sb.y));
#pragma unroll
        for (int c4 = 0; c4 < 2; ++c4) {
          const uint4 wv = lds128(wsm + c4 * 16);
          const uint32_t w[4] = {wv.x, wv.y, wv.z, wv.w};
#pragma unroll
          for (int i = 0; i < 2; ++i) {  // two words -> 8 values -> one 16-byte chunk
            uint32_t o[4];
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              const uint32_t ww = w[2 * i + (e >> 1)];
              o[e] = pack_bf16(fmaf(lift_f32<8>(ww, 16 * (e & 1)), sc, bb), fmaf(lift_f32<8>(ww, 16 * (e & 1) + 8), sc, bb));
            }
            asm volatile("st.shared.v4.b32 [%0], {%1,%2,%3,%4};" ::"r"(dA + (((uint32_t)(cbase + c4 * 2 + i) ^ swz) << 4)),
                         "r"(o[0]), "r"(o[1]), "r"(o[2]), "r"(o[3]) : "memory");
          }
        }
      }
      if (!(a.dbg & 4)) fence_proxy_async();  // st.shared writes -> visible to the tensor core's async proxy
      // ONE arrive per warp: 512 same-address mbarrier arrives per stage serialise in shared memory (measured: 0.5 us
      // per stage of pure barrier traffic).  Every lane has fenced its own writes; __syncwarp orders them before lane 0.
      __syncwarp();
      const bool slot_done = half == SPS - 1 || f == total - 1;
      if (lane == 0) {
        mbar_arrive(&a_full[as]);
        if (slot_done) mbar_arrive(&raw_empty[sl]);  // the warp is done with the slot's words, scales and activations
      }
      if (slot_done) {
        half = 0;
        if (++sl == Rn) { sl = 0; slot = ring; ++ring_use; } else { slot += G::SLOT; }
      } else {
        ++half;
      }
      if (++as == kQdNA) { as = 0; ++use; }
    }
  } else if (warp >= kQdLoaderWarp0) {
    // ================= loaders: 4 warps stream the ring with cp.async; completion arrives on raw_full[slot] =========
    // A pass has an even number of stages when SPS == 2 (host: kstages % SPS == 0), so a slot never spans two passes.
    constexpr int SPS = G::SPS;
    const int l = tid - kQdLoaderWarp0 * 32;           // 0 .. 127
    const uint8_t* wp[8];                               // 1024 16-byte chunks per slot: 8 per thread, 8 chunks per row
    const uint16_t* sbp[2];
    const uint16_t* xp[4 * SPS];
#pragma unroll
    for (int j = 0; j < 4 * SPS; ++j) {
      const int id = l + 128 * j, m = id / (16 * SPS), c = id % (16 * SPS);
      xp[j] = a.x + (size_t)min(m, a.M - 1) * a.x_ld + c * 8;
    }
    uint32_t slot = ring;
    int sl = 0, ring_use = 0;
    for (int ui = 0; ui < my_units; ++ui) {
      const int u = blockIdx.x + ui * gridDim.x;
      const int tile = u / a.ksplit, slice = u % a.ksplit;
      for (int pass = 0; pass < passes; ++pass) {
        int si, nb;
        qd_tile_seg(a, tile, pass, si, nb);
        int k0 = slice * a.kstages * kQdBK;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const int id = l + 128 * j, row = id >> 3, c = id & 7;
          const int n = min(nb + row, a.seg[si].N - 1);
          wp[j] = reinterpret_cast<const uint8_t*>(a.seg[si].wq) + ((size_t)n * a.K + k0) * BITS / 8 + c * 16;
        }
#pragma unroll
        for (int j = 0; j < 2; ++j) {
          const int id = l + 128 * j, row = id >> 1;
          const int n = min(nb + row, a.seg[si].N - 1);
          sbp[j] = ((id & 1) ? a.seg[si].biases : a.seg[si].scales) + (size_t)n * (a.K / 64) + k0 / 64;
        }
        for (int ks = 0; ks < a.kstages; ks += SPS) {
          if (ring_use > 0) mbar_wait(&raw_empty[sl], (uint32_t)(ring_use - 1) & 1);
#pragma unroll
          for (int j = 0; j < 8; ++j) { cp_async16(slot + G::XB + (l + 128 * j) * 16, wp[j]); wp[j] += 128; }
#pragma unroll
          for (int j = 0; j < 2; ++j) {  // row id >> 1: scales at + 0, biases at + SPS * 4
            const int id = l + 128 * j;
            const uint32_t dst = slot + G::XB + G::RAW + (id >> 1) * (SPS * 8) + (id & 1) * (SPS * 4);
            if (SPS == 2) cp_async8(dst, sbp[j]); else cp_async4(dst, sbp[j]);
            sbp[j] += 2 * SPS;
          }
#pragma unroll
          for (int j = 0; j < 4 * SPS; ++j) cp_async16(slot + (l + 128 * j) * 16, xp[j] + k0);
          asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(smem_u32(&raw_full[sl])) : "memory");
          k0 += SPS * kQdBK;
          if (++sl == Rn) { sl = 0; slot = ring; ++ring_use; } else { slot += G::SLOT; }
        }
      }
    }
    asm volatile("cp.async.wait_all;" ::: "memory");
  } else if (warp == kQdProducers / 32) {
    // ================= MMA issuer ====================================================================
    int as = 0, ause = 0;
    for (int f = 0; f < total; ++f) {
      const int ui = f / spu, i = f % spu, buf = ui & 1;
      const int pass = i / a.kstages, ks = i % a.kstages;
      if (i == 0 && ui >= 2) mbar_wait(&acc_empty[buf], (uint32_t)((ui >> 1) - 1) & 1);  // epilogue has read it
      mbar_wait(&a_full[as], (uint32_t)ause & 1);
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      if (lane == 0) {
        const uint32_t d = tmem + buf * 64 + pass * 32;
#pragma unroll
        for (int at = 0; at < 2; ++at) {
          const uint64_t da = umma_desc_sw128(smA + as * kQdAStage + at * (kQdRows * 128));
          const uint64_t db = umma_desc_sw128(smB + as * kQdXBytes + at * (kQdMb * 128));
#pragma unroll
          for (int k4 = 0; k4 < 4; ++k4)
            if (!(a.dbg & 32)) umma_bf16(d, da + 2 * k4, db + 2 * k4, kQdIdesc, (ks > 0 || at > 0 || k4 > 0) ? 1u : 0u);
        }
        umma_commit(&a_empty[as]);
        if (i == spu - 1) umma_commit(&acc_full[buf]);
      }
      __syncwarp();
      if (++as == kQdNA) { as = 0; ++ause; }
    }
  } else if (warp < kQdLoaderWarp0) {
    // ================= epilogue: TMEM lane quadrant warp % 4 ========================================
    const int quad = warp & 3;
    for (int ui = 0; ui < my_units; ++ui) {
      const int buf = ui & 1;
      const int u = blockIdx.x + ui * gridDim.x;
      const int tile = u / a.ksplit, slice = u % a.ksplit;
      mbar_wait(&acc_full[buf], (uint32_t)(ui >> 1) & 1);
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      uint32_t v[32], v2[32];
      tmem_ld32(tmem + (uint32_t)(buf * 64) + ((uint32_t)(quad * 32) << 16), v);
      if (a.dual) tmem_ld32(tmem + (uint32_t)(buf * 64 + 32) + ((uint32_t)(quad * 32) << 16), v2);
      asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
      __syncwarp();
      if (lane == 0) mbar_arrive(&acc_empty[buf]);
      const int row = quad * 32 + lane;
      if (a.ksplit > 1) {  // (never with dual) partial -> workspace; the last slice to arrive sums them in slice order
        float* part = a.ws_part + ((size_t)tile * a.ksplit) * (kQdMb * kQdRows) + row;
#pragma unroll
        for (int m = 0; m < kQdMb; ++m) part[((size_t)slice * kQdMb + m) * kQdRows] = __uint_as_float(v[m]);
        __threadfence();
        __syncwarp();
        unsigned last = 0;
        if (lane == 0) last = atomicAdd(&a.ws_cnt[tile * 4 + quad], 1u) == (unsigned)a.ksplit - 1;
        last = __shfl_sync(0xffffffffu, last, 0);
        if (!last) continue;
        __threadfence();
#pragma unroll
        for (int m = 0; m < kQdMb; ++m) v[m] = 0u;
        for (int s = 0; s < a.ksplit; ++s) {
#pragma unroll
          for (int m = 0; m < kQdMb; ++m)
            v[m] = __float_as_uint(__uint_as_float(v[m]) + ldcg_f32(part + ((size_t)s * kQdMb + m) * kQdRows));
        }
        if (lane == 0) a.ws_cnt[tile * 4 + quad] = 0;
      }
      int si, nb;
      qd_tile_seg(a, tile, 0, si, nb);
      const int n = nb + row;
      if (n < a.seg[si].N) {
        if (a.epi == EPI_SILUMUL) {
#pragma unroll
          for (int m = 0; m < kQdMb; ++m) {
            if (m < a.M) {
              const float gt = bfround(__uint_as_float(v[m])), up = bfround(__uint_as_float(v2[m]));
              const float sg = bfround(1.f / (1.f + expf(-gt)));  // nn.silu = x * sigmoid(x), each op rounded
              a.out[(size_t)m * a.out_ld + n] = f2bf(bfround(gt * sg) * up);
            }
          }
        } else {
          const size_t col = (size_t)a.seg[si].out_off + n;
#pragma unroll
          for (int m = 0; m < kQdMb; ++m) {
            if (m < a.M) {
              float res = bfround(__uint_as_float(v[m]));
              if (a.epi == EPI_RESADD) res = bf16lo((uint32_t)a.residual[(size_t)m * a.out_ld + col]) + res;
              a.out[(size_t)m * a.out_ld + col] = f2bf(res);
            }
          }
        }
      }
      __syncwarp();  // tcgen05.ld of the next unit needs the warp converged
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == kQdProducers / 32)
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(kQdTmemCols) : "memory");
}

// ---- host side -----------------------------------------------------------------------------------------------
// K slices per tile: among the divisors d <= 8 of the stage count that leave >= 4 stages per unit, the one that fills
// the SMs best over whole rounds (ties: fewer slices).
int qmd_pick_ksplit(int ntiles, int K, int dual, int bits) {
  if (dual) return 1;
  const int stages = K / kQdBK, sms = num_sms(), sps = bits == 4 ? 2 : 1;
  int best = 1;
  double best_eff = 0.0;
  for (int d = 1; d <= 8; ++d) {
    if (stages % d || (d > 1 && stages / d < 4) || (stages / d) % sps) continue;
    const long long units = (long long)ntiles * d;
    const double eff = (double)units / (double)(((units + sms - 1) / sms) * sms);
    if (eff > best_eff + 0.02) { best_eff = eff; best = d; }
  }
  return best;
}

size_t qmd_ws_bytes(int M, int ntiles, int K, int ksplit, bool norm) {
  size_t b = norm ? (((size_t)M * K * 2 + 255) & ~(size_t)255) : 0;
  if (ksplit > 1) b += (((size_t)ntiles * 4 * 4 + 255) & ~(size_t)255) + (size_t)ntiles * ksplit * kQdMb * kQdRows * 4;
  return b;
}

bool qmd_supported(int mb, int K, int x_ld, int group_size, int bits) {
  return mb > 8 && mb <= kQdMb && group_size == 64 && (bits == 4 || bits == 8) && K % (bits == 4 ? 256 : 128) == 0 &&
         x_ld % 8 == 0;
}

// Fill and launch.  x must be normalised already (the caller runs launch_rms_norm into the workspace when needed).
cudaError_t launch_qmd(const QmvArgs& g, const uint16_t* x, int x_ld, int bits, float* ws_part, unsigned int* ws_cnt,
                       int ksplit, cudaStream_t st) {
  QmdArgs a;
  memset(&a, 0, sizeof(a));
  int tile = 0;
  for (int i = 0; i < g.nseg; ++i) {
    a.seg[i] = g.seg[i];
    a.seg[i].tile0 = tile;
    tile += (g.seg[i].N + kQdRows - 1) / kQdRows;
  }
  a.nseg = g.nseg; a.dual = g.dual; a.M = g.M; a.K = g.K;
  a.ntiles = g.dual ? (g.seg[0].N + kQdRows - 1) / kQdRows : tile;
  a.ksplit = ksplit; a.kstages = g.K / kQdBK / ksplit;
  a.x = x; a.x_ld = x_ld; a.out = g.out; a.out_ld = g.out_ld; a.residual = g.residual; a.epi = g.epi;
  a.ws_part = ws_part; a.ws_cnt = ws_cnt;
  { const char* e = getenv("PIE_QMD_DBG"); a.dbg = e ? atoi(e) : 0; }
  const int smem = bits == 4 ? QdGeom<4>::SMEM : QdGeom<8>::SMEM;
  static unsigned long long configured[2] = {0, 0};   // per device
  const int idx = bits == 4 ? 0 : 1;
  if (first_use_on_device(configured[idx])) {
    cudaError_t e = bits == 4 ? cudaFuncSetAttribute(qmd_kernel<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem)
                              : cudaFuncSetAttribute(qmd_kernel<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
    if (e != cudaSuccess) return e;
  }
  const int units = a.ntiles * a.ksplit;
  const int grid = units < num_sms() ? units : num_sms();
  count_launch();
  if (bits == 4) qmd_kernel<4><<<grid, kQdThreads, smem, st>>>(a);
  else qmd_kernel<8><<<grid, kQdThreads, smem, st>>>(a);
  return cudaGetLastError();
}

}  // namespace pie
