// Prefill attention (L > 1, causal) on tensor cores, sm_100a.
//
// Replaces mx.fast.scaled_dot_product_attention with the additive causal mask of create_attention_mask
// (PIE/models/base.py:18-53, 92-113; Metal steel_attention, MLX/mlx/backend/metal/scaled_dot_product_attention.cpp:
// 16-131): out[b, l, h] = softmax(scale * q[b,h,l] . K[b, h / n_rep, :offset+l+1]) . V.  The mask value -1e9 makes
// masked probabilities exactly 0 in fp32, so masked keys are skipped (tiles beyond the diagonal are never loaded).
//
// Flash-attention forward: a CTA of 4 warps owns 64 query rows of one head (16 per warp, Q fragments in registers);
// K / V tiles of 64 keys stream through a double-buffered, XOR-swizzled shared-memory stage (cp.async, straight from
// the contiguous or paged cache); S = Q K^T and O += P V run on mma.sync.m16n8k16 (bf16 in, fp32 accumulate) with
// ldmatrix / ldmatrix.trans operand loads; the softmax is online in fp32 (exp2 on log2e-prescaled scores); P is
// rounded to bf16 for the second product.  GQA: the n_rep query heads of a kv head are separate CTAs; their K/V tiles
// come from L2.
#include "pie_internal.h"
#include "pie_ptx.cuh"

namespace pie {

constexpr int kFaBQ = 64, kFaBKV = 64, kFaThreads = 128, kFaMaxSplits = 32, kFaDecStages = 3;

__device__ __forceinline__ size_t fa_kv_addr(const KvLayout& kv, int b, int h, int t) {
  if (kv.page_table) {
    const int page = __ldg(kv.page_table + (size_t)b * kv.pt_stride + t / kv.page_size);
    return ((size_t)page * kv.page_size + t % kv.page_size) * kv.tok_stride + (size_t)h * kv.head_stride;
  }
  return (size_t)b * kv.batch_stride + (size_t)h * kv.head_stride + (size_t)t * kv.tok_stride;
}
__device__ __forceinline__ void ldsm_x4(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr) : "memory");
}
__device__ __forceinline__ void ldsm_x4_trans(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr) : "memory");
}

struct FaArgs {
  uint16_t* out;
  const uint16_t* q;
  const uint16_t* kc;
  const uint16_t* vc;
  KvLayout kv;
  int Hq, Hkv, L, offset;
  float scale_log2e;
};

// shared-memory tile [64 keys][D] bf16; the 16-byte chunk c of row r sits at chunk (c ^ (r & 7)): ldmatrix reads of
// 8 consecutive rows at one logical chunk hit 8 different bank groups
template <int D>
__device__ __forceinline__ uint32_t fa_tile_off(int r, int c) {
  return (uint32_t)(r * (D * 2) + ((c ^ (r & 7)) << 4));
}

template <int D>
__global__ void __launch_bounds__(kFaThreads) fa_prefill_kernel(const __grid_constant__ FaArgs a) {
  constexpr int KC = D / 16;   // k chunks of the S product (over head features)
  constexpr int NT = kFaBKV / 8;  // n tiles of S (keys)
  constexpr int DT = D / 8;    // n tiles of O (head features)
  constexpr int CPR = D / 8;   // 16-byte chunks per tile row
  constexpr int TILE = kFaBKV * D * 2;
  extern __shared__ __align__(128) uint8_t smem[];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, g = lane >> 2, t4 = lane & 3;
  // causal: query tile i needs i + 1 key tiles.  Blocks are dispatched in index order, so the LONGEST tiles go first
  // (the last CTAs to start are then the one-tile ones and the grid's tail is short)
  const int q0 = (int)(gridDim.x - 1 - blockIdx.x) * kFaBQ, h = blockIdx.y, b = blockIdx.z;
  const int kvh = h / (a.Hq / a.Hkv);
  const int T = a.offset + a.L;
  const uint32_t sK = smem_u32(smem), sV = sK + 2 * TILE;

  // keys this CTA needs: 0 .. offset + min(q0 + 63, L - 1)
  const int last_key = a.offset + min(q0 + kFaBQ, a.L) - 1;
  const int ntiles = last_key / kFaBKV + 1;

  // a thread copies chunk `cc` of rows rr, rr + RSTEP, ...: when the 64 keys of a stage lie in one page (or the cache is
  // contiguous) and all exist, one address computation serves the stage (as in fa_decode_kernel)
  constexpr int RSTEP = kFaThreads / CPR, NIT = kFaBKV / RSTEP;
  const int cc = tid % CPR, rr = tid / CPR;
  const bool one_page = !a.kv.page_table || a.kv.page_size % kFaBKV == 0;
  const size_t row_step = (size_t)RSTEP * a.kv.tok_stride;
  const uint32_t dst0 = fa_tile_off<D>(rr, cc);   // (rr + k RSTEP) & 7 == rr & 7: the swizzle term does not move
  auto issue = [&](int j) {
    const uint32_t dk = sK + (j & 1) * TILE, dv = sV + (j & 1) * TILE;
    if (one_page && j * kFaBKV + kFaBKV <= T) {
      const size_t off = fa_kv_addr(a.kv, b, kvh, j * kFaBKV + rr) + cc * 8;
#pragma unroll
      for (int k = 0; k < NIT; ++k) {
        cp_async16(dk + dst0 + k * (RSTEP * D * 2), a.kc + off + k * row_step);
        cp_async16(dv + dst0 + k * (RSTEP * D * 2), a.vc + off + k * row_step);
      }
    } else {
      for (int i = tid; i < kFaBKV * CPR; i += kFaThreads) {
        const int r = i / CPR, c = i % CPR;
        const int t = min(j * kFaBKV + r, T - 1);
        const size_t off = fa_kv_addr(a.kv, b, kvh, t) + c * 8;
        cp_async16(dk + fa_tile_off<D>(r, c), a.kc + off);
        cp_async16(dv + fa_tile_off<D>(r, c), a.vc + off);
      }
    }
    cp_async_commit();
  };
  issue(0);

  // Q fragments (A operand, 16 rows x D per warp): rows q0 + 16 warp + g and + 8, straight from global memory
  const int r0 = q0 + warp * 16 + g, r1 = r0 + 8;
  const uint16_t* qb = a.q + ((size_t)b * a.Hq + h) * a.L * D;
  const uint32_t* qr0 = reinterpret_cast<const uint32_t*>(qb + (size_t)min(r0, a.L - 1) * D);
  const uint32_t* qr1 = reinterpret_cast<const uint32_t*>(qb + (size_t)min(r1, a.L - 1) * D);
  uint32_t qf[KC][4];
#pragma unroll
  for (int kc = 0; kc < KC; ++kc) {
    qf[kc][0] = __ldg(qr0 + kc * 8 + t4);
    qf[kc][1] = __ldg(qr1 + kc * 8 + t4);
    qf[kc][2] = __ldg(qr0 + kc * 8 + 4 + t4);
    qf[kc][3] = __ldg(qr1 + kc * 8 + 4 + t4);
  }
  float o[DT][4];
#pragma unroll
  for (int i = 0; i < DT; ++i) { o[i][0] = o[i][1] = o[i][2] = o[i][3] = 0.f; }
  float m0 = -INFINITY, m1 = -INFINITY, l0 = 0.f, l1 = 0.f;
  const int lim0 = a.offset + r0, lim1 = a.offset + r1;  // last key each of this thread's two rows may see

  for (int j = 0; j < ntiles; ++j) {
    if (j + 1 < ntiles) { issue(j + 1); cp_async_wait<1>(); } else { cp_async_wait<0>(); }
    __syncthreads();
    const uint32_t tk = sK + (j & 1) * TILE, tv = sV + (j & 1) * TILE;
    // ---- S = Q K^T : 16 x 64 per warp ----------------------------------------------------------------------
    float s[NT][4];
#pragma unroll
    for (int n = 0; n < NT; ++n) { s[n][0] = s[n][1] = s[n][2] = s[n][3] = 0.f; }
#pragma unroll
    for (int n = 0; n < NT; ++n) {
#pragma unroll
      for (int kc = 0; kc < KC; kc += 2) {
        // 4 matrices: (keys 8n.., features 16kc..+8), (.. +8..+16), (16(kc+1)..+8), (.. +8..+16)
        uint32_t b0, b1, b2, b3;
        ldsm_x4(tk + fa_tile_off<D>(n * 8 + (lane & 7), kc * 2 + (lane >> 3)), b0, b1, b2, b3);
        mma_bf16_16816(s[n], qf[kc][0], qf[kc][1], qf[kc][2], qf[kc][3], b0, b1);
        mma_bf16_16816(s[n], qf[kc + 1][0], qf[kc + 1][1], qf[kc + 1][2], qf[kc + 1][3], b2, b3);
      }
    }
    // ---- scale, causal mask, online softmax ------------------------------------------------------------------
    const int key0 = j * kFaBKV + 2 * t4;
    const bool diag = j * kFaBKV + kFaBKV - 1 > a.offset + q0 + warp * 16;  // some key of the tile may be masked
    float mx0 = -INFINITY, mx1 = -INFINITY;
#pragma unroll
    for (int n = 0; n < NT; ++n) {
#pragma unroll
      for (int e = 0; e < 4; ++e) s[n][e] *= a.scale_log2e;
      if (diag) {
        const int k = key0 + n * 8;
        if (k > lim0) s[n][0] = -INFINITY;
        if (k + 1 > lim0) s[n][1] = -INFINITY;
        if (k > lim1) s[n][2] = -INFINITY;
        if (k + 1 > lim1) s[n][3] = -INFINITY;
      }
      mx0 = fmaxf(mx0, fmaxf(s[n][0], s[n][1]));
      mx1 = fmaxf(mx1, fmaxf(s[n][2], s[n][3]));
    }
    mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 1)); mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 2));
    mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 1)); mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 2));
    // the reference point (m0, m1) of the running sums only moves when a score exceeds it by 8 (log2 units): same sums,
    // and the rescaling multiplies of the accumulators leave the steady-state loop (probabilities stay below 2^8)
    if (mx0 > m0 + 8.f) {
      const float c0 = exp2f(m0 - mx0);   // m0 == -inf: 0, and everything it multiplies is 0
      l0 *= c0;
#pragma unroll
      for (int i = 0; i < DT; ++i) { o[i][0] *= c0; o[i][1] *= c0; }
      m0 = mx0;
    }
    if (mx1 > m1 + 8.f) {
      const float c1 = exp2f(m1 - mx1);
      l1 *= c1;
#pragma unroll
      for (int i = 0; i < DT; ++i) { o[i][2] *= c1; o[i][3] *= c1; }
      m1 = mx1;
    }
    const float ms0 = m0 == -INFINITY ? 0.f : m0, ms1 = m1 == -INFINITY ? 0.f : m1;
    float ps0 = 0.f, ps1 = 0.f;
    uint32_t pf[NT][2];
#pragma unroll
    for (int n = 0; n < NT; ++n) {
      const float p0 = exp2f(s[n][0] - ms0), p1 = exp2f(s[n][1] - ms0);
      const float p2 = exp2f(s[n][2] - ms1), p3 = exp2f(s[n][3] - ms1);
      ps0 += p0 + p1; ps1 += p2 + p3;
      pf[n][0] = pack_bf16(p0, p1);
      pf[n][1] = pack_bf16(p2, p3);
    }
    l0 += ps0; l1 += ps1;
    // ---- O += P V : contraction over the tile's 64 keys ------------------------------------------------------
#pragma unroll
    for (int kk = 0; kk < kFaBKV / 16; ++kk) {
      const uint32_t a0 = pf[2 * kk][0], a1 = pf[2 * kk][1], a2 = pf[2 * kk + 1][0], a3 = pf[2 * kk + 1][1];
#pragma unroll
      for (int dj = 0; dj < DT; dj += 2) {
        // transposed 8x8 blocks: (keys 16kk..+8, features 8dj..), (keys +8..+16, 8dj..), (keys lo, 8(dj+1)..), (keys hi, ..)
        uint32_t b0, b1, b2, b3;
        ldsm_x4_trans(tv + fa_tile_off<D>(kk * 16 + (lane & 7) + ((lane >> 3) & 1) * 8, dj + (lane >> 4)), b0, b1, b2, b3);
        mma_bf16_16816(o[dj], a0, a1, a2, a3, b0, b1);
        mma_bf16_16816(o[dj + 1], a0, a1, a2, a3, b2, b3);
      }
    }
    __syncthreads();  // the stage is refilled two iterations from now
  }
  // ---- finish: row sums over the 4 lanes of a row, normalise, bf16 store [B, L, Hq * D] --------------------------
  l0 += __shfl_xor_sync(0xffffffffu, l0, 1); l0 += __shfl_xor_sync(0xffffffffu, l0, 2);
  l1 += __shfl_xor_sync(0xffffffffu, l1, 1); l1 += __shfl_xor_sync(0xffffffffu, l1, 2);
  const float i0 = 1.f / l0, i1 = 1.f / l1;
  uint16_t* ob = a.out + ((size_t)b * a.L) * a.Hq * D + (size_t)h * D;
#pragma unroll
  for (int i = 0; i < DT; ++i) {
    const int d = i * 8 + 2 * t4;
    if (r0 < a.L) *reinterpret_cast<uint32_t*>(ob + (size_t)r0 * a.Hq * D + d) = pack_bf16(o[i][0] * i0, o[i][1] * i0);
    if (r1 < a.L) *reinterpret_cast<uint32_t*>(ob + (size_t)r1 * a.Hq * D + d) = pack_bf16(o[i][2] * i1, o[i][3] * i1);
  }
}

// ---- decode attention (L = 1) for batched / long-context steps: the n_rep query heads of a kv head are the M rows ------
// Same operand path as the prefill kernel; the 4 warps of a CTA share each 64-key stage and take 16 keys of it each,
// every warp with its own running (max, sum, O) that a shared-memory pass merges at the end.  grid = (splits of the
// context, kv heads, sequences); split partials go through the same workspace records and last-arriver merge as
// decode_attn_kernel ([B * Hq][nsplit][D + 2] floats + one counter per (sequence, kv head)).  K / V bytes are read
// once per kv head (the SIMT kernel reads them once per pair of query heads) and the dot products leave the FMA pipe.
struct FaDecArgs {
  uint16_t* out;
  const uint16_t* q;     // [B, Hq, D], rope applied
  const uint16_t* kc;
  const uint16_t* vc;
  KvLayout kv;
  const int32_t* pos;    // keys 0 .. pos[b] are attended
  int Hq, Hkv, nsplit;
  float scale_log2e;
  float* ws_part;
  unsigned int* ws_cnt;
};

template <int D>
__global__ void __launch_bounds__(kFaThreads) fa_decode_kernel(const __grid_constant__ FaDecArgs a) {
  constexpr int KC = D / 16, DT = D / 8, CPR = D / 8, TILE = kFaBKV * D * 2, PW = D + 2, NS = kFaDecStages;
  extern __shared__ __align__(128) uint8_t smem[];
  __shared__ int s_last;
  __shared__ float s_m[16][kFaMaxSplits], s_l[16][kFaMaxSplits], s_w[16][kFaMaxSplits];
  pdl_launch_dependents();
  pdl_wait();
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, g = lane >> 2, t4 = lane & 3;
  const int split = blockIdx.x, kvh = blockIdx.y, b = blockIdx.z;
  const int n_rep = a.Hq / a.Hkv, qh0 = kvh * n_rep;
  const int p = a.pos[b], ctx = p + 1;
  const int tiles = (ctx + kFaBKV - 1) / kFaBKV, per = (tiles + a.nsplit - 1) / a.nsplit;
  const int j0 = split * per, j1 = min(tiles, j0 + per);
  const uint32_t sK = smem_u32(smem), sV = sK + NS * TILE;

  // a thread copies chunk `cc` of rows rr, rr + RSTEP, ...: when the 64 keys of a stage lie in one page (or the cache is
  // contiguous) and all exist, one address computation serves the whole stage and the rest are constant offsets
  constexpr int RSTEP = kFaThreads / CPR, NIT = kFaBKV / RSTEP;
  const int cc = tid % CPR, rr = tid / CPR;
  const bool one_page = !a.kv.page_table || a.kv.page_size % kFaBKV == 0;
  const size_t row_step = (size_t)RSTEP * a.kv.tok_stride;
  const uint32_t dst0 = fa_tile_off<D>(rr, cc);   // (rr + k RSTEP) & 7 == rr & 7: the swizzle term does not move
  auto issue = [&](int j) {  // always commits one group (empty past the end): the wait below counts groups
    const uint32_t dk = sK + ((j - j0) % NS) * TILE, dv = sV + ((j - j0) % NS) * TILE;
    if (j < j1 && one_page && j * kFaBKV + kFaBKV - 1 <= p) {
      const size_t off = fa_kv_addr(a.kv, b, kvh, j * kFaBKV + rr) + cc * 8;
#pragma unroll
      for (int k = 0; k < NIT; ++k) {
        cp_async16(dk + dst0 + k * (RSTEP * D * 2), a.kc + off + k * row_step);
        cp_async16(dv + dst0 + k * (RSTEP * D * 2), a.vc + off + k * row_step);
      }
    } else {
      for (int i = tid; j < j1 && i < kFaBKV * CPR; i += kFaThreads) {
        const int r = i / CPR, c = i % CPR;
        const int t = min(j * kFaBKV + r, p);
        const size_t off = fa_kv_addr(a.kv, b, kvh, t) + c * 8;
        cp_async16(dk + fa_tile_off<D>(r, c), a.kc + off);
        cp_async16(dv + fa_tile_off<D>(r, c), a.vc + off);
      }
    }
    cp_async_commit();
  };
  for (int s = 0; s < NS - 1; ++s) issue(j0 + s);

  // Q fragments: row g = query head qh0 + g (rows >= n_rep are zero)
  uint32_t qf[KC][4];
  {
    const uint32_t* q0 = reinterpret_cast<const uint32_t*>(a.q + ((size_t)b * a.Hq + qh0 + min(g, n_rep - 1)) * D);
    const uint32_t* q1 = reinterpret_cast<const uint32_t*>(a.q + ((size_t)b * a.Hq + qh0 + min(g + 8, n_rep - 1)) * D);
    const bool v0 = g < n_rep, v1 = g + 8 < n_rep;
#pragma unroll
    for (int kc = 0; kc < KC; ++kc) {
      qf[kc][0] = v0 ? ldcg32(q0 + kc * 8 + t4) : 0u;
      qf[kc][1] = v1 ? ldcg32(q1 + kc * 8 + t4) : 0u;
      qf[kc][2] = v0 ? ldcg32(q0 + kc * 8 + 4 + t4) : 0u;
      qf[kc][3] = v1 ? ldcg32(q1 + kc * 8 + 4 + t4) : 0u;
    }
  }
  float o[DT][4];
#pragma unroll
  for (int i = 0; i < DT; ++i) { o[i][0] = o[i][1] = o[i][2] = o[i][3] = 0.f; }
  float m0 = -INFINITY, m1 = -INFINITY, l0 = 0.f, l1 = 0.f;

  for (int j = j0; j < j1; ++j) {
    issue(j + NS - 1);
    cp_async_wait<NS - 1>();
    __syncthreads();
    const uint32_t tk = sK + ((j - j0) % NS) * TILE, tv = sV + ((j - j0) % NS) * TILE;
    float s[2][4];
#pragma unroll
    for (int n = 0; n < 2; ++n) {
      s[n][0] = s[n][1] = s[n][2] = s[n][3] = 0.f;
#pragma unroll
      for (int kc = 0; kc < KC; kc += 2) {
        uint32_t b0, b1, b2, b3;
        ldsm_x4(tk + fa_tile_off<D>(warp * 16 + n * 8 + (lane & 7), kc * 2 + (lane >> 3)), b0, b1, b2, b3);
        mma_bf16_16816(s[n], qf[kc][0], qf[kc][1], qf[kc][2], qf[kc][3], b0, b1);
        mma_bf16_16816(s[n], qf[kc + 1][0], qf[kc + 1][1], qf[kc + 1][2], qf[kc + 1][3], b2, b3);
      }
    }
    const int key0 = j * kFaBKV + warp * 16 + 2 * t4;
    float mx0 = -INFINITY, mx1 = -INFINITY;
#pragma unroll
    for (int n = 0; n < 2; ++n) {
#pragma unroll
      for (int e = 0; e < 4; ++e) s[n][e] *= a.scale_log2e;
      const int k = key0 + n * 8;
      if (k > p) { s[n][0] = -INFINITY; s[n][2] = -INFINITY; }
      if (k + 1 > p) { s[n][1] = -INFINITY; s[n][3] = -INFINITY; }
      mx0 = fmaxf(mx0, fmaxf(s[n][0], s[n][1]));
      mx1 = fmaxf(mx1, fmaxf(s[n][2], s[n][3]));
    }
    mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 1)); mx0 = fmaxf(mx0, __shfl_xor_sync(0xffffffffu, mx0, 2));
    mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 1)); mx1 = fmaxf(mx1, __shfl_xor_sync(0xffffffffu, mx1, 2));
    // the reference point (m0, m1) of the running sums only moves when a score exceeds it by 8 (log2 units): same sums,
    // and the 64 rescaling multiplies of the accumulators leave the steady-state loop (probabilities stay below 2^8)
    if (mx0 > m0 + 8.f) {
      const float c0 = exp2f(m0 - mx0);   // m0 == -inf: 0, and everything it multiplies is 0
      l0 *= c0;
#pragma unroll
      for (int i = 0; i < DT; ++i) { o[i][0] *= c0; o[i][1] *= c0; }
      m0 = mx0;
    }
    if (mx1 > m1 + 8.f) {
      const float c1 = exp2f(m1 - mx1);
      l1 *= c1;
#pragma unroll
      for (int i = 0; i < DT; ++i) { o[i][2] *= c1; o[i][3] *= c1; }
      m1 = mx1;
    }
    const float ms0 = m0 == -INFINITY ? 0.f : m0, ms1 = m1 == -INFINITY ? 0.f : m1;
    const float p00 = exp2f(s[0][0] - ms0), p01 = exp2f(s[0][1] - ms0), p02 = exp2f(s[0][2] - ms1), p03 = exp2f(s[0][3] - ms1);
    const float p10 = exp2f(s[1][0] - ms0), p11 = exp2f(s[1][1] - ms0), p12 = exp2f(s[1][2] - ms1), p13 = exp2f(s[1][3] - ms1);
    l0 += (p00 + p01 + p10 + p11);
    l1 += (p02 + p03 + p12 + p13);
    // P as a bf16 pair (hi + lo): 16 significant bits, so this kernel agrees with the fp32-probability SIMT kernel to
    // ~1e-4 and greedy tokens do not depend on which of the two the engine picked; the extra 16 MMAs ride under the
    // K/V stream (the kernel is HBM-bound)
    const uint32_t a0 = pack_bf16(p00, p01), a1 = pack_bf16(p02, p03), a2 = pack_bf16(p10, p11), a3 = pack_bf16(p12, p13);
    const uint32_t e0 = pack_bf16(p00 - bf16lo(a0), p01 - bf16hi(a0)), e1 = pack_bf16(p02 - bf16lo(a1), p03 - bf16hi(a1));
    const uint32_t e2 = pack_bf16(p10 - bf16lo(a2), p11 - bf16hi(a2)), e3 = pack_bf16(p12 - bf16lo(a3), p13 - bf16hi(a3));
#pragma unroll
    for (int dj = 0; dj < DT; dj += 2) {
      uint32_t b0, b1, b2, b3;
      ldsm_x4_trans(tv + fa_tile_off<D>(warp * 16 + (lane & 7) + ((lane >> 3) & 1) * 8, dj + (lane >> 4)), b0, b1, b2, b3);
      mma_bf16_16816(o[dj], a0, a1, a2, a3, b0, b1);
      mma_bf16_16816(o[dj + 1], a0, a1, a2, a3, b2, b3);
      mma_bf16_16816(o[dj], e0, e1, e2, e3, b0, b1);
      mma_bf16_16816(o[dj + 1], e0, e1, e2, e3, b2, b3);
    }
    __syncthreads();
  }
  l0 += __shfl_xor_sync(0xffffffffu, l0, 1); l0 += __shfl_xor_sync(0xffffffffu, l0, 2);
  l1 += __shfl_xor_sync(0xffffffffu, l1, 1); l1 += __shfl_xor_sync(0xffffffffu, l1, 2);
  // per-warp partials -> shared memory (the stages are free: the loop ended with a barrier), rows < n_rep only
  float* rec = reinterpret_cast<float*>(smem);  // [4 warps][n_rep][PW]
  if (g < n_rep) {
    float* r0 = rec + ((size_t)warp * n_rep + g) * PW;
#pragma unroll
    for (int i = 0; i < DT; ++i) { r0[i * 8 + 2 * t4] = o[i][0]; r0[i * 8 + 2 * t4 + 1] = o[i][1]; }
    if (t4 == 0) { r0[D] = m0; r0[D + 1] = l0; }
  }
  if (g + 8 < n_rep) {
    float* r1 = rec + ((size_t)warp * n_rep + g + 8) * PW;
#pragma unroll
    for (int i = 0; i < DT; ++i) { r1[i * 8 + 2 * t4] = o[i][2]; r1[i * 8 + 2 * t4 + 1] = o[i][3]; }
    if (t4 == 0) { r1[D] = m1; r1[D + 1] = l1; }
  }
  __syncthreads();
  const size_t rec0 = ((size_t)b * a.Hq + qh0) * a.nsplit;
  for (int i = tid; i < n_rep * D; i += kFaThreads) {
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
  const int cidx = b * a.Hkv + kvh;
  if (tid == 0) s_last = (atomicAdd(&a.ws_cnt[cidx], 1u) == (unsigned)a.nsplit - 1);
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  // merge of the splits: (max, sum) of every record first (one parallel round trip), then the weighted rows with the
  // loads of 8 splits in flight per thread
  for (int i = tid; i < n_rep * a.nsplit; i += kFaThreads) {
    const int r = i / a.nsplit, sp = i % a.nsplit;
    const float* pr = a.ws_part + (rec0 + (size_t)r * a.nsplit + sp) * PW;
    s_m[r][sp] = ldcg_f32(pr + D);
    s_l[r][sp] = ldcg_f32(pr + D + 1);
  }
  __syncthreads();
  for (int i = tid; i < n_rep * a.nsplit; i += kFaThreads) {
    const int r = i / a.nsplit, sp = i % a.nsplit;
    float mn = -INFINITY;
    for (int k = 0; k < a.nsplit; ++k) mn = fmaxf(mn, s_m[r][k]);
    s_w[r][sp] = s_m[r][sp] == -INFINITY ? 0.f : exp2f(s_m[r][sp] - mn);
  }
  __syncthreads();
  for (int i = tid; i < n_rep * D; i += kFaThreads) {
    const int r = i / D, d = i % D;
    const float* pr = a.ws_part + (rec0 + (size_t)r * a.nsplit) * PW + d;
    float lt = 0.f, at = 0.f;
#pragma unroll 8
    for (int sp = 0; sp < a.nsplit; ++sp) {  // split order: deterministic
      lt += s_l[r][sp] * s_w[r][sp];
      at += ldcg_f32(pr + (size_t)sp * PW) * s_w[r][sp];
    }
    a.out[((size_t)b * a.Hq + qh0 + r) * D + d] = f2bf(at / lt);
  }
  if (tid == 0) a.ws_cnt[cidx] = 0;
}

cudaError_t launch_decode_attn_tc(uint16_t* out, const uint16_t* q, const uint16_t* kc, const uint16_t* vc,
                                  const KvLayout& kv, const int32_t* pos, int B, int Hq, int Hkv, int D, float scale,
                                  int nsplit, float* ws, cudaStream_t st) {
  if (nsplit > kFaMaxSplits) return cudaErrorInvalidValue;
  FaDecArgs a;
  a.out = out; a.q = q; a.kc = kc; a.vc = vc; a.kv = kv; a.pos = pos; a.Hq = Hq; a.Hkv = Hkv; a.nsplit = nsplit;
  a.scale_log2e = scale * 1.4426950408889634f;
  const size_t cnt_bytes = (((size_t)B * Hq * sizeof(unsigned)) + 255) & ~(size_t)255;
  a.ws_cnt = reinterpret_cast<unsigned int*>(ws);
  a.ws_part = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(ws) + cnt_bytes);
  const size_t smem = (size_t)2 * kFaDecStages * kFaBKV * D * 2;
  {  // per launch, not once per process: the attribute is per device
    cudaError_t e = D == 128 ? cudaFuncSetAttribute(fa_decode_kernel<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)
                             : cudaFuncSetAttribute(fa_decode_kernel<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
  }
  const dim3 grid(nsplit, Hkv, B);
  count_launch();
  return D == 128 ? launch_pdl(fa_decode_kernel<128>, grid, dim3(kFaThreads), smem, st, a)
                  : launch_pdl(fa_decode_kernel<64>, grid, dim3(kFaThreads), smem, st, a);
}

cudaError_t launch_prefill_attn_tc(uint16_t* out, const uint16_t* q, const uint16_t* kc, const uint16_t* vc,
                                   const KvLayout& kv, int B, int Hq, int Hkv, int L, int D, int offset, float scale,
                                   cudaStream_t st) {
  FaArgs a;
  a.out = out; a.q = q; a.kc = kc; a.vc = vc; a.kv = kv; a.Hq = Hq; a.Hkv = Hkv; a.L = L; a.offset = offset;
  a.scale_log2e = scale * 1.4426950408889634f;
  const size_t smem = (size_t)4 * kFaBKV * D * 2;  // K and V, double-buffered
  if (D == 128) {  // per launch, not once per process: the attribute is per device (head_dim 64 fits the default 48 KB)
    cudaError_t e = cudaFuncSetAttribute(fa_prefill_kernel<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
  }
  const dim3 grid((L + kFaBQ - 1) / kFaBQ, Hq, B);
  count_launch();
  if (D == 128) fa_prefill_kernel<128><<<grid, kFaThreads, smem, st>>>(a);
  else fa_prefill_kernel<64><<<grid, kFaThreads, smem, st>>>(a);
  return cudaGetLastError();
}

}  // namespace pie
