// Main loop, epilogues and launcher of the decode GEMV (included by pie_qmv.cu).
namespace pie {

template <int BITS>
__global__ void __launch_bounds__(256) qmv_kernel(const __grid_constant__ QmvArgs a) {
  using G = QmvGeom<BITS>;
  extern __shared__ __align__(128) uint8_t smem[];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, W = blockDim.x >> 5;
  const int g = lane >> 2, t = lane & 3;
  const int tile = blockIdx.x, ks = blockIdx.y;
  const int k0 = ks * a.kslice;
  const int klen = min(a.K, k0 + a.kslice) - k0;
  const int nchunks = klen / G::CK;

  const int xrow = qmv_xrow_bytes(a.kslice);
  uint8_t* xs = smem;
  float* sumx = reinterpret_cast<float*>(smem + (((size_t)a.M * xrow + 127) & ~(size_t)127));
  float* red = sumx + (a.kslice / kGroup) * 8;
  uint8_t* ring = reinterpret_cast<uint8_t*>(red + W * 128) + (size_t)warp * kQmvStages * G::STAGE_BYTES;
  ring = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(ring) + 127) & ~(uintptr_t)127);
  const uint32_t ring_s = smem_u32(ring);

  // ---- which weight rows this tile covers -------------------------------------------------------
  int sidx = 0;
  if (!a.dual)
    for (int s = 1; s < a.nseg; ++s)
      if (tile >= a.seg[s].tile0) sidx = s;
  const int nbase = a.dual ? tile * 8 : (tile - a.seg[sidx].tile0) * kTileRows;
  const int gpr = a.K / kGroup;  // groups per weight row
  auto row_of = [&](int r, int& seg_i) -> int {  // tile row r -> (segment, clamped row index)
    int n;
    if (a.dual) { seg_i = r >> 3; n = nbase + (r & 7); } else { seg_i = sidx; n = nbase + r; }
    return min(n, a.seg[seg_i].N - 1);
  };
  // cp.async sources: piece p = lane + 32 i -> tile row p >> 3, 16-byte chunk p & 7
  const uint8_t* wsrc[4];
  uint32_t wdst[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int p = lane + 32 * i, r = p >> 3, c = p & 7;
    int si;
    const int n = row_of(r, si);
    wsrc[i] = reinterpret_cast<const uint8_t*>(a.seg[si].wq) + ((size_t)n * a.K + k0) * BITS / 8 + c * 16;
    wdst[i] = r * kRowBytes + swz_chunk<BITS>(r, c);
  }
  const uint8_t* sbsrc;
  uint32_t sbdst;
  {
    const int r = lane & 15, which = lane >> 4;
    int si;
    const int n = row_of(r, si);
    const uint16_t* base = which ? a.seg[si].biases : a.seg[si].scales;
    sbsrc = reinterpret_cast<const uint8_t*>(base + (size_t)n * gpr + k0 / kGroup);
    sbdst = kTileRows * kRowBytes + which * G::SB_BYTES + r * G::GPC * 2;
  }
  auto issue = [&](int chunk, int stage) {
    const uint32_t sbase = ring_s + stage * G::STAGE_BYTES;
#pragma unroll
    for (int i = 0; i < 4; ++i) cp_async16(sbase + wdst[i], wsrc[i] + (size_t)chunk * kRowBytes);
    if (BITS == 4) cp_async8(sbase + sbdst, sbsrc + (size_t)chunk * G::GPC * 2);
    else cp_async4(sbase + sbdst, sbsrc + (size_t)chunk * G::GPC * 2);
  };

  // ---- prologue: weights first (they never depend on the previous kernel), then x ------------------
  int my_n = 0;
  for (int c = warp; c < nchunks; c += W) ++my_n;
#pragma unroll
  for (int s = 0; s < kQmvStages - 1; ++s) {
    if (s < my_n) issue(warp + s * W, s);
    cp_async_commit();
  }
  pdl_launch_dependents();
  pdl_wait();
  stage_x<BITS>(a, k0, klen, xs, sumx, red);
  __syncthreads();

  float y[4] = {0.f, 0.f, 0.f, 0.f};
  const uint32_t xs_s = smem_u32(xs) + g * xrow + t * 32;
  const bool has_x = g < a.M;
  for (int i = 0; i < my_n; ++i) {
    const int nxt = i + kQmvStages - 1;
    if (nxt < my_n) issue(warp + nxt * W, nxt % kQmvStages);
    cp_async_commit();
    cp_async_wait<kQmvStages - 1>();
    __syncwarp();
    const int chunk = warp + i * W;
    const uint32_t sbase = ring_s + (i % kQmvStages) * G::STAGE_BYTES;
    const uint32_t sb = sbase + kTileRows * kRowBytes;
    float sc[2][G::GPC], bi[2][G::GPC];
    if (BITS == 4) {
      uint2 s0 = lds64(sb + g * 8), s1 = lds64(sb + (g + 8) * 8);
      uint2 b0 = lds64(sb + G::SB_BYTES + g * 8), b1 = lds64(sb + G::SB_BYTES + (g + 8) * 8);
      sc[0][0] = bf16lo(s0.x); sc[0][1] = bf16hi(s0.x); sc[0][2 % G::GPC] = bf16lo(s0.y); sc[0][3 % G::GPC] = bf16hi(s0.y);
      sc[1][0] = bf16lo(s1.x); sc[1][1] = bf16hi(s1.x); sc[1][2 % G::GPC] = bf16lo(s1.y); sc[1][3 % G::GPC] = bf16hi(s1.y);
      bi[0][0] = bf16lo(b0.x); bi[0][1] = bf16hi(b0.x); bi[0][2 % G::GPC] = bf16lo(b0.y); bi[0][3 % G::GPC] = bf16hi(b0.y);
      bi[1][0] = bf16lo(b1.x); bi[1][1] = bf16hi(b1.x); bi[1][2 % G::GPC] = bf16lo(b1.y); bi[1][3 % G::GPC] = bf16hi(b1.y);
    } else {
      uint32_t s0 = lds32(sb + g * 4), s1 = lds32(sb + (g + 8) * 4);
      uint32_t b0 = lds32(sb + G::SB_BYTES + g * 4), b1 = lds32(sb + G::SB_BYTES + (g + 8) * 4);
      sc[0][0] = bf16lo(s0); sc[0][1] = bf16hi(s0); sc[1][0] = bf16lo(s1); sc[1][1] = bf16hi(s1);
      bi[0][0] = bf16lo(b0); bi[0][1] = bf16hi(b0); bi[1][0] = bf16lo(b1); bi[1][1] = bf16hi(b1);
    }
#pragma unroll
    for (int j = 0; j < G::GPC; ++j) {
      const int kg = chunk * G::GPC + j;  // group index inside the slice
      uint4 x0 = make_uint4(0, 0, 0, 0), x1 = x0;
      if (has_x) { x0 = lds128(xs_s + kg * 128); x1 = lds128(xs_s + kg * 128 + 16); }
      const float2 sx = *reinterpret_cast<const float2*>(sumx + kg * 8 + 2 * t);
      const uint32_t xb[8] = {x0.x, x0.y, x0.z, x0.w, x1.x, x1.y, x1.z, x1.w};
      float acc[4] = {0.f, 0.f, 0.f, 0.f};
      float off;
      if (BITS == 4) {
        const uint32_t wa_addr = sbase + g * kRowBytes + ((j ^ (g & 3)) * 32) + t * 8;
        const uint2 wa = lds64(wa_addr), wb = lds64(wa_addr + 8 * kRowBytes);
#pragma unroll
        for (int m = 0; m < 4; ++m)
          mma_bf16_16816(acc, lift4(wa.x, 4 * m), lift4(wb.x, 4 * m), lift4(wa.y, 4 * m), lift4(wb.y, 4 * m),
                         xb[2 * m], xb[2 * m + 1]);
        off = 128.f;
      } else {
        const uint32_t wa_addr = sbase + g * kRowBytes + ((j ^ (g & 1)) * 64) + t * 16;
        const uint4 wa = lds128(wa_addr), wb = lds128(wa_addr + 8 * kRowBytes);
        const uint32_t ua[4] = {wa.x, wa.y, wa.z, wa.w}, ub[4] = {wb.x, wb.y, wb.z, wb.w};
        float hi[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
        for (int m = 0; m < 4; ++m) {
          mma_bf16_16816(acc, lift4(ua[m], 0), lift4(ub[m], 0), lift4(ua[m], 8), lift4(ub[m], 8), xb[2 * m],
                         xb[2 * m + 1]);
          mma_bf16_16816(hi, lift4(ua[m], 4), lift4(ub[m], 4), lift4(ua[m], 12), lift4(ub[m], 12), xb[2 * m],
                         xb[2 * m + 1]);
        }
#pragma unroll
        for (int q = 0; q < 4; ++q) acc[q] = fmaf(16.f, hi[q], acc[q]);
        off = 128.f + 2048.f;
      }
      const float bb0 = fmaf(-off, sc[0][j], bi[0][j]), bb1 = fmaf(-off, sc[1][j], bi[1][j]);
      y[0] += sc[0][j] * acc[0] + bb0 * sx.x;
      y[1] += sc[0][j] * acc[1] + bb0 * sx.y;
      y[2] += sc[1][j] * acc[2] + bb1 * sx.x;
      y[3] += sc[1][j] * acc[3] + bb1 * sx.y;
    }
    __syncwarp();
  }
  cp_async_wait<0>();

  // ---- cross-warp reduction (fixed order) ---------------------------------------------------------
  float* myred = red + warp * 128;
  *reinterpret_cast<float2*>(myred + g * 8 + 2 * t) = make_float2(y[0], y[1]);
  *reinterpret_cast<float2*>(myred + (g + 8) * 8 + 2 * t) = make_float2(y[2], y[3]);
  __syncthreads();
  const int r = tid & 15, m = (tid >> 4) & 7;
  float v = 0.f;
  if (tid < 128)
    for (int w = 0; w < W; ++w) v += red[w * 128 + r * 8 + m];

  if (a.ksplit > 1) {  // deterministic split-K fix-up: last CTA of the tile sums the slices in order
    __shared__ int s_last;
    if (tid < 128) a.ws_part[((size_t)tile * a.ksplit + ks) * 128 + tid] = v;
    __threadfence();
    __syncthreads();
    if (tid == 0) s_last = (atomicAdd(&a.ws_cnt[tile], 1u) == (unsigned)a.ksplit - 1);
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    if (tid < 128) {
      v = 0.f;
      for (int s = 0; s < a.ksplit; ++s) v += ldcg_f32(a.ws_part + ((size_t)tile * a.ksplit + s) * 128 + tid);
    }
    if (tid == 0) a.ws_cnt[tile] = 0;
  }

  // ---- epilogue -----------------------------------------------------------------------------------
  if (a.epi == EPI_SILUMUL) {
    __syncthreads();
    if (tid < 128) red[tid] = v;  // [m][16]: rows 0-7 gate, 8-15 up
    __syncthreads();
    if (tid < 128 && r < 8 && m < a.M) {
      const int n = nbase + r;
      if (n < a.seg[0].N) {
        const float gt = bfround(red[m * 16 + r]), up = bfround(red[m * 16 + r + 8]);
        const float sg = bfround(1.f / (1.f + expf(-gt)));  // nn.silu = x * sigmoid(x), each op rounded
        a.out[(size_t)m * a.out_ld + n] = f2bf(bfround(gt * sg) * up);
      }
    }
    return;
  }
  if (tid < 128 && m < a.M) {
    const int n = nbase + r;
    if (n < a.seg[sidx].N) {
      const size_t o = (size_t)m * a.out_ld + a.seg[sidx].out_off + n;
      float res = bfround(v);
      if (a.epi == EPI_RESADD) res = bf16lo((uint32_t)a.residual[o]) + res;
      a.out[o] = f2bf(res);
    }
  }
}

}  // namespace pie
#include "pie_qmv_launch.inl"
