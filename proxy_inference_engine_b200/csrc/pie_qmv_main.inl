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

template <int BITS, int MT, bool ONECOL>
__global__ void __launch_bounds__(MT == 4 ? 256 : 512, 1) qmv_kernel(const __grid_constant__ QmvArgs a) {
  using G = QmvGeom<BITS>;
  constexpr int S = kQmvStages;
  constexpr int NS = kQmvSlots;
  extern __shared__ __align__(128) uint8_t smem[];
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
  const int my_tiles = tile_first < a.ntiles ? (a.ntiles - 1 - tile_first) / tstride + 1 : 0;
  const int total = my_tiles * nch;

  // ---- shared-memory carve-up (must match qmv_smem_bytes) ----------------------------------------
  const int xrow = qmv_xrow_bytes(a.kslice);
  uint8_t* xs = smem;
  float* sumx = reinterpret_cast<float*>(smem + (((size_t)a.M * xrow + 127) & ~(size_t)127));
  float4* exch = reinterpret_cast<float4*>(sumx + (a.kslice / kGroup) * 8 * MT);   // [W][NS][MT][32]
  uint64_t* bars = reinterpret_cast<uint64_t*>(exch + (size_t)W * NS * MT * 32);   // [nteams][2][NS]
  uint8_t* ring = reinterpret_cast<uint8_t*>(bars + 64) + (size_t)warp * S * G::STAGE_BYTES;
  ring = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(ring) + 127) & ~(uintptr_t)127);
  const uint32_t ring_s = smem_u32(ring);
  uint64_t* full_bar = bars + team * 2 * NS;   // finisher waits: ks_w - 1 arrivals
  uint64_t* empty_bar = full_bar + NS;         // writers wait: 1 arrival (the finisher has read the slot)
  if (ks_w > 1 && wt == 0 && lane == 0) {
    for (int i = 0; i < NS; ++i) { mbar_init(full_bar + i, ks_w - 1); mbar_init(empty_bar + i, 1); }
    fence_mbar_init();
  }

  const int gpr = a.K / kGroup;  // groups per weight row
  // cp.async pieces of a stage: piece p = lane + 32 i -> tile row p >> 3, 16-byte chunk p & 7
  uint32_t wdst[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int p = lane + 32 * i, r = p >> 3, c = p & 7;
    wdst[i] = r * kRowBytes + swz_chunk<BITS>(r, c);
  }
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
  const uint8_t* wsrc[4];
  const uint8_t* sbsrc;
  auto set_tile_ptrs = [&](int tile) {
    int sidx, nbase;
    tile_seg(tile, sidx, nbase);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int p = lane + 32 * i, r = p >> 3, c = p & 7;
      const int si = a.dual ? (r >> 3) : sidx;
      const int n = min(a.dual ? nbase + (r & 7) : nbase + r, a.seg[si].N - 1);
      wsrc[i] = reinterpret_cast<const uint8_t*>(a.seg[si].wq) + ((size_t)n * a.K + kw0) * BITS / 8 + c * 16;
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
  auto issue_next = [&]() {
    if (p_todo > 0) {
      const uint32_t sbase = ring_s + p_soff;
#pragma unroll
      for (int i = 0; i < 4; ++i) { cp_async16(sbase + wdst[i], wsrc[i]); wsrc[i] += kRowBytes; }
      if (BITS == 4) cp_async8(sbase + sbdst, sbsrc); else cp_async4(sbase + sbdst, sbsrc);
      sbsrc += G::GPC * 2;
      p_soff = p_soff == (S - 1) * G::STAGE_BYTES ? 0u : p_soff + G::STAGE_BYTES;
      --p_todo;
      if (--p_left == 0) {
        p_left = nch;
        p_tile += tstride;
        if (p_todo > 0) set_tile_ptrs(p_tile);
      }
    }
    cp_async_commit();
  };

  // ---- prologue: weights first (they never depend on the previous kernel), then x --------------------
#pragma unroll
  for (int s = 0; s < S - 1; ++s) issue_next();
  pdl_launch_dependents();
  pdl_wait();
  stage_x<BITS, MT>(a, k0, a.kslice, xs, sumx, reinterpret_cast<float*>(exch));
  __syncthreads();
#include "pie_qmv_loop.inl"
}

}  // namespace pie
#include "pie_qmv_launch.inl"
