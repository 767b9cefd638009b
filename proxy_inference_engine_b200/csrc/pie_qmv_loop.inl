  // ---- body of qmv_kernel (included inside the function by pie_qmv_body.cuh): consume the ring ----------
  float y[MT][4];
#pragma unroll
  for (int mt = 0; mt < MT; ++mt) { y[mt][0] = y[mt][1] = y[mt][2] = y[mt][3] = 0.f; }
  // columns >= M of the MMA are never stored: they may read any staged row (row 0)
  uint32_t xs_row[MT];
#pragma unroll
  for (int mt = 0; mt < MT; ++mt)
    xs_row[mt] = smem_u32(xs) + (g + 8 * mt < a.M ? g + 8 * mt : 0) * xrow + t * 32 + ((kw0 - k0) >> 6) * 128;
  const uint32_t sumx_s0 = smem_u32(sumx) + (((kw0 - k0) >> 6) * MT * 8 + 2 * t) * 4;
  const uint32_t lmask = a.lift_mask, lmagic = a.lift_magic;  // in registers: one LOP3 per lift
  // per-lane offsets of the weight words of group j inside a stage (swizzled, see swz_chunk)
  uint32_t woff[G::GPC];
#pragma unroll
  for (int j = 0; j < G::GPC; ++j)
    woff[j] = g * kRowBytes + (BITS == 4 ? ((j ^ (g & 3)) * 32) + t * 8 : ((j ^ (g & 1)) * 64) + t * 16);
  const uint32_t sboff = kTileRows * kRowBytes + g * G::GPC * 2;
  const uint32_t exch_s = smem_u32(exch);

  int c_tile = tile_first, c_left = nch, c_iter = 0;
  uint32_t c_soff = 0, xoff = 0;  // xoff: byte offset of the current group block in x (128 B per group)
  uint32_t res_raw[MT][4];        // residual values of the finisher, fetched one stage early
  for (int f = 0; f < total; ++f) {
    issue_next();
    cp_async_wait<S - 1>();
    __syncwarp();
    const bool last_stage = c_left == 1;
    const bool finisher = (c_iter & (ks_w - 1)) == wt;
    int sidx = 0, nbase = 0;
    if (last_stage) {
      tile_seg(c_tile, sidx, nbase);
      if (finisher && a.epi == EPI_RESADD) {
#pragma unroll
        for (int mt = 0; mt < MT; ++mt)
#pragma unroll
          for (int q = 0; q < 4; ++q) {
            const int m = mt * 8 + 2 * t + (q & 1), n = nbase + g + (q >> 1) * 8;
            res_raw[mt][q] = 0;
            if (m < a.M && n < a.seg[sidx].N && !(ONECOL && (q & 1)))
              res_raw[mt][q] = COHERENT ? (uint32_t)ldcg16(a.residual + (size_t)m * a.out_ld + a.seg[sidx].out_off + n)
                                        : (uint32_t)a.residual[(size_t)m * a.out_ld + a.seg[sidx].out_off + n];
          }
      }
    }
    const uint32_t sbase = ring_s + c_soff;
    const uint32_t sb = sbase + sboff;
    // packed bf16 scales / biases of tile rows g (index 0) and g + 8 (index 1): unpacked per group, not up front
    uint32_t scp[2][2], bip[2][2];
    if (BITS == 4) {
      const uint2 s0 = lds64(sb), s1 = lds64(sb + 64);
      const uint2 b0 = lds64(sb + G::SB_BYTES), b1 = lds64(sb + G::SB_BYTES + 64);
      scp[0][0] = s0.x; scp[0][1] = s0.y; scp[1][0] = s1.x; scp[1][1] = s1.y;
      bip[0][0] = b0.x; bip[0][1] = b0.y; bip[1][0] = b1.x; bip[1][1] = b1.y;
    } else {
      scp[0][0] = lds32(sb); scp[1][0] = lds32(sb + 32);
      bip[0][0] = lds32(sb + G::SB_BYTES); bip[1][0] = lds32(sb + G::SB_BYTES + 32);
      scp[0][1] = scp[1][1] = bip[0][1] = bip[1][1] = 0;
    }
#pragma unroll
    for (int j = 0; j < G::GPC; ++j) {
      uint32_t xb[MT][8];
      float2 sx[MT];
#pragma unroll
      for (int mt = 0; mt < MT; ++mt) {
        const uint4 x0 = lds128(xs_row[mt] + xoff + j * 128), x1 = lds128(xs_row[mt] + xoff + j * 128 + 16);
        xb[mt][0] = x0.x; xb[mt][1] = x0.y; xb[mt][2] = x0.z; xb[mt][3] = x0.w;
        xb[mt][4] = x1.x; xb[mt][5] = x1.y; xb[mt][6] = x1.z; xb[mt][7] = x1.w;
        const uint2 sxr = lds64(sumx_s0 + (xoff >> 2) * MT + (j * MT + mt) * 32);  // sumx[(kg * MT + mt) * 8 + 2 t]
        sx[mt] = make_float2(__uint_as_float(sxr.x), __uint_as_float(sxr.y));
      }
      float acc[MT][4];
#pragma unroll
      for (int mt = 0; mt < MT; ++mt) { acc[mt][0] = acc[mt][1] = acc[mt][2] = acc[mt][3] = 0.f; }
      float off;
      if (BITS == 4) {
        const uint2 wa = lds64(sbase + woff[j]), wb = lds64(sbase + woff[j] + 8 * kRowBytes);
#pragma unroll
        for (int m = 0; m < 4; ++m) {
          const uint32_t a0 = lift4(wa.x, 4 * m, lmask, lmagic), a1 = lift4(wb.x, 4 * m, lmask, lmagic),
                         a2 = lift4(wa.y, 4 * m, lmask, lmagic), a3 = lift4(wb.y, 4 * m, lmask, lmagic);
#pragma unroll
          for (int mt = 0; mt < MT; ++mt) mma_bf16_16816(acc[mt], a0, a1, a2, a3, xb[mt][2 * m], xb[mt][2 * m + 1]);
        }
        off = 128.f;
      } else {
        const uint4 wa = lds128(sbase + woff[j]), wb = lds128(sbase + woff[j] + 8 * kRowBytes);
        const uint32_t ua[4] = {wa.x, wa.y, wa.z, wa.w}, ub[4] = {wb.x, wb.y, wb.z, wb.w};
        float hi[MT][4];
#pragma unroll
        for (int mt = 0; mt < MT; ++mt) { hi[mt][0] = hi[mt][1] = hi[mt][2] = hi[mt][3] = 0.f; }
#pragma unroll
        for (int m = 0; m < 4; ++m) {
          const uint32_t l0 = lift4(ua[m], 0, lmask, lmagic), l1 = lift4(ub[m], 0, lmask, lmagic),
                         l2 = lift4(ua[m], 8, lmask, lmagic), l3 = lift4(ub[m], 8, lmask, lmagic);
          const uint32_t h0 = lift4(ua[m], 4, lmask, lmagic), h1 = lift4(ub[m], 4, lmask, lmagic),
                         h2 = lift4(ua[m], 12, lmask, lmagic), h3 = lift4(ub[m], 12, lmask, lmagic);
#pragma unroll
          for (int mt = 0; mt < MT; ++mt) {
            mma_bf16_16816(acc[mt], l0, l1, l2, l3, xb[mt][2 * m], xb[mt][2 * m + 1]);
            mma_bf16_16816(hi[mt], h0, h1, h2, h3, xb[mt][2 * m], xb[mt][2 * m + 1]);
          }
        }
#pragma unroll
        for (int mt = 0; mt < MT; ++mt)
#pragma unroll
          for (int q = 0; q < 4; ++q) acc[mt][q] = fmaf(16.f, hi[mt][q], acc[mt][q]);
        off = 128.f + 2048.f;
      }
      const float sc0 = j & 1 ? bf16hi(scp[0][j >> 1]) : bf16lo(scp[0][j >> 1]);
      const float sc1 = j & 1 ? bf16hi(scp[1][j >> 1]) : bf16lo(scp[1][j >> 1]);
      const float bi0 = j & 1 ? bf16hi(bip[0][j >> 1]) : bf16lo(bip[0][j >> 1]);
      const float bi1 = j & 1 ? bf16hi(bip[1][j >> 1]) : bf16lo(bip[1][j >> 1]);
      const float bb0 = fmaf(-off, sc0, bi0), bb1 = fmaf(-off, sc1, bi1);
#pragma unroll
      for (int mt = 0; mt < MT; ++mt) {
        y[mt][0] = fmaf(bb0, sx[mt].x, fmaf(sc0, acc[mt][0], y[mt][0]));
        y[mt][2] = fmaf(bb1, sx[mt].x, fmaf(sc1, acc[mt][2], y[mt][2]));
        if (!ONECOL) {
          y[mt][1] = fmaf(bb0, sx[mt].y, fmaf(sc0, acc[mt][1], y[mt][1]));
          y[mt][3] = fmaf(bb1, sx[mt].y, fmaf(sc1, acc[mt][3], y[mt][3]));
        }
      }
    }
    __syncwarp();
    c_soff = c_soff == (S - 1) * G::STAGE_BYTES ? 0u : c_soff + G::STAGE_BYTES;
    xoff += G::GPC * 128;
    if (--c_left > 0) continue;

    // ---- end of a tile ------------------------------------------------------------------------------
    c_left = nch;
    xoff = 0;
    const int slot = c_iter & (NS - 1);
    const uint32_t use = (uint32_t)c_iter / NS;  // how many times this slot was used before
    const int tile = c_tile;
    c_tile += tstride;
    ++c_iter;
    if (!finisher) {
      if (use > 0) mbar_wait(empty_bar + slot, (use - 1) & 1);
      const uint32_t dst = exch_s + ((((uint32_t)warp * NS + slot) * MT) * 32 + lane) * 16;
#pragma unroll
      for (int mt = 0; mt < MT; ++mt) {
        asm volatile("st.shared.v4.f32 [%0], {%1,%2,%3,%4};" ::"r"(dst + mt * 512), "f"(y[mt][0]), "f"(y[mt][1]),
                     "f"(y[mt][2]), "f"(y[mt][3]) : "memory");
        y[mt][0] = y[mt][1] = y[mt][2] = y[mt][3] = 0.f;
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(full_bar + slot);
      continue;
    }
    float v[MT][4];
    if (ks_w > 1) {
      // try_wait.parity only tells the barrier's current phase from the one before it, so this warp must not get a
      // phase ahead of full_bar: once the slot's previous tile has been released (empty phase use - 1), its full phase
      // is complete as well and full_bar is in phase `use` or later
      if (use > 0) mbar_wait(empty_bar + slot, (use - 1) & 1);
      mbar_wait(full_bar + slot, use & 1);
#pragma unroll
      for (int mt = 0; mt < MT; ++mt) { v[mt][0] = v[mt][1] = v[mt][2] = v[mt][3] = 0.f; }
      for (int w = 0; w < ks_w; ++w) {  // warp order, own partial in its place: deterministic
        if (w == wt) {
#pragma unroll
          for (int mt = 0; mt < MT; ++mt)
#pragma unroll
            for (int q = 0; q < 4; ++q) v[mt][q] += y[mt][q];
        } else {
          const uint32_t src = exch_s + ((((uint32_t)(warp - wt + w) * NS + slot) * MT) * 32 + lane) * 16;
#pragma unroll
          for (int mt = 0; mt < MT; ++mt) {
            const uint4 p = lds128(src + mt * 512);
            v[mt][0] += __uint_as_float(p.x); v[mt][1] += __uint_as_float(p.y);
            v[mt][2] += __uint_as_float(p.z); v[mt][3] += __uint_as_float(p.w);
          }
        }
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(empty_bar + slot);
    } else {
#pragma unroll
      for (int mt = 0; mt < MT; ++mt)
#pragma unroll
        for (int q = 0; q < 4; ++q) v[mt][q] = y[mt][q];
    }
#pragma unroll
    for (int mt = 0; mt < MT; ++mt) { y[mt][0] = y[mt][1] = y[mt][2] = y[mt][3] = 0.f; }

    if (a.ksplit > 1) {  // cross-CTA split-K: the last CTA to arrive sums the slices in slice order (one warp does it all)
      float4* part = reinterpret_cast<float4*>(a.ws_part) + ((size_t)tile * a.ksplit) * MT * 32 + lane;
#pragma unroll
      for (int mt = 0; mt < MT; ++mt)
        part[((size_t)ksg * MT + mt) * 32] = make_float4(v[mt][0], v[mt][1], v[mt][2], v[mt][3]);
      __threadfence();
      __syncwarp();
      unsigned last = 0;
      if (lane == 0) last = atomicAdd(&a.ws_cnt[tile], 1u) == (unsigned)a.ksplit - 1;
      last = __shfl_sync(0xffffffffu, last, 0);
      if (!last) continue;
      __threadfence();
#pragma unroll
      for (int mt = 0; mt < MT; ++mt) { v[mt][0] = v[mt][1] = v[mt][2] = v[mt][3] = 0.f; }
      for (int s = 0; s < a.ksplit; ++s) {
#pragma unroll
        for (int mt = 0; mt < MT; ++mt) {
          const uint4 p = ldcg128(part + ((size_t)s * MT + mt) * 32);
          v[mt][0] += __uint_as_float(p.x); v[mt][1] += __uint_as_float(p.y);
          v[mt][2] += __uint_as_float(p.z); v[mt][3] += __uint_as_float(p.w);
        }
      }
      if (lane == 0) a.ws_cnt[tile] = 0;
    }

    // ---- epilogue from registers: lane (g, t) holds rows g, g + 8 x activation rows 2t, 2t + 1 (per mt) -------
#pragma unroll
    for (int mt = 0; mt < MT; ++mt) {
#pragma unroll
      for (int c = 0; c < (ONECOL ? 1 : 2); ++c) {
        const int m = mt * 8 + 2 * t + c;
        if (m >= a.M) continue;
        if (a.epi == EPI_SILUMUL) {  // row g = gate, row g + 8 = up of the same output column
          const int n = nbase + g;
          if (n < a.seg[0].N) {
            const float gt = bfround(v[mt][c]), up = bfround(v[mt][2 + c]);
            const float sg = bfround(1.f / (1.f + expf(-gt)));  // nn.silu = x * sigmoid(x), each op rounded
            a.out[(size_t)m * a.out_ld + n] = f2bf(bfround(gt * sg) * up);
          }
        } else {
#pragma unroll
          for (int h = 0; h < 2; ++h) {
            const int n = nbase + g + 8 * h;
            if (n < a.seg[sidx].N) {
              float res = bfround(v[mt][2 * h + c]);
              if (a.epi == EPI_RESADD) res = bf16lo(res_raw[mt][2 * h + c]) + res;
              a.out[(size_t)m * a.out_ld + a.seg[sidx].out_off + n] = f2bf(res);
            }
          }
        }
      }
    }
  }
  cp_async_wait<0>();
