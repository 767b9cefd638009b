// Launchers for decode attention and the (bring-up) causal prefill attention.
namespace pie {

template <int D>
static cudaError_t launch_decode_attn_d(const DecAttnArgs& a0, int B, int n_rep, cudaStream_t st) {
  DecAttnArgs a = a0;
  const int R = n_rep % 4 == 0 ? 4 : (n_rep % 2 == 0 ? 2 : 1);
  a.rg = n_rep / R;
  const dim3 grid(a.nsplit, a.Hkv * a.rg, B);
  count_launch();
  if (R == 4) decode_attn_kernel<D, 4><<<grid, 128, 0, st>>>(a);
  else if (R == 2) decode_attn_kernel<D, 2><<<grid, 128, 0, st>>>(a);
  else decode_attn_kernel<D, 1><<<grid, 128, 0, st>>>(a);
  return cudaGetLastError();
}

cudaError_t launch_decode_attn(uint16_t* out, const uint16_t* q, const uint16_t* kc, const uint16_t* vc,
                               const KvLayout& kv, const int32_t* pos, int B, int Hq, int Hkv, int D,
                               float scale, int nsplit, float* ws, cudaStream_t st) {
  if (B == 0) return cudaSuccess;
  // tensor-core kernel (pie_fa.cu) unless PIE_DECODE_ATTN=simt; it takes the n_rep query heads of a kv head as MMA rows
  static const bool simt = [] { const char* e = getenv("PIE_DECODE_ATTN"); return e && e[0] == 's'; }();
  if (!simt && Hq / Hkv <= 16 && nsplit <= 32)
    return launch_decode_attn_tc(out, q, kc, vc, kv, pos, B, Hq, Hkv, D, scale, nsplit, ws, st);
  DecAttnArgs a;
  a.out = out; a.q = q; a.kc = kc; a.vc = vc; a.kv = kv; a.pos = pos;
  a.Hq = Hq; a.Hkv = Hkv; a.nsplit = nsplit; a.rg = 1;
  a.scale_log2e = scale * 1.4426950408889634f;
  const size_t cnt_bytes = (((size_t)B * Hq * sizeof(unsigned)) + 255) & ~(size_t)255;
  a.ws_cnt = reinterpret_cast<unsigned int*>(ws);
  a.ws_part = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(ws) + cnt_bytes);
  if (D == 128) return launch_decode_attn_d<128>(a, B, Hq / Hkv, st);
  return launch_decode_attn_d<64>(a, B, Hq / Hkv, st);
}

// ---- a10: SDPA for L > 1 with the causal mask of create_attention_mask (PIE/models/base.py:18-53) ---------
// Bring-up kernel (one CTA per query row, fp32, online softmax).  The additive -1e9 mask makes masked
// probabilities underflow to exactly 0 in fp32, so masked keys are skipped.  q [B,Hq,L,D] -> out [B,L,Hq*D].
template <int D>
__global__ void __launch_bounds__(128) prefill_attn_kernel(uint16_t* out, const uint16_t* q, const uint16_t* kc,
                                                           const uint16_t* vc, const KvLayout kv, int Hq, int Hkv,
                                                           int L, int offset, float scale_log2e) {
  constexpr int E = D / 32;
  __shared__ float sm[4][D + 2];
  const int l = blockIdx.x, h = blockIdx.y, b = blockIdx.z;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int kvh = h / (Hq / Hkv);
  const uint16_t* qr = q + (((size_t)b * Hq + h) * L + l) * D;
  float qf[E], acc[E];
#pragma unroll
  for (int e = 0; e < E; ++e) { qf[e] = bf16lo((uint32_t)qr[lane * E + e]) * scale_log2e; acc[e] = 0.f; }
  float m = -INFINITY, ls = 0.f;
  const int T = offset + l + 1;
  for (int t = warp; t < T; t += 4) {
    const size_t off = kv_addr_a(kv, b, kvh, t);
    float s = 0.f, vf[E];
#pragma unroll
    for (int e = 0; e < E; ++e) {
      s = fmaf(qf[e], bf16lo((uint32_t)kc[off + lane * E + e]), s);
      vf[e] = bf16lo((uint32_t)vc[off + lane * E + e]);
    }
    s = warp_sum(s);
    const float mn = fmaxf(m, s), corr = exp2f(m - mn), p = exp2f(s - mn);
    ls = ls * corr + p;
#pragma unroll
    for (int e = 0; e < E; ++e) acc[e] = fmaf(p, vf[e], acc[e] * corr);
    m = mn;
  }
#pragma unroll
  for (int e = 0; e < E; ++e) sm[warp][lane * E + e] = acc[e];
  if (lane == 0) { sm[warp][D] = m; sm[warp][D + 1] = ls; }
  __syncthreads();
  for (int d = threadIdx.x; d < D; d += 128) {
    float mn = -INFINITY;
    for (int w = 0; w < 4; ++w) mn = fmaxf(mn, sm[w][D]);
    float lt = 0.f, at = 0.f;
    for (int w = 0; w < 4; ++w) {
      const float cw = sm[w][D] == -INFINITY ? 0.f : exp2f(sm[w][D] - mn);
      lt += sm[w][D + 1] * cw;
      at += sm[w][d] * cw;
    }
    out[((size_t)b * L + l) * Hq * D + (size_t)h * D + d] = f2bf(at / lt);
  }
}

cudaError_t launch_prefill_attn(uint16_t* out, const uint16_t* q, const uint16_t* kc, const uint16_t* vc,
                                const KvLayout& kv, int B, int Hq, int Hkv, int L, int D, int offset,
                                float scale, cudaStream_t st) {
  if (B == 0 || L == 0) return cudaSuccess;
  static const bool naive = [] { const char* e = getenv("PIE_PREFILL_ATTN_NAIVE"); return e && e[0] == '1'; }();
  if (!naive) return launch_prefill_attn_tc(out, q, kc, vc, kv, B, Hq, Hkv, L, D, offset, scale, st);
  const float sl = scale * 1.4426950408889634f;
  const dim3 grid(L, Hq, B);
  count_launch();
  if (D == 128) prefill_attn_kernel<128><<<grid, 128, 0, st>>>(out, q, kc, vc, kv, Hq, Hkv, L, offset, sl);
  else prefill_attn_kernel<64><<<grid, 128, 0, st>>>(out, q, kc, vc, kv, Hq, Hkv, L, offset, sl);
  return cudaGetLastError();
}

}  // namespace pie
