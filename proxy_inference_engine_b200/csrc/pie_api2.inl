// Second half of the C ABI (included by pie_api.cu): qmm, norm, rope, KV cache, attention, glue, sampler.
static pie::KvLayout to_kv(const pie_kv_layout* k) {
  pie::KvLayout o;
  o.batch_stride = k->batch_stride; o.head_stride = k->head_stride; o.tok_stride = k->tok_stride;
  o.page_table = k->page_table; o.pt_stride = k->pt_stride; o.page_size = k->page_size;
  return o;
}
static int check_kv(const pie_kv_layout* k, int D) {
  if (!k) return PIE_ENULL;
  if (k->tok_stride < D || k->tok_stride % 8 || k->head_stride % 8 || k->batch_stride % 8) return PIE_EINVAL_SHAPE;
  if (k->page_table && (k->page_size <= 0 || k->pt_stride <= 0)) return PIE_EINVAL_SHAPE;
  return PIE_OK;
}

extern "C" {

// M >= 64: the tcgen05 / TMEM tile kernel (pie_qmm.cu), weights dequantised once to bf16 into shared memory.
// Smaller M, or shapes off its grid (K % 128, unaligned pointers): the decode GEMV in passes of <= 32 rows.
int pie_qmm(void* y, const void* x, const uint32_t* wq, const void* scales, const void* biases, int M, int N,
            int K, int group_size, int bits, void* workspace, size_t workspace_bytes, void* stream) {
  if (M == 0 || N == 0) return PIE_OK;
  if (!y || !x || !wq || !scales || !biases) return PIE_ENULL;
  if (M < 0 || N < 0 || K <= 0) return PIE_EINVAL_SHAPE;
  const bool tc = M >= 64 && quant_supported(group_size, bits) && K % 128 == 0 && aligned16(x) && aligned16(wq) &&
                  aligned16(y) && !(reinterpret_cast<uintptr_t>(scales) & 3) &&
                  !(reinterpret_cast<uintptr_t>(biases) & 3) && env_int("PIE_QMM_TC", 1) != 0;
  if (tc)
    return cuda_status(launch_qmm_tc((uint16_t*)y, (const uint16_t*)x, wq, (const uint16_t*)scales,
                                     (const uint16_t*)biases, M, N, K, bits, (cudaStream_t)stream));
  return pie_qmv(y, x, wq, scales, biases, M, N, K, group_size, bits, workspace, workspace_bytes, stream);
}

int pie_rms_norm(void* out, const void* x, const void* weight, long long rows, int dims, float eps, void* stream) {
  if (rows == 0) return PIE_OK;
  if (!out || !x) return PIE_ENULL;
  if (rows < 0 || dims <= 0) return PIE_EINVAL_SHAPE;
  return cuda_status(launch_rms_norm((uint16_t*)out, (const uint16_t*)x, (const uint16_t*)weight, rows, dims, eps,
                                     (cudaStream_t)stream));
}

int pie_rope(void* out, const void* x, long long n_batch, int L, int D, int dims, int traditional, float base,
             float scale, int offset, const float* freqs, void* stream) {
  if (n_batch == 0 || L == 0) return PIE_OK;
  if (!out || !x) return PIE_ENULL;
  if (n_batch < 0 || L < 0 || D <= 0 || dims <= 0 || dims > D || dims % 2) return PIE_EINVAL_SHAPE;
  if (!freqs && !(base > 0.f)) return PIE_EINVAL_SHAPE;
  return cuda_status(launch_rope((uint16_t*)out, (const uint16_t*)x, n_batch, L, D, dims, traditional, base, scale,
                                 offset, freqs, (cudaStream_t)stream));
}

int pie_rope_kv_append(void* q_out, const void* qkv, void* kcache, void* vcache, const pie_kv_layout* kv,
                       const int32_t* positions, const float* freqs, int B, int Hq, int Hkv, int D, void* stream) {
  if (B == 0) return PIE_OK;
  if (!q_out || !qkv || !kcache || !vcache || !positions || !freqs) return PIE_ENULL;
  if (B < 0 || Hq <= 0 || Hkv <= 0 || Hq % Hkv || (D != 64 && D != 128)) return PIE_EINVAL_SHAPE;
  if (int s = check_kv(kv, D)) return s;
  return cuda_status(launch_rope_kv_append((uint16_t*)q_out, (const uint16_t*)qkv, (uint16_t*)kcache,
                                           (uint16_t*)vcache, to_kv(kv), positions, freqs, B, Hq, Hkv, D,
                                           (cudaStream_t)stream));
}

int pie_kv_write(void* kcache, void* vcache, const pie_kv_layout* kv, const void* k, const void* v, int B, int Hkv,
                 int L, int D, int offset, void* stream) {
  if (B == 0 || L == 0) return PIE_OK;
  if (!kcache || !vcache || !k || !v) return PIE_ENULL;
  if (B < 0 || Hkv <= 0 || L < 0 || D <= 0 || offset < 0) return PIE_EINVAL_SHAPE;
  if (int s = check_kv(kv, D)) return s;
  return cuda_status(launch_kv_write((uint16_t*)kcache, (uint16_t*)vcache, to_kv(kv), (const uint16_t*)k,
                                     (const uint16_t*)v, B, Hkv, L, D, offset, (cudaStream_t)stream));
}

size_t pie_decode_attn_workspace_bytes(int B, int Hq, int D, int nsplit) {
  if (B <= 0 || Hq <= 0 || D <= 0 || nsplit <= 0) return 0;
  const size_t cnt = (((size_t)B * Hq * 4) + 255) & ~(size_t)255;
  return cnt + (size_t)B * Hq * nsplit * (D + 2) * sizeof(float);
}

int pie_decode_attn_pick_splits(int B, int Hq, int Hkv, int max_context) {
  if (B <= 0 || Hq <= 0 || Hkv <= 0) return 1;
  const int n_rep = Hq / Hkv;
  int ctas, want;
  if (n_rep <= 16) {  // tensor-core kernel: one CTA per (sequence, kv head, split), 2 resident per SM (96 KB of stages)
    ctas = B * Hkv;
    want = 148 * 2;
  } else {
    const int rg = n_rep % 4 == 0 ? n_rep / 4 : (n_rep % 2 == 0 ? n_rep / 2 : n_rep);
    ctas = B * Hkv * rg;
    want = 148 * 4;
  }
  // ONE wave of resident CTAs, never a wave and a fraction: measured on B200 (8B shapes, paged, us per layer at
  // 1 / 2 / 3 / 4 splits): bs32 ctx2048 (256 CTAs per split) 45 / 59 / 62 / 66; bs16 ctx4096 (128) 70 / 50 / 65 / 57;
  // bs8 ctx2048 (64), 2 / 4 / 5 / 8 splits: 27 / 23 / 29 / 24
  int ns = want / ctas;
  const int cap = (max_context + 63) / 64;  // >= 64 tokens (one stage) per split
  if (ns > cap) ns = cap;
  if (ns > 32) ns = 32;  // kFaMaxSplits (pie_fa.cu)
  return ns < 1 ? 1 : ns;
}

int pie_decode_attn(void* out, const void* q, const void* kcache, const void* vcache, const pie_kv_layout* kv,
                    const int32_t* positions, int B, int Hq, int Hkv, int D, float scale, int nsplit,
                    void* workspace, size_t workspace_bytes, void* stream) {
  if (B == 0) return PIE_OK;
  if (!out || !q || !kcache || !vcache || !positions) return PIE_ENULL;
  if (B < 0 || Hq <= 0 || Hkv <= 0 || Hq % Hkv) return PIE_EINVAL_SHAPE;
  if (D != 64 && D != 128) return PIE_EUNSUPPORTED;
  if (int s = check_kv(kv, D)) return s;
  if (!aligned16(q) || !aligned16(kcache) || !aligned16(vcache)) return PIE_EALIGN;
  if (nsplit <= 0) nsplit = 1;
  if (nsplit > 1 && (!workspace || workspace_bytes < pie_decode_attn_workspace_bytes(B, Hq, D, nsplit)))
    return PIE_EWORKSPACE;
  return cuda_status(launch_decode_attn((uint16_t*)out, (const uint16_t*)q, (const uint16_t*)kcache,
                                        (const uint16_t*)vcache, to_kv(kv), positions, B, Hq, Hkv, D, scale, nsplit,
                                        (float*)workspace, (cudaStream_t)stream));
}

int pie_prefill_attn(void* out, const void* q, const void* kcache, const void* vcache, const pie_kv_layout* kv, int B,
                     int Hq, int Hkv, int L, int D, int offset, float scale, void* stream) {
  if (B == 0 || L == 0) return PIE_OK;
  if (!out || !q || !kcache || !vcache) return PIE_ENULL;
  if (B < 0 || Hq <= 0 || Hkv <= 0 || Hq % Hkv || L < 0 || offset < 0) return PIE_EINVAL_SHAPE;
  if (D != 64 && D != 128) return PIE_EUNSUPPORTED;
  if (int s = check_kv(kv, D)) return s;
  return cuda_status(launch_prefill_attn((uint16_t*)out, (const uint16_t*)q, (const uint16_t*)kcache,
                                         (const uint16_t*)vcache, to_kv(kv), B, Hq, Hkv, L, D, offset, scale,
                                         (cudaStream_t)stream));
}

int pie_silu_mul(void* out, const void* gate, const void* up, long long n, void* stream) {
  if (n == 0) return PIE_OK;
  if (!out || !gate || !up) return PIE_ENULL;
  return cuda_status(launch_silu_mul((uint16_t*)out, (const uint16_t*)gate, (const uint16_t*)up, n, (cudaStream_t)stream));
}
int pie_add(void* out, const void* a, const void* b, long long n, void* stream) {
  if (n == 0) return PIE_OK;
  if (!out || !a || !b) return PIE_ENULL;
  return cuda_status(launch_add((uint16_t*)out, (const uint16_t*)a, (const uint16_t*)b, n, (cudaStream_t)stream));
}
int pie_argmax(int32_t* out_idx, const void* logits, int rows, int vocab, int32_t* positions_to_advance,
               int32_t* seq_out, int seq_ld, void* stream) {
  if (rows == 0) return PIE_OK;
  if (!out_idx || !logits) return PIE_ENULL;
  if (rows < 0 || vocab <= 0) return PIE_EINVAL_SHAPE;
  return cuda_status(launch_argmax(out_idx, (const uint16_t*)logits, rows, vocab, positions_to_advance, seq_out, seq_ld,
                                   (cudaStream_t)stream));
}

}  // extern "C"
