/* TEST INFRASTRUCTURE ONLY.
 *
 * CPU restatement (plain C) of the reference algorithms on the quantized-Llama decode path.
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
 * load this library; the product path never does and fails loudly without its CUDA library.
 *
 * Parity status: PINNED.  orc_quantize / orc_qmm_t_seq are checked bit-for-bit against the
 * reference's own code compiled from /root/reference (oracle/ref_harness.cpp -> oracle/_ref),
 * and every routine is checked against the MLX property tests it restates
 * (tests/test_oracle_*.py; MLX/python/tests/test_quantized.py, test_fast.py, test_fast_sdpa.py).
 *
 * Citations: MLX = /root/reference/src/pie_core/external/mlx, PIE = /root/reference/src/proxy_inference_engine.
 *
 * Conventions.  Every float tensor crosses this ABI as float32.  A "bf16 tensor" is a float32
 * array whose values are bf16-representable; "mode B" entry points take `round_out` and round
 * their result to bf16 (RNE, NaN -> 0x7FC0, MLX/mlx/types/bf16.h:41-57) exactly where the MLX
 * graph has an op boundary.  Compile with -ffp-contract=off so no FMA contraction changes the
 * reference's operation order.
 */
#include <math.h>
#include <stddef.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

/* ---- bf16 (MLX/mlx/types/bf16.h:41-72) ---------------------------------------------------- */
static inline uint16_t f32_to_bf16_bits(float x) {
  if (isnan(x)) return 0x7FC0;
  union { float f; uint32_t u; } in;
  in.f = x;
  in.u += ((in.u >> 16) & 1u) + 0x7FFFu;
  return (uint16_t)(in.u >> 16);
}
static inline float bf16_bits_to_f32(uint16_t h) {
  union { float f; uint32_t u; } out;
  out.u = ((uint32_t)h) << 16;
  return out.f;
}
static inline float rbf(float x) { return bf16_bits_to_f32(f32_to_bf16_bits(x)); }

void orc_f32_to_bf16(const float* x, uint16_t* out, size_t n) {
  for (size_t i = 0; i < n; ++i) out[i] = f32_to_bf16_bits(x[i]);
}
void orc_bf16_to_f32(const uint16_t* x, float* out, size_t n) {
  for (size_t i = 0; i < n; ++i) out[i] = bf16_bits_to_f32(x[i]);
}
void orc_round_bf16(float* x, size_t n) {
  for (size_t i = 0; i < n; ++i) x[i] = rbf(x[i]);
}
int orc_max_threads(void) {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}
void orc_set_threads(int n) {
#ifdef _OPENMP
  if (n > 0) omp_set_num_threads(n);
#else
  (void)n;
#endif
}

/* ---- a1: affine group quantisation (MLX/mlx/backend/cpu/quantized.cpp:603-661) ------------- *
 * w: [w_size] row-major, groups of `group_size` consecutive elements.  bits in {2,4,8}.
 * store_bf16 != 0 restates T = bfloat16_t: the inputs are bf16 values and scale/bias are rounded
 * to bf16 when stored (the packing itself uses the unrounded fp32 scale/bias, as the reference). */
void orc_quantize(const float* w, uint32_t* out, float* scales, float* biases, int bits,
                  int group_size, size_t w_size, int store_bf16) {
  const float n_bins = (float)((1 << bits) - 1);
  const float eps = 1e-7f;
  const int el_per_int = 32 / bits;
  const int int_per_group = group_size / el_per_int;
  const size_t n_groups = w_size / (size_t)group_size;
#pragma omp parallel for schedule(static)
  for (size_t i = 0; i < n_groups; ++i) {
    const size_t w_idx = i * (size_t)group_size;
    float w_min = INFINITY, w_max = -INFINITY;
    for (int j = 0; j < group_size; ++j) {
      float v = w[w_idx + j];
      w_max = (w_max < v) ? v : w_max; /* std::max(w_max, v) */
      w_min = (v < w_min) ? v : w_min; /* std::min(w_min, v) */
    }
    const int mask = fabsf(w_min) > fabsf(w_max);
    float scale = (w_max - w_min) / n_bins;
    scale = (scale < eps) ? eps : scale; /* std::max(scale, eps) */
    scale = mask ? scale : -scale;
    const float edge = mask ? w_min : w_max;
    const float q0 = rintf(edge / scale);
    float bias = 0.0f;
    if (q0 != 0.0f) {
      scale = edge / q0;
      bias = edge;
    }
    const size_t out_idx = i * (size_t)int_per_group;
    for (int j = 0; j < int_per_group; ++j) {
      uint32_t out_el = 0;
      for (int k = 0; k < el_per_int; ++k) {
        float w_el = w[w_idx + (size_t)j * el_per_int + k];
        w_el = rintf((w_el - bias) / scale);
        w_el = fminf(fmaxf(w_el, 0.0f), n_bins);
        out_el |= ((uint32_t)w_el) << (k * bits);
      }
      out[out_idx + j] = out_el;
    }
    scales[i] = store_bf16 ? rbf(scale) : scale;
    biases[i] = store_bf16 ? rbf(bias) : bias;
  }
}

/* ---- unpack: element k of a row = bits [bits*(k%epi), +bits) of word k/epi, LSB first -------- *
 * (the `<< (k * bits)` at quantized.cpp:648; the byte walk of _qmm_t :129-143 reads the same). */
static inline uint32_t unpack_q(const uint32_t* wrow, int k, int bits) {
  const int epi = 32 / bits;
  return (wrow[k / epi] >> ((k % epi) * bits)) & ((1u << bits) - 1u);
}
void orc_unpack(const uint32_t* w, uint8_t* q, size_t rows, int K, int bits) {
  const size_t wrow = (size_t)K * bits / 32;
#pragma omp parallel for schedule(static)
  for (size_t r = 0; r < rows; ++r)
    for (int k = 0; k < K; ++k) q[r * K + k] = (uint8_t)unpack_q(w + r * wrow, k, bits);
}

/* ---- a4: dequantize (MLX/mlx/fast.cpp:972-1017 fallback) --------------------------------------- *
 * w_hat = T(T(q) * scale) + bias, every op in T.  t_bf16 != 0 -> T = bf16 (two roundings);
 * else T = float.  Optional row gather (QuantizedEmbedding.__call__, quantized.py:100-107). */
void orc_dequantize_rows(const uint32_t* w, const float* scales, const float* biases,
                         const int32_t* ids, size_t n_ids, int K, int bits, int group_size,
                         int t_bf16, float* out) {
  const size_t wrow = (size_t)K * bits / 32;
  const size_t grow = (size_t)K / group_size;
#pragma omp parallel for schedule(static)
  for (size_t i = 0; i < n_ids; ++i) {
    const size_t r = ids ? (size_t)ids[i] : i;
    for (int k = 0; k < K; ++k) {
      const float q = (float)unpack_q(w + r * wrow, k, bits);
      const float s = scales[r * grow + k / group_size];
      const float b = biases[r * grow + k / group_size];
      float v = q * s;
      if (t_bf16) v = rbf(v);
      v = v + b;
      if (t_bf16) v = rbf(v);
      out[i * (size_t)K + k] = v;
    }
  }
}

/* ---- a2: quantized matmul, transpose=True ------------------------------------------------------ *
 * (1) orc_qmm_t_seq: the x86 scalar loop `_qmm_t<T>` (quantized.cpp:96-153):
 *        sum += x[k] * (scale * T(q) + bias), strictly sequential in k, every op rounded to T.
 *     t_bf16 = 0 -> T=float (bit-exact vs the compiled reference), 1 -> T=bfloat16_t. */
void orc_qmm_t_seq(float* y, const float* x, const uint32_t* w, const float* scales,
                   const float* biases, int M, int N, int K, int bits, int group_size, int t_bf16) {
  const size_t wrow = (size_t)K * bits / 32;
  const size_t grow = (size_t)K / group_size;
#pragma omp parallel for collapse(2) schedule(static)
  for (int m = 0; m < M; ++m) {
    for (int n = 0; n < N; ++n) {
      const float* xr = x + (size_t)m * K;
      const uint32_t* wr = w + (size_t)n * wrow;
      float sum = 0.0f;
      for (int k = 0; k < K; ++k) {
        const float s = scales[n * grow + k / group_size];
        const float b = biases[n * grow + k / group_size];
        const float q = (float)unpack_q(wr, k, bits);
        if (t_bf16) {
          float t = rbf(s * q);
          t = rbf(t + b);
          t = rbf(xr[k] * t);
          sum = rbf(sum + t);
        } else {
          float t = s * q;
          t = t + b;
          t = xr[k] * t;
          sum = sum + t;
        }
      }
      y[(size_t)m * N + n] = sum;
    }
  }
}

/* (2) orc_qmm_t: the fp32-inside semantics of `_qmm_t_simd` (quantized.cpp:180-225, S = 8):
 *        wf = float(q) * scale + bias ; acc[lane] += x * wf ; y = T(sum(acc))
 *     which is also what Metal qmv_fast computes (fp32 accumulate, one rounding to T at the store,
 *     MLX/mlx/backend/metal/kernels/quantized.h:647-688).  round_out -> bf16 rounding of y (mode B).
 *     8 lanes; lane j accumulates elements k = j (mod 8); lanes are summed 0..7 left to right. */
void orc_qmm_t(float* y, const float* x, const uint32_t* w, const float* scales,
               const float* biases, int M, int N, int K, int bits, int group_size, int round_out) {
  const size_t wrow = (size_t)K * bits / 32;
  const size_t grow = (size_t)K / group_size;
  const int epi = 32 / bits;
  const uint32_t qmask = (1u << bits) - 1u;
  /* thread wake-ups cost milliseconds on some VMs: stay serial for decode-sized calls */
  const int par = ((double)M * (double)N * (double)K) > 2.0e8;
#pragma omp parallel if (par)
  {
    float* wf = (float*)malloc(sizeof(float) * (size_t)K);
#pragma omp for schedule(static)
    for (int n = 0; n < N; ++n) {
      const uint32_t* wr = w + (size_t)n * wrow;
      for (int k = 0; k < K; k += epi) {
        const uint32_t word = wr[k / epi];
        const float s = scales[n * grow + k / group_size];
        const float b = biases[n * grow + k / group_size];
        for (int p = 0; p < epi; ++p) {
          float t = (float)(int32_t)((word >> (p * bits)) & qmask) * s;
          wf[k + p] = t + b;
        }
      }
      for (int m = 0; m < M; ++m) {
        const float* xr = x + (size_t)m * K;
        float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        for (int k = 0; k < K; k += 8) {
          for (int j = 0; j < 8; ++j) {
            float t = xr[k + j] * wf[k + j];
            acc[j] = acc[j] + t;
          }
        }
        float sum = acc[0];
        for (int j = 1; j < 8; ++j) sum = sum + acc[j];
        y[(size_t)m * N + n] = round_out ? rbf(sum) : sum;
      }
    }
    free(wf);
  }
}

/* ---- a6: RMSNorm (MLX/mlx/fast.cpp:93-111 fallback; metal rms_norm.metal:29-71) ---------------- *
 * x32 * rsqrt(mean(x32^2) + eps) -> astype(T) -> * weight (in T). */
void orc_rms_norm(const float* x, const float* weight, float* out, size_t rows, int dims, float eps,
                  int t_bf16) {
#pragma omp parallel for schedule(static)
  for (size_t r = 0; r < rows; ++r) {
    const float* xr = x + r * (size_t)dims;
    float* o = out + r * (size_t)dims;
    double ss = 0.0; /* reduction order is unspecified in MLX; double keeps it order-free */
    for (int i = 0; i < dims; ++i) ss += (double)xr[i] * (double)xr[i];
    const float mean = (float)(ss / (double)dims);
    const float inv = 1.0f / sqrtf(mean + eps);
    for (int i = 0; i < dims; ++i) {
      float v = xr[i] * inv;
      if (t_bf16) v = rbf(v);
      if (weight) {
        v = v * weight[i];
        if (t_bf16) v = rbf(v);
      }
      o[i] = v;
    }
  }
}

/* ---- a7: RoPE with explicit freqs (MLX/mlx/fast.cpp:397-471; PIE/models/llama/utils.py:42-50) --- *
 * x: [n_batch, L, dims_total] (the op flattens leading dims; for [B,H,L,D] pass n_batch = B*H).
 * theta[p,i] = ((p + offset) * scale) * (1 / freqs[i]); non-traditional pairs (i, i + dims/2);
 * traditional pairs (2i, 2i+1).  All fp32; round_out rounds the result to bf16 (the Metal kernel
 * keeps fp32 inside and casts once, rope.metal:16-46).                                            */
void orc_rope(const float* x, float* out, size_t n_batch, int L, int dims_total, int dims,
              int traditional, float scale, int offset, const float* freqs, float base,
              int round_out) {
  const int half = dims / 2;
  float* inv = (float*)malloc(sizeof(float) * (size_t)half);
  for (int i = 0; i < half; ++i) {
    if (freqs) inv[i] = 1.0f / freqs[i];
    else inv[i] = expf((float)(-i) * (logf(base) / (float)half));
  }
#pragma omp parallel for collapse(2) schedule(static)
  for (size_t b = 0; b < n_batch; ++b) {
    for (int p = 0; p < L; ++p) {
      const float* xr = x + (b * (size_t)L + p) * (size_t)dims_total;
      float* o = out + (b * (size_t)L + p) * (size_t)dims_total;
      const float pos = ((float)p + (float)offset) * scale;
      for (int i = 0; i < half; ++i) {
        const float theta = pos * inv[i];
        const float c = cosf(theta), s = sinf(theta);
        const int i1 = traditional ? 2 * i : i;
        const int i2 = traditional ? 2 * i + 1 : i + half;
        const float x1 = xr[i1], x2 = xr[i2];
        float a = x1 * c, bq = x2 * s, cq = x1 * s, d = x2 * c;
        float o1 = a - bq, o2 = cq + d;
        o[i1] = round_out ? rbf(o1) : o1;
        o[i2] = round_out ? rbf(o2) : o2;
      }
      for (int i = dims; i < dims_total; ++i) o[i] = xr[i];
    }
  }
  free(inv);
}

/* Llama3RoPE.__init__ (PIE/models/llama/utils.py:22-39), float32 arithmetic as mx does. */
void orc_llama3_freqs(float* freqs, int dims, float base, float factor, float low_freq_factor,
                      float high_freq_factor, float max_embedding_length,
                      float global_embedding_length) {
  const float low_freq_wavelen = global_embedding_length / low_freq_factor;
  const float high_freq_wavelen = global_embedding_length / high_freq_factor;
  const float two_pi = 2.0f * 3.14159265358979323846f;
  for (int i = 0; i < dims / 2; ++i) {
    const float e = (float)(2 * i) / (float)dims;
    float f = powf(base, e);
    const float wavelen = two_pi * f;
    const float f_scaled = (wavelen > low_freq_wavelen) ? f * factor : f;
    const int is_medium = (wavelen > high_freq_wavelen) && (wavelen < low_freq_wavelen);
    const float smooth = (max_embedding_length / wavelen - low_freq_factor) /
                         (high_freq_factor - low_freq_factor);
    const float smooth_f = f_scaled / ((1.0f - smooth) / factor + smooth);
    freqs[i] = is_medium ? smooth_f : f_scaled;
  }
}

/* ---- a9/a10: SDPA (MLX/mlx/fast.cpp:678-727 fallback; PIE/models/base.py:18-53 mask) ----------- *
 * q [B,Hq,L,D], k,v [B,Hkv,T,D] (row stride of k/v given as kv_stride_t tokens so a cache buffer
 * with capacity > T can be passed directly), out [B,Hq,L,D].
 * scores = (scale*q) . k^T (+ mask) ; softmax in fp32 (precise=true) ; out = P . v.
 * mask: NULL, or additive [L,T] float (the -1e9 causal mask of create_causal_mask).
 * causal_offset >= 0 builds that mask on the fly: masked where (offset + row) < col.
 * Everything fp32 inside (Metal sdpa_vector semantics); round_out rounds the output to bf16. */
void orc_sdpa(const float* q, const float* k, const float* v, float* out, int B, int Hq, int Hkv,
              int L, int T, int D, size_t kv_cap, float scale, const float* mask,
              int causal_offset, int round_out) {
  const int n_rep = Hq / Hkv;
#pragma omp parallel
  {
    float* sc = (float*)malloc(sizeof(float) * (size_t)T);
    float* acc = (float*)malloc(sizeof(float) * (size_t)D);
#pragma omp for collapse(3) schedule(static)
    for (int b = 0; b < B; ++b)
      for (int h = 0; h < Hq; ++h)
        for (int l = 0; l < L; ++l) {
          const float* qr = q + (((size_t)b * Hq + h) * L + l) * D;
          const float* kb = k + ((size_t)b * Hkv + h / n_rep) * kv_cap * D;
          const float* vb = v + ((size_t)b * Hkv + h / n_rep) * kv_cap * D;
          float mx = -INFINITY;
          for (int t = 0; t < T; ++t) {
            float s = 0.0f;
            for (int d = 0; d < D; ++d) s += (scale * qr[d]) * kb[(size_t)t * D + d];
            if (mask) s += mask[(size_t)l * T + t];
            if (causal_offset >= 0 && (causal_offset + l) < t) s += -1e9f;
            sc[t] = s;
            mx = (s > mx) ? s : mx;
          }
          float denom = 0.0f;
          for (int t = 0; t < T; ++t) {
            sc[t] = expf(sc[t] - mx);
            denom += sc[t];
          }
          for (int d = 0; d < D; ++d) acc[d] = 0.0f;
          for (int t = 0; t < T; ++t) {
            const float p = sc[t];
            for (int d = 0; d < D; ++d) acc[d] += p * vb[(size_t)t * D + d];
          }
          float* o = out + (((size_t)b * Hq + h) * L + l) * D;
          const float invd = 1.0f / denom;
          for (int d = 0; d < D; ++d) {
            float r = acc[d] * invd;
            o[d] = round_out ? rbf(r) : r;
          }
        }
    free(sc);
    free(acc);
  }
}

/* ---- a11: MLP glue: silu(g) * u with T-rounded op boundaries ----------------------------------- *
 * nn.silu = x * sigmoid(x) (MLX/python/mlx/nn/layers/activations.py:125-132), then `* up`
 * (PIE/models/llama/language.py:126-127).  t_bf16: sigmoid, product and final product each
 * rounded to bf16.                                                                               */
void orc_silu_mul(const float* g, const float* u, float* out, size_t n, int t_bf16) {
#pragma omp parallel for schedule(static)
  for (size_t i = 0; i < n; ++i) {
    float sg = 1.0f / (1.0f + expf(-g[i]));
    if (t_bf16) sg = rbf(sg);
    float a = g[i] * sg;
    if (t_bf16) a = rbf(a);
    float r = a * u[i];
    out[i] = t_bf16 ? rbf(r) : r;
  }
}
void orc_add(const float* a, const float* b, float* out, size_t n, int t_bf16) {
#pragma omp parallel for schedule(static)
  for (size_t i = 0; i < n; ++i) {
    float r = a[i] + b[i];
    out[i] = t_bf16 ? rbf(r) : r;
  }
}
