/* pie_b200.h -- C ABI of libpie_b200.so: the B200 (sm_100a) implementation of the quantized Llama
 * decode path of TheProxyCompany/proxy-inference-engine (PIE).
 *
 * Every entry point replaces one operator the reference reaches through MLX; the citation beside each
 * declaration is the reference interface it stands in for.  Shorthands: PIE = src/proxy_inference_engine,
 * MLX = src/pie_core/external/mlx (both relative to the reference root).
 *
 * Conventions
 *   - plain pointers and sizes only; every pointer is a DEVICE pointer unless the name ends in _host;
 *   - activations, scales, biases and norm weights are bf16 (uint16 bit patterns); packed weights are
 *     uint32 in MLX's affine layout: element k of a row lives in bits [bits*(k % (32/bits)), +bits) of
 *     word k / (32/bits), LSB first (MLX/mlx/backend/cpu/quantized.cpp:648); one (scale, bias) pair per
 *     `group_size` consecutive elements of a row;
 *   - all calls are asynchronous on `stream` (a cudaStream_t passed as void*), allocate nothing, keep no
 *     state, never synchronise: they can be captured in CUDA graphs.  Workspaces are caller-owned;
 *   - the return value is a pie_status; PIE_ECUDA means a CUDA runtime error (pie_last_cuda_error()).
 *   - supported set: bits in {4, 8}, group_size 64.  Anything else -> PIE_EUNSUPPORTED (the Python
 *     wrapper raises ValueError before the call, like MLX/mlx/ops.cpp:82-134 and fast.cpp:839-866).
 */
#ifndef PIE_B200_H_
#define PIE_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum pie_status {
  PIE_OK = 0,
  PIE_EINVAL_SHAPE = 1,  /* a dimension breaks the operator's contract                    */
  PIE_EINVAL_DTYPE = 2,  /* dtype code not understood                                     */
  PIE_EUNSUPPORTED = 3,  /* valid for MLX, not built here (bits/group_size/head_dim ...)  */
  PIE_ECUDA = 4,         /* CUDA error; see pie_last_cuda_error()                          */
  PIE_ENULL = 5,         /* a required pointer is NULL                                    */
  PIE_EALIGN = 6,        /* a pointer is not 16-byte aligned                               */
  PIE_EWORKSPACE = 7,    /* caller-owned workspace missing or smaller than *_workspace_bytes() */
  PIE_ESTATIC_LIMIT = 8  /* the request needs more than the library's static scratch (PIE_STATIC_SCRATCH_BYTES) */
} pie_status;

enum { PIE_DTYPE_BF16 = 0, PIE_DTYPE_F32 = 1 };

/* epilogues of pie_qmv_fused */
enum {
  PIE_EPI_STORE = 0,   /* out[m, off + n] = bf16(y)                                               */
  PIE_EPI_RESADD = 1,  /* out[m, n] = bf16(residual[m, n] + bf16(y))     (language.py:150-154)    */
  PIE_EPI_SILUMUL = 2  /* out[m, n] = silu(bf16(y_gate)) * bf16(y_up)    (language.py:126-127)    */
};

int pie_abi_version(void);
const char* pie_status_string(int status);
const char* pie_last_cuda_error(void);
/* number of kernels this library has launched since load (the bench's gpu_launches claim) */
unsigned long long pie_launch_count(void);

/* ---- a1  mx.quantize  (MLX/mlx/ops.cpp:4023-4029 -> MLX/mlx/backend/cpu/quantized.cpp:603-661) -------
 * w [rows, K] (dtype) -> wq [rows, K*bits/32] u32, scales/biases [rows, K/group_size] (dtype). Bit-exact. */
int pie_quantize(const void* w, int dtype, uint32_t* wq, void* scales, void* biases, long long rows, int K,
                 int group_size, int bits, void* stream);

/* ---- a4  mx.dequantize, with the row gather of QuantizedEmbedding.__call__ ---------------------------
 * (MLX/python/mlx/nn/layers/quantized.py:100-107; MLX/mlx/fast.cpp:972-1017)
 * out[i, :] = bf16(bf16(q[r] * scales[r]) + biases[r]),  r = ids ? ids[i] : i.   out bf16 [n_rows, K].  */
int pie_dequantize(void* out, const uint32_t* wq, const void* scales, const void* biases, const int32_t* ids,
                   long long n_rows, long long table_rows, int K, int group_size, int bits, void* stream);

/* ---- a2/a5  mx.quantized_matmul(x, w, scales, biases, transpose=True) for decode-sized M -------------
 * (MLX/mlx/ops.cpp:3984-4021; QuantizedLinear.__call__ MLX/python/mlx/nn/layers/quantized.py:204-216;
 *  replaces Metal qmv_fast, MLX/mlx/backend/metal/kernels/quantized.h:623-689)
 * y[m, n] = bf16( sum_k x[m, k] * (scales[n, k/gs] * q[n, k] + biases[n, k/gs]) ), fp32 accumulate.
 * x bf16 [M, K], y bf16 [M, N].  M > 32 runs as passes of 32 rows.
 * Split-K workspace: pie_qmv_workspace_bytes() bytes, caller-owned, ZEROED ONCE at allocation (the
 * kernel leaves its counters at zero), never shared by two streams at a time.  `dual` = 1 for the SwiGLU
 * form of pie_qmv_fused (n_total = N of one of the two matrices), else n_total = sum of the segments' N. */
size_t pie_qmv_workspace_bytes(int M, long long n_total, int K, int bits, int dual, int ksplit_hint);
int pie_qmv(void* y, const void* x, const uint32_t* wq, const void* scales, const void* biases, int M, int N,
            int K, int group_size, int bits, void* workspace, size_t workspace_bytes, void* stream);

/* Fused form used by the decode step: up to 3 weight matrices sharing x (q/k/v), optional RMSNorm of x on
 * the way in (a6: norm_weight != NULL -> x <- rms_norm(x, norm_weight, eps)) and residual-add / SwiGLU on
 * the way out (a11).  For PIE_EPI_SILUMUL pass exactly two segments (gate, up) of equal N; the output row
 * is then N wide.  residual (ld = out_ld) may alias out.  x_ld = elements between rows of x.
 * pdl != 0 launches with programmatic stream serialization: the kernel prefetches its weights while the
 * previous kernel on the stream drains, and reads x only after that kernel has completed.               */
typedef struct pie_qmv_seg {
  const uint32_t* wq;
  const void* scales;
  const void* biases;
  int N;
} pie_qmv_seg;
int pie_qmv_fused(void* out, int out_ld, const void* x, int x_ld, int M, int K, const pie_qmv_seg* segs_host,
                  int nseg, int group_size, int bits, const void* norm_weight, float eps, int epilogue,
                  const void* residual, int ksplit_hint, void* workspace, size_t workspace_bytes, int pdl,
                  void* stream);

/* ---- a3  mx.quantized_matmul for prefill-sized M (replaces Metal qmm_t, kernels/quantized.h:954-1078) -
 * Same contraction and workspace contract as pie_qmv.  x bf16 [M, K] row-major, y bf16 [M, N].           */
int pie_qmm(void* y, const void* x, const uint32_t* wq, const void* scales, const void* biases, int M, int N,
            int K, int group_size, int bits, void* workspace, size_t workspace_bytes, void* stream);

/* ---- a6  mx.fast.rms_norm (MLX/mlx/fast.cpp:56-123; kernels/rms_norm.metal:13-81) -------------------- */
int pie_rms_norm(void* out, const void* x, const void* weight, long long rows, int dims, float eps,
                 void* stream);

/* ---- a7  mx.fast.rope (MLX/mlx/fast.cpp:353-511) as called by Llama3RoPE (PIE/models/llama/utils.py:42-50)
 * x, out bf16 [n_batch, L, row_stride >= dims] (n_batch = B*H for [B,H,L,D]); rotates the first `dims`
 * features; freqs f32 [dims/2] (wavelength-like: theta = pos/freqs[i]) or NULL -> base.                 */
int pie_rope(void* out, const void* x, long long n_batch, int L, int D, int dims, int traditional, float base,
             float scale, int offset, const float* freqs, void* stream);

/* ---- a7+a8 fused for decode (L == 1): rotate q and k at position pos[b], append k,v to the cache --------
 * (Attention.__call__ PIE/models/llama/language.py:84-98; ReusableKVCache.update_and_fetch reusable.py:96-142)
 * qkv bf16 [B, (Hq + 2 Hkv) * D] as produced by the fused q/k/v projection; q_out bf16 [B, Hq, D].
 * KV geometry in elements: addr(b, h, t) = base + b*batch_stride + h*head_stride + t*tok_stride, or with
 * page_table (int32 [B, pt_stride], page_size tokens per page):
 *   addr = (page_table[b, t / page_size] * page_size + t % page_size) * tok_stride + h * head_stride.   */
typedef struct pie_kv_layout {
  long long batch_stride, head_stride, tok_stride;
  const int32_t* page_table;
  int pt_stride, page_size;
} pie_kv_layout;
int pie_rope_kv_append(void* q_out, const void* qkv, void* kcache, void* vcache, const pie_kv_layout* kv_host,
                       const int32_t* positions, const float* freqs, int B, int Hq, int Hkv, int D,
                       void* stream);

/* ---- a8  cache.update_and_fetch for L >= 1 (prefill): copy k,v [B, Hkv, L, D] into the cache at offset - */
int pie_kv_write(void* kcache, void* vcache, const pie_kv_layout* kv_host, const void* k, const void* v, int B,
                 int Hkv, int L, int D, int offset, void* stream);

/* ---- a9  mx.fast.scaled_dot_product_attention, L == 1, mask None (PIE/models/base.py:92-113) ----------
 * split-K flash-decode (the scheme of Metal sdpa_vector_2pass, kernels/sdpa_vector.h:168-392).
 * q bf16 [B, Hq, D]; context length of row b = positions[b] + 1; out bf16 [B, Hq * D].
 * workspace (nsplit > 1): pie_decode_attn_workspace_bytes(B, Hq, D, nsplit) bytes, caller-owned,
 * ZEROED ONCE at allocation (counters are left at zero).  nsplit is fixed per launch so the call can be
 * captured in a CUDA graph; splits beyond the current context are empty and cost nothing.              */
size_t pie_decode_attn_workspace_bytes(int B, int Hq, int D, int nsplit);
int pie_decode_attn_pick_splits(int B, int Hq, int Hkv, int max_context);
int pie_decode_attn(void* out, const void* q, const void* kcache, const void* vcache,
                    const pie_kv_layout* kv_host, const int32_t* positions, int B, int Hq, int Hkv, int D,
                    float scale, int nsplit, void* workspace, size_t workspace_bytes, void* stream);

/* ---- pie_decode_attn without a caller-owned workspace and without any allocation -------------------------------
 * NO COUNTERPART IN THE REFERENCE: added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  Same arguments and results as
 * pie_decode_attn; the split counters and per-split partials live in a file-scope device array of
 * PIE_STATIC_SCRATCH_BYTES (4 MiB), held by every GPU that loads the library until exit.  A request that needs more
 * (pie_decode_attn_workspace_bytes(B, Hq, D, nsplit) > PIE_STATIC_SCRATCH_BYTES) returns PIE_ESTATIC_LIMIT and
 * launches nothing.  All calls share the array: the counters are cleared on `stream` before each launch, and calls that
 * overlap on the GPU (different streams without ordering) race -- serialise them or use pie_decode_attn.           */
#define PIE_STATIC_SCRATCH_BYTES (4u << 20)
int pie_decode_attn_static(void* out, const void* q, const void* kcache, const void* vcache,
                           const pie_kv_layout* kv_host, const int32_t* positions, int B, int Hq, int Hkv, int D,
                           float scale, int nsplit, void* stream);

/* ---- a7+a8+a9 in one launch: rotate q and k at positions[b], append k,v to the cache, flash-decode ---------------
 * (Attention.__call__ PIE/models/llama/language.py:84-104 for L == 1).  qkv bf16 [B, (Hq + 2 Hkv) D] un-rotated, as
 * the fused q/k/v projection writes it; out bf16 [B, Hq D].  One CTA of 16 warps per (row, kv head, pair of query
 * heads, context split); the appended token is read from qkv, older ones from the cache.  Workspace and nsplit as
 * for pie_decode_attn (pie_decode_step_pick_splits sizes nsplit for this kernel).
 * rope_table (optional, pie_rope_table): sin / cos of the angles of positions[], computed once per step for all
 * layers; NULL = every launch computes them from freqs.  Position-only work (angles, the cache rows of older tokens)
 * runs before the launch's dependency wait: positions[] and the cache rows below positions[b] must not be written by
 * the kernel launched just before this one.                                                                        */
int pie_rope_table(float* table /* f32 [B][2][half_dims] */, const int32_t* positions, const float* freqs, int B,
                   int half_dims, void* stream);
int pie_decode_attn_fused(void* out, const void* qkv, void* kcache, void* vcache, const pie_kv_layout* kv_host,
                          const int32_t* positions, const float* freqs, const float* rope_table, int B, int Hq,
                          int Hkv, int D, float scale, int nsplit, void* workspace, size_t workspace_bytes, int pdl,
                          void* stream);

/* ---- a10 SDPA for L > 1 with the additive causal mask of create_attention_mask (PIE/models/base.py:18-53)
 * q bf16 [B, Hq, L, D]; k,v read from the cache (T = offset + L tokens); out bf16 [B, L, Hq * D].       */
int pie_prefill_attn(void* out, const void* q, const void* kcache, const void* vcache,
                     const pie_kv_layout* kv_host, int B, int Hq, int Hkv, int L, int D, int offset, float scale,
                     void* stream);

/* ---- (f3) quantised KV cache: QuantizedKVCache.update_and_fetch (PIE/cache/kv_cache/quantized.py:38-108) and
 * quantized_scaled_dot_product_attention for L == 1 (PIE/models/base.py:56-89), 8 bits, groups of 64 along head_dim.
 * The cache is the reference's state tuple per K and V: quants u32 [slots, D / 4] (byte e of a row = element e, the
 * packing of mx.quantize), scales and biases bf16 [slots, D / 64]; slot of (row b, kv head h, token t) =
 * (b * Hkv + h) * capacity + t, or with a page table (page_table[b * pt_stride + t / page_size] * Hkv + h) * page_size
 * + t % page_size.  The struct lives on the host and is read at the call.
 *   pie_kvq_append : pie_rope_kv_append with a quantised cache -- q heads rotated into q_out, k heads rotated, rounded
 *                    to bf16 and quantised into the cache at positions[b], v heads quantised into the cache;
 *                    bit-exact with mx.quantize of the bf16 rows.
 *   pie_kvq_write  : pie_kv_write with a quantised cache (prefill; k already rotated).
 *   pie_decode_attn_q8 : pie_decode_attn over the quantised cache; K, V and the scores are never materialised
 *                    (q.k = s * sum(q kq) + b * sum(q) per group), fp32 throughout, one bf16 rounding at the store.
 *                    Workspace and its zero-at-allocation contract as pie_decode_attn; nsplit from
 *                    pie_decode_attn_q8_pick_splits.                                                              */
typedef struct pie_kvq_cache {
  uint32_t* k_quant; void* k_scales; void* k_biases;
  uint32_t* v_quant; void* v_scales; void* v_biases;
  long long capacity;          /* tokens per (row, kv head), even; unused with a page table (page_size: a power of two) */
  const int32_t* page_table;   /* device int32 [B, pt_stride] or NULL */
  int pt_stride, page_size;
  int group_size, bits;        /* 64, 8 */
} pie_kvq_cache;
int pie_kvq_append(void* q_out, const void* qkv, const pie_kvq_cache* cache_host, const int32_t* positions,
                   const float* freqs, int B, int Hq, int Hkv, int D, void* stream);
int pie_kvq_write(const pie_kvq_cache* cache_host, const void* k, const void* v, int B, int Hkv, int L, int D,
                  int offset, void* stream);
int pie_decode_attn_q8_pick_splits(int B, int Hq, int Hkv, int max_context);
int pie_decode_attn_q8(void* out, const void* q, const pie_kvq_cache* cache_host, const int32_t* positions, int B,
                       int Hq, int Hkv, int D, float scale, int nsplit, void* workspace, size_t workspace_bytes,
                       void* stream);

/* ---- a11 glue: nn.silu(g) * u and h + r, bf16 with a rounding at each MLX op boundary ------------------ */
int pie_silu_mul(void* out, const void* gate, const void* up, long long n, void* stream);
int pie_add(void* out, const void* a, const void* b, long long n, void* stream);

/* ---- greedy sampler (temp == 0): first arg-max over the vocabulary (PIE/samplers/__init__.py make_sampler).
 * positions_to_advance (optional, int32 [rows]) is incremented by one per row: the device-side
 * `cache.offset += 1` of the decode loop (PIE/engine/inference_engine.py:282-297), so that a captured
 * step can be replayed without host involvement.  seq_out (optional, int32 [rows, seq_ld]) is the
 * device-side token list of the engine: the sampled token is also stored at seq_out[row, new position].  */
int pie_argmax(int32_t* out_idx, const void* logits, int rows, int vocab, int32_t* positions_to_advance,
               int32_t* seq_out, int seq_ld, void* stream);


/* ---- (f1) the rest of the sampler, on the device: logprobs = f32(logits) - logsumexp (PIE/engine/inference_engine.py:
 * 268-269) and make_sampler's temperature branches (PIE/samplers/__init__.py:11-45: top_p.py, min_p.py, top_k.py,
 * categorical.py; precedence top_p, then min_p, then top_k, else plain categorical), one launch per step.
 * temp must be > 0 (temp == 0 is pie_argmax).  The categorical draw is the arg-max of logprobs / temp + Gumbel noise over
 * the kept tokens.  Noise: Philox4x32-10 keyed by rng_state (DEVICE memory, 32 bytes: u64 seed, u64 draw counter, 16
 * bytes of scratch -- zero at start; the kernel advances the counter, so a captured graph draws fresh numbers on every
 * replay), or `uniforms` (f32 [rows, vocab] in (0, 1)) when given.  logprobs_out (optional) f32 [rows, vocab].
 * positions_to_advance / seq_out as for pie_argmax.  logits bf16 [rows, ld >= vocab]. */
int pie_sample(int32_t* tokens, const void* logits, int rows, int vocab, int ld, float temp, float top_p, float min_p,
               int min_tokens_to_keep, int top_k, void* rng_state, const float* uniforms, float* logprobs_out,
               int32_t* positions_to_advance, int32_t* seq_out, int seq_ld, void* stream);

/* ---- (e) tensor-parallel exchange steps, one-shot over NVLink peer memory ----------------------------------------
 * Replace the collectives of MLX's sharded linears at decode sizes: `all_sum` after a row-split matrix
 * (QuantizedShardedToAllLinear, MLX/python/mlx/nn/layers/distributed.py:459-566) and the gather of vocabulary-split
 * logits.  Every rank owns one SYMMETRIC buffer of pie_tp_buffer_bytes() bytes, zeroed once, mapped into every peer
 * (CUDA IPC / fabric handles, e.g. torch.distributed._symmetric_memory); base[r] is rank r's buffer as mapped in THIS
 * process.  `epoch`: 8 local device bytes holding {1, 0} initially; the kernels advance it themselves, so the calls can
 * be captured in a CUDA graph and replayed.  All ranks must issue the same sequence of exchanges.
 *   pie_tp_allreduce: out[rows, width] = bf16(residual + bf16(sum over ranks of partial)), fp32 sum in rank order
 *                     (identical bits on every rank); rows * width <= max_elems and <= 131072.
 *   pie_tp_allgather: every rank's slice [rows, width] lands at columns [rank * width, (rank + 1) * width) of the
 *                     [rows, world * width] bf16 matrix at byte offset pie_tp_gather_offset() of EVERY rank's buffer. */
typedef struct pie_tp_peers {
  void* base[8];
  int rank, world, max_elems;
  void* epoch;
} pie_tp_peers;
/* The all-reduce FUSED into the matrices on either side of it (batch 1): mode 1 = producer -- the o / down projection
 * stores its partial row straight into slot `rank` of every peer's buffer instead of `out`; mode 2 = consumer -- the
 * next matrix's x load waits for the `world` flags and takes x = x + bf16(sum of the slots) (x = the residual stream),
 * stores that new residual stream to h_out (which must not alias x) and goes on with its RMS norm.  No kernel of its own,
 * no extra pass over the row.  A buffer used this way must not be used by pie_tp_allreduce / pie_tp_allgather (separate
 * exchange numbers): give the fused path its own buffer and epoch (16 bytes: {1, 0, 0, 0}). */
typedef struct pie_tp_fused {
  pie_tp_peers peers;
  int mode;      /* 0 none, 1 producer, 2 consumer */
  void* h_out;   /* mode 2: bf16 [K], receives the new residual stream */
} pie_tp_fused;
int pie_qmv_fused_tp(void* out, int out_ld, const void* x, int x_ld, int M, int K, const pie_qmv_seg* segs_host,
                     int nseg, int group_size, int bits, const void* norm_weight, float eps, int epilogue,
                     const void* residual, int ksplit_hint, void* workspace, size_t workspace_bytes, int pdl,
                     const pie_tp_fused* tp_host, void* stream);
size_t pie_tp_buffer_bytes(int world, int max_elems, size_t gather_bytes);
size_t pie_tp_gather_offset(int world, int max_elems);
int pie_tp_allreduce(void* out, const void* partial, const void* residual, int rows, int width,
                     const pie_tp_peers* peers_host, void* stream);
int pie_tp_allgather(const void* slice, int rows, int width, const pie_tp_peers* peers_host, void* stream);


/* ---- a12  one whole decode step (L == 1) in ONE cooperative launch ---------------------------------------
 * Model.__call__ for a single new token per sequence (PIE/models/llama/language.py:199-210 -> LlamaModel.__call__
 * :170-187 -> TransformerBlock.__call__ :144-154 -> Attention.__call__ :75-108 / MLP.__call__ :126-127) followed by
 * the temp == 0 sampler and `cache.offset += 1` of generate_step (PIE/engine/inference_engine.py:282-297):
 *   h = embed[tokens]; per layer: qkv = qmv(rms_norm(h)); rope(q,k) at positions[b]; append k,v; flash-decode;
 *   h += qmv(attn; o); act = silu(gate)*up of qmv(rms_norm(h)); h += qmv(act; down);  logits = qmv(rms_norm(h); head);
 *   tokens[b] = argmax(logits[b]); positions[b] += 1; seq[b, positions[b]] = tokens[b].
 * A persistent kernel (one CTA per SM) walks the step's matrices in order: each warp's TMA weight ring goes on with the
 * next matrix's units across the grid barriers that separate the matrices, so HBM never idles between them.  The GEMV
 * arithmetic is that of the TMA-fed per-operator kernel (fp16 staging); attention, embedding and sampler as the
 * per-operator entry points above.  Built for B == 1 and 4-bit weights with K % 512 == 0 (PIE_EUNSUPPORTED otherwise:
 * use the per-operator calls).
 * Protocol: pie_decode_step_plan() (host only, no CUDA call) turns the description into `plan_host`
 * (pie_decode_step_plan_bytes(n_layers) bytes); the caller copies it to the device once; pie_decode_step() launches.
 * `sync_ws`: 256 device bytes, zeroed once at allocation.  attn_ws: pie_decode_step_attn_ws_bytes(), zeroed once. */
typedef struct pie_step_layer {
  pie_qmv_seg q, k, v, o, gate, up, down;
  const void* input_norm;      /* input_layernorm.weight            bf16 [hidden] */
  const void* post_norm;       /* post_attention_layernorm.weight   bf16 [hidden] */
  void* kcache;
  void* vcache;
  pie_kv_layout kv;
} pie_step_layer;
typedef struct pie_step_model {
  int n_layers, hidden, ffn, Hq, Hkv, D, vocab, bits, group_size;
  float rms_eps, attn_scale;
  const float* rope_freqs;     /* f32 [D/2], Llama3RoPE._freqs */
  pie_qmv_seg embed;           /* embed_tokens (N = vocab) */
  pie_qmv_seg head;            /* lm_head, or embed_tokens again when tied */
  const void* final_norm;
} pie_step_model;
typedef struct pie_step_io {
  int B;
  int32_t* tokens;             /* [B] in: token to consume; out: token sampled */
  int32_t* positions;          /* [B] in: cache slot of the consumed token; out: +1 */
  int32_t* seq; int seq_ld;    /* optional token list [B, seq_ld] */
  void *h, *qkv, *attn, *act, *logits;   /* bf16 scratch: [B,hidden] [B,(Hq+2Hkv)D] [B,Hq*D] [B,ffn] [B,vocab] */
  void* attn_ws; size_t attn_ws_bytes; int nsplit;  /* nsplit from pie_decode_step_pick_splits */
  void* qmv_ws; size_t qmv_ws_bytes;     /* split-K workspace of pie_qmv_fused (max over the step's matrices) */
  unsigned long long* trace;             /* optional profiling aid: u64 [2 + 3 * (4 n_layers + 1)] receiving CTA 0's
                                            %globaltimer at kernel start / end, and per matrix: input ready / x staged /
                                            matrix done */
} pie_step_io;
size_t pie_decode_step_plan_bytes(int n_layers);
int pie_decode_step_pick_splits(int B, int Hq, int Hkv, int max_context);
size_t pie_decode_step_attn_ws_bytes(int B, int Hq, int D, int nsplit);
int pie_decode_step_plan(const pie_step_model* model, const pie_step_layer* layers, const pie_step_io* io,
                         void* plan_host, size_t plan_bytes);
int pie_decode_step(const void* plan_dev, const void* plan_host, size_t plan_bytes, void* sync_ws, void* stream);

/* ---- f2  paged-KV page allocator: fixed pool of 64-token pages, reference counts, lock-free LIFO free list ------
 * (PageAllocator, CORE/include/engine/page_allocator.hpp:17-72, CORE/src/engine/page_allocator.cpp:8-158; KVPage's
 *  counters, CORE/include/engine/page.hpp:84-123; tests CORE-level tests/cpp/test_page_allocator.cpp).
 * Host-side: a page id indexes the pool tensors of cache.PagedKVCache in HBM and goes into the page tables the attention
 * kernels read (pie_kv_layout.page_table).  create returns NULL for num_pages == 0, num_heads <= 0 or head_dim <= 0
 * (std::invalid_argument in the reference); allocate returns -1 when the pool is exhausted (std::nullopt); free / add_ref
 * return PIE_EINVAL_SHAPE for an id outside the pool (std::out_of_range).  All calls are thread-safe and lock-free;
 * pools of up to 2^24 - 2 pages (the free count and the list head share one 64-bit word).                          */
#define PIE_PAGE_TOKENS 64
typedef struct pie_page_allocator pie_page_allocator;
pie_page_allocator* pie_page_allocator_create(size_t num_pages, int num_heads, int head_dim);
void pie_page_allocator_destroy(pie_page_allocator* a);
long long pie_page_allocate(pie_page_allocator* a);                 /* allocate_page(): ref count 1, num_tokens 0 */
int pie_page_free(pie_page_allocator* a, uint32_t page_id);         /* free_page(): back on the free list at count 0 */
int pie_page_add_ref(pie_page_allocator* a, uint32_t page_id);      /* add_ref(): sharing (prefix reuse) */
long long pie_page_ref_count(const pie_page_allocator* a, uint32_t page_id);   /* -1: id outside the pool */
long long pie_page_num_tokens(const pie_page_allocator* a, uint32_t page_id);
int pie_page_set_num_tokens(pie_page_allocator* a, uint32_t page_id, uint32_t n);
size_t pie_page_allocator_size(const pie_page_allocator* a);
size_t pie_page_allocator_num_free(const pie_page_allocator* a);    /* get_num_free_pages() */
/* whole sequences at a time (no reference counterpart): n pages, all or nothing (PIE_EWORKSPACE when the pool is short) */
int pie_page_allocate_n(pie_page_allocator* a, int32_t* out_ids, int n);
int pie_page_free_n(pie_page_allocator* a, const int32_t* ids, int n);
int pie_page_add_ref_n(pie_page_allocator* a, const int32_t* ids, int n);
/* the reference's own scenario benchmark and contention test, run natively on this allocator:
 * BM_PageAllocator_ReferenceCountingScenario (benchmarks/cpp/core/page_allocator_benchmark.cpp:191-300) -> ms per
 * iteration; HighContentionPushPopStress (tests/cpp/test_page_allocator.cpp:278-330) -> pages lost or handed out twice */
double pie_page_allocator_bench_refcount(pie_page_allocator* a, int iterations);
long long pie_page_allocator_stress(pie_page_allocator* a, int threads, int ops_per_thread);

/* Profiling aid (no reference counterpart): after this call every TMA GEMV and fused-attention launch writes up to
 * sixteen %globaltimer stamps per CTA (entry, first loads issued, dependency wait returned, activations staged, first
 * weights landed, main loop done, CTA reduction, exit, ...) into the next 32 KB slot of `buf` (device memory).  NULL
 * switches it off.  A launch takes its slot when it is ENQUEUED, so a captured CUDA graph keeps writing the slots it was
 * captured with.  tools/timeline_step.py prints the timeline of a decode step from it. */
int pie_debug_stamps(void* buf, size_t bytes);

#ifdef __cplusplus
}
#endif
#endif /* PIE_B200_H_ */
