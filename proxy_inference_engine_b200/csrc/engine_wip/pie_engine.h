// Private descriptors shared by the host API (pie_api.cu) and the decode engine kernel
// (pie_engine.cu).  Nothing here is part of the public C ABI (include/pie_b200.h).
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace pie {

// ---- geometry of the weight-streaming ring -------------------------------------------------
constexpr int kTileRows = 16;         // one mma.sync m16 tile of weight rows
constexpr int kStageGroups = 8;       // quantisation groups (64 elements) per stage along K
constexpr int kGroup = 64;            // the only group size on the fast path
constexpr int kStageK = kStageGroups * kGroup;  // 512 elements of K per stage
constexpr int kConsumerWarps = 15;           // 15 + 1 producer = 512 threads -> 128 registers each
constexpr int kThreads = 32 * (kConsumerWarps + 1);  // warp 0 = TMA producer
constexpr int kMaxBatch = 8;          // mma.sync n8
constexpr int kMaxStages = 64;
constexpr int kTileSlots = 64;        // >= stages in flight (see DESIGN.md "tile slots")
constexpr int kMaxSeg = 3;

__host__ __device__ constexpr int stage_w_bytes(int bits) { return kTileRows * kStageK * bits / 8; }
__host__ __device__ constexpr int stage_sb_bytes() { return kTileRows * kStageGroups * 2; }
__host__ __device__ constexpr int stage_boxes(int bits) { return stage_w_bytes(bits) / (kTileRows * 128); }

enum XMode : int {
  X_PLAIN = 0,   // x = bf16 vector(s) from global
  X_NORM = 1,    // x = rms_norm(h) * w_norm
  X_EMBED = 2,   // h = dequantize(embed[token]) ; x = rms_norm(h) * w_norm   (first phase of a step)
};
enum Epilogue : int {
  EPI_STORE = 0,    // out[m, n] = bf16(y)
  EPI_RESADD = 1,   // h[m, n] = bf16(h[m, n] + bf16(y))
  EPI_SILUMUL = 2,  // dual tile (rows 0-7 gate, 8-15 up): out[m, n] = bf16(bf16(g*bf16(sigmoid g)) * u)
  EPI_LOGITS = 3,   // out[m, n] = bf16(y) (if out) and arg-max over n of bf16(y)
};

struct GemvSeg {
  int map_w, map_s, map_b;  // indices into the tensor-map array
  int N;                    // rows of this matrix
  int tile0;                // first tile index of this segment inside the phase
  int out_off;              // column offset of this segment in the output row
};

struct GemvDesc {
  GemvSeg seg[kMaxSeg];
  int nseg;
  int dual;        // 1: seg[0] = gate, seg[1] = up, each tile = 8 gate rows + 8 up rows
  int K;           // reduction length (multiple of 64)
  int chunks;      // ceil(K / kStageK)
  int ntiles;      // tiles over all segments
  int xmode;
  int epi;
  int out_ld;      // leading dimension (elements) of the output / residual rows
  float eps;
  const void* x_in;     // X_PLAIN: bf16 [M, K]; X_NORM: bf16 h [M, K]; X_EMBED: unused
  const void* norm_w;   // bf16 [K]
  void* out;            // bf16 [M, out_ld]  (EPI_RESADD: the residual stream h, updated in place)
};

struct AttnDesc {
  void* kcache;   // bf16
  void* vcache;
};

// One decode launch.  Everything the kernel needs, passed by value (< 4 KB).
struct EngineParams {
  // model
  int n_layers, hidden, ffn, n_heads, n_kv, head_dim, vocab, bits;
  float attn_scale;
  const CUtensorMap* maps;     // global memory
  const GemvDesc* gemv;        // [n_layers * 4 + 1]  (qkv, o, gate/up, down) per layer + head
  const AttnDesc* attn;        // [n_layers]
  const float* rope_freqs;     // [head_dim / 2], wavelength-like (theta = pos / freqs[i])
  // embedding (row gather + dequantize), MLX layout
  const uint32_t* embed_w; const uint16_t* embed_s; const uint16_t* embed_b;
  // kv geometry (elements)
  long long kv_batch_stride, kv_head_stride, kv_tok_stride;
  const int* page_table;  // optional [batch, pt_stride]
  int pt_stride, page_size;
  // io
  int batch, nsteps;
  int* tokens;            // [batch] in/out
  int* positions;         // [batch] in/out
  int* out_tokens;        // [nsteps, batch] or null
  void* logits;           // bf16 [batch, vocab] or null
  // activations (global scratch, bf16)
  void* h;                // [batch, hidden]
  void* qkv;              // [batch, (n_heads + 2 n_kv) * head_dim]
  void* attn_out;         // [batch, n_heads * head_dim]
  void* act;              // [batch, ffn]
  // sync / reduction scratch
  unsigned int* grid_bar;       // [1]   zeroed before launch
  unsigned int* tile_cnt;       // [max tiles]   self-cleaning
  float* tile_part;             // [max tiles][2][16*kMaxBatch]
  unsigned int* attn_cnt;       // [batch * n_kv] self-cleaning
  float* attn_part;             // [batch * n_heads][nsplit][head_dim + 2]
  unsigned long long* argmax;   // [2][batch]  zeroed before launch
  int* error_flag;
  // config
  int nsplit;            // attention splits per (batch, kv head)
  int nstages;           // ring depth
  int tile_aligned;      // 1: CTA ranges aligned to whole tiles (no cross-CTA fix-up)
  // shared-memory carve-up (bytes from the dynamic smem base, 1024-aligned)
  int smem_ring_w, smem_ring_s, smem_ring_b, smem_x, smem_sumx, smem_slots, smem_attn, smem_total;
  int x_ld;              // bf16 elements per x row in smem (K or 2K for 8-bit)
};

// Standalone quantized matvec (pie_qmv): one GEMV phase, maps passed as grid constants.
struct QmvParams {
  GemvDesc d;
  int batch, bits, nstages, tile_aligned;
  unsigned int* tile_cnt;
  float* tile_part;
  int* error_flag;
  int smem_ring_w, smem_ring_s, smem_ring_b, smem_x, smem_sumx, smem_slots, smem_total;
  int x_ld;
};

}  // namespace pie
