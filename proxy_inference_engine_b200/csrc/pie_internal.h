// Private launch descriptors shared by the C-ABI layer (pie_api.cu) and the kernels.
// Nothing here is part of the public ABI (include/pie_b200.h).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/pie_b200.h"

namespace pie {

constexpr int kGroup = 64;      // the only group size on the fast path
constexpr int kTileRows = 16;   // one mma.sync m16 tile of weight rows
constexpr int kMaxBatch = 32;   // activation rows per qmv pass: up to 4 mma.sync n8 column tiles
constexpr int kMaxSeg = 3;      // q / k / v   or   gate / up
constexpr int kRowBytes = 128;  // bytes of one weight row per ring stage (256 q4 / 128 q8 elements)
constexpr int kQmvStages = 4;   // per-warp cp.async ring depth
constexpr int kQmvSlots = 2;    // exchange slots per warp for the end-of-tile hand-off (power of two)
constexpr int kQmvBarBytes = 512;  // mbarriers at the head of a qmv CTA's shared memory (63 barriers + a live count)
constexpr int kQmvProdWords = 8;   // per-thread producer words a fused caller parks in shared memory between stages
constexpr int kQmvMaxSmem = 227 * 1024;  // dynamic shared memory a qmv CTA may use (one CTA per SM)

enum Epilogue : int { EPI_STORE = 0, EPI_RESADD = 1, EPI_SILUMUL = 2 };

struct QmvSeg {
  const uint32_t* wq;
  const uint16_t* scales;
  const uint16_t* biases;
  int N;        // rows of this matrix
  int tile0;    // first tile index of this segment
  int out_off;  // column offset of this segment in the output row
};

struct QmvArgs {
  QmvSeg seg[kMaxSeg];
  int nseg;
  int dual;      // 1: seg[0] = gate, seg[1] = up; a tile is 8 gate rows + 8 up rows (SwiGLU epilogue)
  int M, K;
  int ntiles;
  int ksplit;    // cross-CTA K slices (gridDim.x % ksplit == 0); 1 unless x does not fit in shared memory
  int kslice;    // elements of K per slice (multiple of ks_w * the stage's K extent)
  int ks_w;      // warps per team: a tile's K slice is cut over ks_w warps of one CTA (power of two)
  int ks_w_log2;
  int nch;       // stages per (warp, tile) = kslice / (stage K extent * ks_w)
  uint32_t lift_mask, lift_magic;  // 0x000f000f, 0x43004300 (see lift4)
  const uint16_t* x;  // bf16 [M, K]
  int x_ld;
  // optional fused RMSNorm of x: x <- bf16(bf16(x * inv_rms) * norm_w)   (inv_rms computed per CTA over full K)
  const uint16_t* norm_w;
  float eps;
  uint16_t* out;
  int out_ld;
  const uint16_t* residual;  // EPI_RESADD (ld = out_ld)
  int epi;
  float* ws_part;        // [ntiles][ksplit][128 * mt] fp32, only if ksplit > 1
  unsigned int* ws_cnt;  // [ntiles], zero before the first launch, self-cleaning
};

// tensor-parallel exchange buffers (pie_comm.cu, pie_qmv2.cu): data [2 sets][world][max_elems] bf16, then flags
// [2 sets][world][kTpMaxCtas] u32
constexpr int kTpMaxWorld = 8;
constexpr int kTpMaxCtas = 32;
__host__ __device__ inline size_t tp_flags_off(int world, int max_elems) {
  return (((size_t)2 * world * max_elems * 2) + 255) & ~(size_t)255;
}

struct KvLayout {
  long long batch_stride, head_stride, tok_stride;
  const int32_t* page_table;
  int pt_stride, page_size;
};

// ---- launchers (return cudaError_t of the launch) -------------------------------------------------
cudaError_t launch_qmv(const QmvArgs& a, int bits, int warps, int grid, bool pdl, cudaStream_t st);
cudaError_t launch_qmv_generic(uint16_t* y, const uint16_t* x, const uint32_t* wq, const uint16_t* s,
                               const uint16_t* b, int M, int N, int K, int gs, int bits, cudaStream_t st);
size_t qmv_smem_bytes(int M, int kslice, int bits, int warps, int mt);
cudaError_t launch_qmm_tc(uint16_t* y, const uint16_t* x, const uint32_t* wq, const uint16_t* s, const uint16_t* b,
                          int M, int N, int K, int bits, cudaStream_t st);
// batched-decode GEMM on tcgen05 (pie_qmd.cu): 9 .. 32 activation rows per pass
int qmd_pick_ksplit(int ntiles, int K, int dual, int bits);
size_t qmd_ws_bytes(int M, int ntiles, int K, int ksplit, bool norm);
bool qmd_supported(int mb, int K, int x_ld, int group_size, int bits);
cudaError_t launch_qmd(const QmvArgs& g, const uint16_t* x, int x_ld, int bits, float* ws_part, unsigned int* ws_cnt,
                       int ksplit, cudaStream_t st);
int qmv_mt(int M);
int num_sms();

cudaError_t launch_quantize(const void* w, int dtype, uint32_t* wq, void* s, void* b, long long rows, int K,
                            int gs, int bits, cudaStream_t st);
cudaError_t launch_dequantize(uint16_t* out, const uint32_t* wq, const uint16_t* s, const uint16_t* b,
                              const int32_t* ids, long long n_rows, long long table_rows, int K, int gs,
                              int bits, int* err_flag, cudaStream_t st);
cudaError_t launch_rms_norm(uint16_t* out, const uint16_t* x, const uint16_t* w, long long rows, int dims,
                            float eps, cudaStream_t st);
cudaError_t launch_rope(uint16_t* out, const uint16_t* x, long long n_batch, int L, int D, int dims,
                        int traditional, float base, float scale, int offset, const float* freqs,
                        cudaStream_t st);
cudaError_t launch_rope_kv_append(uint16_t* q_out, const uint16_t* qkv, uint16_t* kc, uint16_t* vc,
                                  const KvLayout& kv, const int32_t* pos, const float* freqs, int B, int Hq,
                                  int Hkv, int D, cudaStream_t st);
cudaError_t launch_kv_write(uint16_t* kc, uint16_t* vc, const KvLayout& kv, const uint16_t* k,
                            const uint16_t* v, int B, int Hkv, int L, int D, int offset, cudaStream_t st);
cudaError_t launch_decode_attn(uint16_t* out, const uint16_t* q, const uint16_t* kc, const uint16_t* vc,
                               const KvLayout& kv, const int32_t* pos, int B, int Hq, int Hkv, int D,
                               float scale, int nsplit, float* ws, cudaStream_t st);
cudaError_t launch_decode_attn_tc(uint16_t* out, const uint16_t* q, const uint16_t* kc, const uint16_t* vc,
                                  const KvLayout& kv, const int32_t* pos, int B, int Hq, int Hkv, int D, float scale,
                                  int nsplit, float* ws, cudaStream_t st);
cudaError_t launch_prefill_attn_tc(uint16_t* out, const uint16_t* q, const uint16_t* kc, const uint16_t* vc,
                                   const KvLayout& kv, int B, int Hq, int Hkv, int L, int D, int offset, float scale,
                                   cudaStream_t st);
cudaError_t launch_prefill_attn(uint16_t* out, const uint16_t* q, const uint16_t* kc, const uint16_t* vc,
                                const KvLayout& kv, int B, int Hq, int Hkv, int L, int D, int offset,
                                float scale, cudaStream_t st);
cudaError_t launch_silu_mul(uint16_t* out, const uint16_t* g, const uint16_t* u, long long n, cudaStream_t st);
cudaError_t launch_add(uint16_t* out, const uint16_t* a, const uint16_t* b, long long n, cudaStream_t st);
cudaError_t launch_argmax(int32_t* out, const uint16_t* logits, int rows, int vocab, int32_t* pos_adv,
                          int32_t* seq_out, int seq_ld, cudaStream_t st);

void count_launch(int n = 1);
// Function attributes (the dynamic shared-memory opt-in) are per device: `mask` remembers the devices already served.
inline bool first_use_on_device(unsigned long long& mask) {
  int dev = 0;
  cudaGetDevice(&dev);
  const unsigned long long bit = 1ull << (dev & 63);
  if (mask & bit) return false;
  mask |= bit;
  return true;
}

// Launch `kernel` as a programmatic dependent of the stream's previous kernel: it may be scheduled while that kernel is
// still running and must execute griddepcontrol.wait (pdl_wait) before touching anything an earlier kernel wrote.  Every
// kernel launched this way waits at its top, so completion stays transitive along the stream.  PIE_PDL_AUX=0: plain launch.
bool pdl_aux_enabled();
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st,
                              Args... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_aux_enabled() ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

struct QmvPlan { int warps, grid, ksplit, kslice, ks_w, ntiles, mt; };
QmvPlan plan_qmv(int M, long long rows_total, int K, int bits, int dual, int ksplit_hint, int force_warps);
size_t qmv_ws_bytes(const QmvPlan& p);
int build_qmv_args(QmvArgs& a, QmvPlan& p, void* out, int out_ld, const void* x, int x_ld, int m0, int mb, int K,
                   const pie_qmv_seg* segs, int nseg, int group_size, int bits, const void* norm_weight, float eps,
                   int epilogue, const void* residual, int ksplit_hint, void* workspace, size_t workspace_bytes,
                   int force_warps);
int status_of(cudaError_t e);  // cudaError_t -> pie_status, recording the error string

// int8-tensor-core batched-decode matmul for 8-bit weights (pie_qmi.cu)
bool qmi_supported(int mb, int K, int x_ld, int group_size, int bits, const pie_qmv_seg* segs, int nseg, int dual);
int qmi_pick_ksplit(int ntiles, int K, int dual);
size_t qmi_ws_bytes(int ntiles, int K, int ksplit);
cudaError_t launch_qmi(void* out, int out_ld, const void* x, int x_ld, int M, int K, const pie_qmv_seg* segs, int nseg,
                       const void* norm_w, float eps, int epi, const void* residual, int ksplit, uint8_t* ws,
                       unsigned int* ws_cnt, cudaStream_t st);

}  // namespace pie
