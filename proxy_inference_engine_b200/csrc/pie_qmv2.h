// Launch descriptor of the TMA-fed decode GEMV (pie_qmv2.cu).  Private to the library.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "pie_internal.h"

namespace pie {

constexpr int kQ2Warps = 16;       // warps per CTA, each with a private TMA ring
constexpr int kQ2Stages = 2;       // ring depth per warp (a stage holds one unit)
constexpr int kQ2UnitK = 512;      // k extent of a unit (8 quantization groups)
constexpr int kQ2WBytes = 4096;    // packed 4-bit weights of a unit: 16 rows x 256 B
constexpr int kQ2SbBytes = 512;    // bf16 scales (256 B) + biases (256 B) of a unit

struct Qmv2Args {
  CUtensorMap wmap[kMaxSeg], smap[kMaxSeg], bmap[kMaxSeg];           // this matrix (up to 3 row segments)
  int N[kMaxSeg], tile0[kMaxSeg], out_off[kMaxSeg];
  int nseg, dual, M, K;
  int ntiles, upt;  // 16-row tiles (dual: 8 gate + 8 up rows); units per tile = ceil(K / 512)
  uint32_t lift_mask, lift_mask2, lift_magic;
  int dbg;  // experiments (PIE_QMV2_DBG): 1 = skip the arithmetic, 2 = load weights only
  int f16;  // 1: activations staged as scaled fp16, weights lifted to fp16 (see q2_stage_x)
  const uint16_t* x;
  int x_ld;
  const uint16_t* norm_w;
  float eps;
  uint16_t* out;
  int out_ld;
  const uint16_t* residual;
  int epi;
  // fused tensor-parallel exchange (one row): 0 none; 1 producer -- the finished rows go into slot `tp_rank` of every
  // peer's buffer instead of `out`; 2 consumer -- x = x (the residual stream) + bf16(sum over ranks of the slots),
  // CTA 0 stores it to tp_h_out, then the usual norm / staging.  See pie_comm.cu for the buffer and the epoch protocol.
  int tp_mode, tp_rank, tp_world, tp_max_elems;
  uint8_t* tp_peer[8];
  unsigned int* tp_epoch;   // local: [0] exchange number, [1] consumer departures, [2] producer departures
  uint16_t* tp_h_out;
};

size_t qmv2_smem_bytes(int M, int K);
bool qmv2_supported(int M, int K, int bits, int group_size, const pie_qmv_seg* segs, int nseg, int dual);
int build_qmv2_args(Qmv2Args& a, void* out, int out_ld, const void* x, int x_ld, int M, int K, const pie_qmv_seg* segs,
                    int nseg, const void* norm_weight, float eps, int epilogue, const void* residual,
                    const pie_tp_fused* tp = nullptr);
cudaError_t launch_qmv2(const Qmv2Args& a, bool pdl, cudaStream_t st);


}  // namespace pie
