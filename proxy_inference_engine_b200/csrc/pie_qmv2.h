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

// (Scalars first, tensor maps last: a launch's parameters are cold in the constant cache, and everything the kernel's
// first microsecond needs then sits in two or three lines.)
struct Qmv2Args {
  int ntiles, upt;  // 16-row tiles (dual: 8 gate + 8 up rows); units per tile = ceil(K / 512)
  int tiles_q, tiles_r;   // stand-alone launch: CTA c owns tiles_q (+1 if c < tiles_r) consecutive tiles
  uint32_t upt_inv;       // ceil(2^32 / upt): n / upt == umulhi(n, upt_inv) for the unit numbers of a launch (upt > 1)
  int nseg, dual, M, K;
  // Copies per unit.  All maps are 4-D: weights (128 B, row, 256-k chunk, matrix), scales / biases (group, row, matrix,
  // scales|biases).  sb_stacked: a matrix's biases lie a fixed stride behind its scales (one copy brings both);
  // gu_stacked (dual): the up matrix lies a fixed stride behind the gate matrix (one copy brings both).  The engine
  // allocates its matrices that way (checkpoint.colocate_for_decode); other callers get one copy per array.
  int sb_stacked, gu_stacked;
  int dbg;  // experiments (PIE_QMV2_DBG): 1 = skip the arithmetic, 4 / 8 / 16 = where the first loads are issued (see the kernel)
  int f16;  // 1: activations staged as scaled fp16, weights lifted to fp16 (see q2_stage_x)
  int N[kMaxSeg], tile0[kMaxSeg], out_off[kMaxSeg];
  uint32_t lift_mask, lift_mask2, lift_magic;
  const uint16_t* x;
  int x_ld;
  const uint16_t* norm_w;
  float eps;
  uint16_t* out;
  int out_ld;
  const uint16_t* residual;
  int epi;
  // profiling aid (pie_debug_stamps): 16 x %globaltimer per CTA of this launch, nullptr = off
  unsigned long long* stamps;
  // fused tensor-parallel exchange (one row): 0 none; 1 producer -- the finished rows go into slot `tp_rank` of every
  // peer's buffer instead of `out`; 2 consumer -- x = x (the residual stream) + bf16(sum over ranks of the slots),
  // CTA 0 stores it to tp_h_out, then the usual norm / staging.  See pie_comm.cu for the buffer and the epoch protocol.
  int tp_mode, tp_rank, tp_world, tp_max_elems;
  unsigned int* tp_epoch;   // local: [0] exchange number, [1] consumer departures, [2] producer departures
  uint16_t* tp_h_out;
  uint8_t* tp_peer[8];
  CUtensorMap wmap[kMaxSeg], smap[kMaxSeg], bmap[kMaxSeg];           // this matrix (up to 3 row segments)
};

size_t qmv2_smem_bytes(int M, int K);
bool qmv2_supported(int M, int K, int bits, int group_size, const pie_qmv_seg* segs, int nseg, int dual);
int build_qmv2_args(Qmv2Args& a, void* out, int out_ld, const void* x, int x_ld, int M, int K, const pie_qmv_seg* segs,
                    int nseg, const void* norm_weight, float eps, int epilogue, const void* residual,
                    const pie_tp_fused* tp = nullptr);
cudaError_t launch_qmv2(const Qmv2Args& a, bool pdl, cudaStream_t st);
// profiling aid (pie_debug_stamps): slots of 256 CTAs x 16 stamps handed to consecutive launches
constexpr size_t kStampSlotBytes = 256 * 16 * 8;
void debug_stamps(void* buf, size_t bytes);
unsigned long long* next_stamp_slot();


}  // namespace pie
