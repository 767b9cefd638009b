// C ABI of libpie_b200.so (include/pie_b200.h): argument validation + dispatch.  No allocation, no state
// beyond a launch counter and the last CUDA error string.
#include "../../include/pie_b200.h"
#include "pie_internal.h"
#include "pie_qmv2.h"

#include <atomic>
#include <cstdio>
#include <cstdlib>
#include <cstring>

namespace pie {
static std::atomic<unsigned long long> g_launches{0};
static thread_local char g_cuda_err[256] = "";
void count_launch(int n) { g_launches.fetch_add((unsigned long long)n, std::memory_order_relaxed); }
bool pdl_aux_enabled() {
  static const bool on = [] { const char* e = getenv("PIE_PDL_AUX"); return !(e && e[0] == '0'); }();
  return on;
}

static int cuda_status(cudaError_t e) {
  if (e == cudaSuccess) return PIE_OK;
  snprintf(g_cuda_err, sizeof(g_cuda_err), "%s: %s", cudaGetErrorName(e), cudaGetErrorString(e));
  return PIE_ECUDA;
}
static inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }
static inline bool quant_supported(int gs, int bits) { return gs == 64 && (bits == 4 || bits == 8); }
static inline int stage_k(int bits) { return kRowBytes * 8 / bits; }


static int env_int(const char* name, int dflt) {
  const char* e = getenv(name);
  return e && *e ? atoi(e) : dflt;
}

// One CTA per SM.  ksplit (cross-CTA K slices) is the smallest divisor of the chunk count for which the CTA's
// slice of x fits beside the weight rings in shared memory; ks_w (warps per team) is the largest power of two that
// still leaves every warp >= 4 stages per tile, raised when there are too few tiles to occupy every team.
QmvPlan plan_qmv(int M, long long rows_total, int K, int bits, int dual, int ksplit_hint, int force_warps) {
  QmvPlan p;
  p.mt = qmv_mt(M);
  p.warps = p.mt == 4 ? 8 : 16;
  { const int w = env_int("PIE_QMV_WARPS", 0); if (w == 4 || w == 8 || (w == 16 && p.mt < 4)) p.warps = w; }
  if (force_warps) p.warps = force_warps;
  p.ntiles = (int)((rows_total + (dual ? 7 : 15)) / (dual ? 8 : 16));
  const int ck = stage_k(bits), chunks = K / ck;
  const int sms = num_sms();
  // the fused decode step keeps per-thread producer words and a little static shared memory beside the GEMV's own
  const size_t reserve = force_warps ? (size_t)force_warps * 32 * kQmvProdWords * 4 + 2048 : 0;
  p.ksplit = 0;
  for (int ks = ksplit_hint > 0 ? ksplit_hint : 1; ks <= chunks && ks <= sms; ++ks) {
    if (chunks % ks) continue;
    if (qmv_smem_bytes(M, K / ks, bits, p.warps, p.mt) + reserve <= (size_t)kQmvMaxSmem) { p.ksplit = ks; break; }
  }
  if (p.ksplit == 0) p.ksplit = chunks;  // cannot happen for M <= 32 (one chunk of x is <= 16 KB)
  p.kslice = K / p.ksplit;
  int ncta = sms * env_int("PIE_QMV_CTAS_PER_SM", 1) / p.ksplit;  // > 1 only with PIE_QMV_WARPS=8 (experiment)
  if (ncta > p.ntiles) ncta = p.ntiles;
  if (ncta < 1) ncta = 1;
  p.grid = ncta * p.ksplit;
  const int cps = chunks / p.ksplit;  // chunks per CTA slice
  const int min_stages = env_int("PIE_QMV_MIN_STAGES", 4);
  int ks_w = 1;
  while (ks_w * 2 <= p.warps && cps % (ks_w * 2) == 0 && cps / (ks_w * 2) >= min_stages) ks_w *= 2;
  int teams_wanted = p.ntiles / ncta;  // tiles per CTA: more teams than that would idle
  if (teams_wanted < 1) teams_wanted = 1;
  // ... but never down to one stage per warp and tile (measured, o_proj at 1 row: 5.9 us with 16 warps per tile, 5.5 with 8)
  while (p.warps / ks_w > teams_wanted && ks_w * 2 <= p.warps && cps % (ks_w * 2) == 0 && cps / (ks_w * 2) >= 2) ks_w *= 2;
  { const int f = env_int("PIE_QMV_KSW", 0); if (f > 0 && f <= p.warps && p.warps % f == 0 && cps % f == 0) ks_w = f; }
  p.ks_w = ks_w;
  return p;
}
// Workspace layout shared by the GEMV (split-K) and the batched-decode GEMM: arrival counters live in a FIXED region at
// the head of the buffer, partial sums and everything else behind it.  The kernels leave their counters at zero, so one
// zero-initialised buffer serves any sequence of shapes -- which only holds because the counter region does not move
// with the shape (a shape-dependent split let a later, larger matrix read an earlier matrix's partial sums as counters).
constexpr size_t kWsCntBytes = 128 << 10;  // 32768 GEMV tiles (524288 weight rows)
size_t qmv_ws_bytes(const QmvPlan& p) {
  if (p.ksplit <= 1) return 0;
  return kWsCntBytes + (size_t)p.ntiles * p.ksplit * 128 * p.mt * sizeof(float);
}

// Validate one fused GEMV and fill its launch descriptor for activation rows m0 .. m0 + mb - 1 (mb <= kMaxBatch).
// force_warps != 0 pins the warp count (the fused decode step runs every matrix with the same CTA shape).
int build_qmv_args(QmvArgs& a, QmvPlan& p, void* out, int out_ld, const void* x, int x_ld, int m0, int mb, int K,
                   const pie_qmv_seg* segs, int nseg, int group_size, int bits, const void* norm_weight, float eps,
                   int epilogue, const void* residual, int ksplit_hint, void* workspace, size_t workspace_bytes,
                   int force_warps) {
  if (!out || !x || !segs) return PIE_ENULL;
  if (!quant_supported(group_size, bits)) return PIE_EUNSUPPORTED;
  if (nseg < 1 || nseg > kMaxSeg || mb <= 0 || mb > kMaxBatch || K <= 0 || K % group_size != 0) return PIE_EINVAL_SHAPE;
  if (K % stage_k(bits) != 0 || x_ld % 8 != 0 || x_ld < K) return PIE_EUNSUPPORTED;
  if (epilogue < PIE_EPI_STORE || epilogue > PIE_EPI_SILUMUL) return PIE_EINVAL_SHAPE;
  if (epilogue == PIE_EPI_RESADD && !residual) return PIE_ENULL;
  const int dual = epilogue == PIE_EPI_SILUMUL;
  if (dual && (nseg != 2 || segs[0].N != segs[1].N)) return PIE_EINVAL_SHAPE;
  if (!aligned16(x) || (norm_weight && !aligned16(norm_weight))) return PIE_EALIGN;
  memset(&a, 0, sizeof(a));
  long long rows_total = 0;
  int tile = 0, off = 0;
  for (int i = 0; i < nseg; ++i) {
    if (!segs[i].wq || !segs[i].scales || !segs[i].biases) return PIE_ENULL;
    if (segs[i].N <= 0) return PIE_EINVAL_SHAPE;
    if (!aligned16(segs[i].wq) || (reinterpret_cast<uintptr_t>(segs[i].scales) & 7) ||
        (reinterpret_cast<uintptr_t>(segs[i].biases) & 7)) return PIE_EALIGN;
    a.seg[i].wq = segs[i].wq;
    a.seg[i].scales = (const uint16_t*)segs[i].scales;
    a.seg[i].biases = (const uint16_t*)segs[i].biases;
    a.seg[i].N = segs[i].N;
    a.seg[i].tile0 = tile;
    a.seg[i].out_off = off;
    tile += (segs[i].N + 15) / 16;
    off += segs[i].N;
    rows_total += segs[i].N;
  }
  if (dual) rows_total = segs[0].N;
  if (out_ld < (dual ? segs[0].N : off)) return PIE_EINVAL_SHAPE;
  a.nseg = nseg; a.dual = dual; a.K = K; a.x_ld = x_ld;
  a.norm_w = (const uint16_t*)norm_weight; a.eps = eps;
  a.out_ld = out_ld; a.epi = epilogue;
  p = plan_qmv(mb, rows_total, K, bits, dual, ksplit_hint, force_warps);
  if (!dual) p.ntiles = tile;
  const size_t need = qmv_ws_bytes(p);
  if (need > workspace_bytes || (need && !workspace)) return PIE_EWORKSPACE;
  if (need && (size_t)p.ntiles * 4 > kWsCntBytes) return PIE_EUNSUPPORTED;
  a.M = mb; a.ntiles = p.ntiles; a.ksplit = p.ksplit; a.kslice = p.kslice; a.ks_w = p.ks_w;
  a.ks_w_log2 = 0; while ((1 << a.ks_w_log2) < p.ks_w) ++a.ks_w_log2;
  a.nch = p.kslice / (stage_k(bits) * p.ks_w);
  a.lift_mask = 0x000f000fu; a.lift_magic = 0x43004300u;
  a.x = (const uint16_t*)x + (size_t)m0 * x_ld;
  a.out = (uint16_t*)out + (size_t)m0 * out_ld;
  a.residual = residual ? (const uint16_t*)residual + (size_t)m0 * out_ld : nullptr;
  a.ws_cnt = (unsigned int*)workspace;
  a.ws_part = (float*)((uint8_t*)workspace + kWsCntBytes);
  return PIE_OK;
}
int status_of(cudaError_t e) { return cuda_status(e); }

// Workspace of the batched-decode GEMM, laid out so that it can share the buffer with the GEMV: the counter region
// comes first and is as large as the GEMV's (both kernels leave their counters zero), then the normalised
// activations, then the split-K partials.
static inline size_t qmd_cnt_bytes(long long) { return kWsCntBytes; }
static size_t qmd_total_bytes(int M, long long rows_total, int K, int dual, int bits) {
  const int mb = M > kMaxBatch ? kMaxBatch : M;
  size_t worst = 0;
  const int t0 = (int)((rows_total + 127) / 128);
  for (int t = t0; t <= t0 + kMaxSeg; ++t) {
    const size_t b = qmd_ws_bytes(mb, t, K, qmd_pick_ksplit(t, K, dual, bits), true);
    if (b > worst) worst = b;
  }
  return qmd_cnt_bytes(rows_total) + worst;
}
}  // namespace pie

using namespace pie;

extern "C" {

int pie_abi_version(void) { return 2; }
const char* pie_status_string(int s) {
  switch (s) {
    case PIE_OK: return "ok";
    case PIE_EINVAL_SHAPE: return "invalid shape";
    case PIE_EINVAL_DTYPE: return "invalid dtype";
    case PIE_EUNSUPPORTED: return "unsupported configuration";
    case PIE_ECUDA: return "CUDA error";
    case PIE_ENULL: return "null pointer";
    case PIE_EALIGN: return "pointer not 16-byte aligned";
    case PIE_EWORKSPACE: return "workspace too small";
    case PIE_ESTATIC_LIMIT: return "request exceeds the library's static scratch (PIE_STATIC_SCRATCH_BYTES = 4 MiB)";
    default: return "unknown";
  }
}
const char* pie_last_cuda_error(void) { return g_cuda_err; }
unsigned long long pie_launch_count(void) { return g_launches.load(); }

int pie_quantize(const void* w, int dtype, uint32_t* wq, void* scales, void* biases, long long rows, int K,
                 int group_size, int bits, void* stream) {
  if (!w || !wq || !scales || !biases) return rows == 0 ? PIE_OK : PIE_ENULL;
  if (dtype != PIE_DTYPE_BF16 && dtype != PIE_DTYPE_F32) return PIE_EINVAL_DTYPE;
  if (group_size != 32 && group_size != 64 && group_size != 128) return PIE_EUNSUPPORTED;
  if (bits != 2 && bits != 4 && bits != 8) return PIE_EUNSUPPORTED;
  if (rows < 0 || K <= 0 || K % group_size != 0) return PIE_EINVAL_SHAPE;
  return cuda_status(launch_quantize(w, dtype, wq, scales, biases, rows, K, group_size, bits, (cudaStream_t)stream));
}

int pie_dequantize(void* out, const uint32_t* wq, const void* scales, const void* biases, const int32_t* ids,
                   long long n_rows, long long table_rows, int K, int group_size, int bits, void* stream) {
  if (n_rows == 0) return PIE_OK;
  if (!out || !wq || !scales || !biases) return PIE_ENULL;
  if (group_size != 32 && group_size != 64 && group_size != 128) return PIE_EUNSUPPORTED;
  if (bits != 2 && bits != 4 && bits != 8) return PIE_EUNSUPPORTED;
  if (n_rows < 0 || table_rows <= 0 || K <= 0 || K % group_size != 0) return PIE_EINVAL_SHAPE;
  return cuda_status(launch_dequantize((uint16_t*)out, wq, (const uint16_t*)scales, (const uint16_t*)biases, ids,
                                       n_rows, table_rows, K, group_size, bits, nullptr, (cudaStream_t)stream));
}

size_t pie_qmv_workspace_bytes(int M, long long n_total, int K, int bits, int dual, int ksplit_hint) {
  if (M <= 0 || n_total <= 0 || K <= 0 || (bits != 4 && bits != 8) || K % stage_k(bits) != 0) return 0;
  QmvPlan p = plan_qmv(M > kMaxBatch ? kMaxBatch : M, n_total, K, bits, dual, ksplit_hint, 0);
  p.ntiles += kMaxSeg;  // segments are tiled separately: up to one ragged tile each
  size_t need = qmv_ws_bytes(p);
  if (M > 8 && K % 128 == 0) {
    const size_t q = qmd_total_bytes(M, n_total, K, dual, bits);
    if (q > need) need = q;
  }
  if (bits == 8 && K % 128 == 0) {   // pie_qmi.cu: planes + group sums + split-K partials behind the counter region
    const int t0 = (int)((dual ? n_total : n_total) / 128);
    size_t worst = 0;
    for (int t = t0; t <= t0 + kMaxSeg; ++t) {
      const size_t b = qmi_ws_bytes(t, K, qmi_pick_ksplit(t, K, dual));
      if (b > worst) worst = b;
    }
    if (kWsCntBytes + worst > need) need = kWsCntBytes + worst;
  }
  return need;
}

int pie_qmv_fused(void* out, int out_ld, const void* x, int x_ld, int M, int K, const pie_qmv_seg* segs, int nseg,
                  int group_size, int bits, const void* norm_weight, float eps, int epilogue, const void* residual,
                  int ksplit_hint, void* workspace, size_t workspace_bytes, int pdl, void* stream) {
  return pie_qmv_fused_tp(out, out_ld, x, x_ld, M, K, segs, nseg, group_size, bits, norm_weight, eps, epilogue, residual,
                          ksplit_hint, workspace, workspace_bytes, pdl, nullptr, stream);
}

int pie_qmv_fused_tp(void* out, int out_ld, const void* x, int x_ld, int M, int K, const pie_qmv_seg* segs, int nseg,
                     int group_size, int bits, const void* norm_weight, float eps, int epilogue, const void* residual,
                     int ksplit_hint, void* workspace, size_t workspace_bytes, int pdl, const pie_tp_fused* tp,
                     void* stream) {
  if (M == 0) return PIE_OK;
  if (tp && !tp->mode) tp = nullptr;
  for (int m0 = 0; m0 < M; m0 += kMaxBatch) {
    const int mb = M - m0 < kMaxBatch ? M - m0 : kMaxBatch;
    QmvArgs a;
    QmvPlan p;
    const int st0 = build_qmv_args(a, p, out, out_ld, x, x_ld, m0, mb, K, segs, nseg, group_size, bits, norm_weight, eps,
                                   epilogue, residual, ksplit_hint, workspace, workspace_bytes, 0);
    if (st0 != PIE_OK) return st0;
    // TMA-fed GEMV (pie_qmv2.cu): 4-bit weights, up to 8 rows, K a multiple of 512 -- every matrix of a Llama decode
    // step.  PIE_QMV2=0 keeps the cp.async kernel below (B200, 8B shapes, one row, us cp.async / TMA: q/k/v 7.8 / 7.7,
    // o 5.5 / 5.3, gate+up 18.9 / 16.5, down 13.1 / 11.5, head 64.2 / 49.7).
    const int q2_mode = env_int("PIE_QMV2", 1);
    if (M <= 8 && x_ld % 8 == 0 && q2_mode && qmv2_supported(mb, K, bits, group_size, segs, nseg, a.dual)) {
      Qmv2Args a2;
      const int st1 = build_qmv2_args(a2, out, out_ld, x, x_ld, mb, K, segs, nseg, norm_weight, eps, epilogue, residual, tp);
      if (st1 == PIE_OK) {
        const int st2 = cuda_status(launch_qmv2(a2, pdl != 0, (cudaStream_t)stream));
        if (st2 != PIE_OK) return st2;
        continue;
      }
      if (tp) return st1;
    }
    if (tp) return PIE_EUNSUPPORTED;  // the fused exchange lives in the TMA GEMV (one row, 4-bit, K % 512 == 0)
    // Batched-decode GEMM on tcgen05 (pie_qmd.cu).  PIE_QMD: 0 never, 1 where it measured faster than the GEMV at 32
    // rows (B200, 8B shapes, us GEMV / tcgen05: q8 gate+up 121 / 98, q8 head 417 / 290, q4 head 233 / 217; the narrow
    // matrices and 4-bit gate+up / down are still faster on the GEMV: DESIGN.md section 4.6), 2 wherever it is supported.
    // 8-bit weights, more than 8 rows: the integer tensor cores take the packed weights as they are (pie_qmi.cu).
    // PIE_QMI=0 keeps the bf16 forms below, 2 also routes 1..8 rows here.
    const int qmi_mode = env_int("PIE_QMI", 1);
    if (qmi_mode && (mb > 8 || qmi_mode == 2) && qmi_supported(mb, K, x_ld, group_size, bits, segs, nseg, a.dual)) {
      int nt = 0;
      for (int i = 0; i < nseg; ++i) nt += segs[i].N / 128;
      if (a.dual) nt = segs[0].N / 128;
      const int ks = qmi_pick_ksplit(nt, K, a.dual);
      if (workspace && kWsCntBytes + qmi_ws_bytes(nt, K, ks) <= workspace_bytes && (size_t)nt * 32 <= kWsCntBytes) {
        const int st2 = cuda_status(launch_qmi(a.out, out_ld, a.x, x_ld, mb, K, segs, nseg, norm_weight, eps, epilogue,
                                               a.residual, ks, (uint8_t*)workspace + kWsCntBytes, (unsigned int*)workspace,
                                               (cudaStream_t)stream));
        if (st2 != PIE_OK) return st2;
        continue;
      }
    }
    const int qmd_mode = env_int("PIE_QMD", 1);  // read per call: the tests switch it
    long long rows_total = 0;
    int nt = 0;
    for (int i = 0; i < nseg; ++i) { rows_total += segs[i].N; nt += (segs[i].N + 127) / 128; }
    if (a.dual) { rows_total = 2LL * segs[0].N; nt = (segs[0].N + 127) / 128; }
    const bool qmd_pays = mb > 16 && ((bits == 8 && rows_total >= 16384) || rows_total >= 65536);
    if (qmd_mode && (qmd_mode == 2 || qmd_pays) && qmd_supported(mb, K, x_ld, group_size, bits) &&
        (!norm_weight || x_ld == K)) {
      if (a.dual) rows_total = segs[0].N;
      const int ks = qmd_pick_ksplit(nt, K, a.dual, bits);
      const size_t cnt = qmd_cnt_bytes(rows_total), xn_bytes = norm_weight ? (((size_t)mb * K * 2 + 255) & ~(size_t)255) : 0;
      if (workspace && cnt + qmd_ws_bytes(mb, nt, K, ks, norm_weight != nullptr) <= workspace_bytes) {
        uint8_t* wsb = (uint8_t*)workspace;
        const uint16_t* xin = a.x;
        int xin_ld = x_ld;
        if (norm_weight) {
          const int st1 = cuda_status(launch_rms_norm((uint16_t*)(wsb + cnt), a.x, a.norm_w, mb, K, eps, (cudaStream_t)stream));
          if (st1 != PIE_OK) return st1;
          xin = (const uint16_t*)(wsb + cnt);
          xin_ld = K;
        }
        float* part = (float*)(wsb + cnt + xn_bytes + (((size_t)nt * 16 + 255) & ~(size_t)255));
        const int st2 = cuda_status(launch_qmd(a, xin, xin_ld, bits, part, (unsigned int*)workspace, ks, (cudaStream_t)stream));
        if (st2 != PIE_OK) return st2;
        continue;
      }
    }
    const int st = cuda_status(launch_qmv(a, bits, p.warps, p.grid, pdl != 0, (cudaStream_t)stream));
    if (st != PIE_OK) return st;
  }
  return PIE_OK;
}

int pie_qmv(void* y, const void* x, const uint32_t* wq, const void* scales, const void* biases, int M, int N,
            int K, int group_size, int bits, void* workspace, size_t workspace_bytes, void* stream) {
  if (M == 0 || N == 0) return PIE_OK;
  if (!y || !x || !wq || !scales || !biases) return PIE_ENULL;
  if (M < 0 || N < 0 || K <= 0) return PIE_EINVAL_SHAPE;
  const bool fast = quant_supported(group_size, bits) && K % stage_k(bits) == 0 && aligned16(x) && aligned16(wq) &&
                    !(reinterpret_cast<uintptr_t>(scales) & 7) && !(reinterpret_cast<uintptr_t>(biases) & 7);
  if (fast) {
    pie_qmv_seg seg = {wq, scales, biases, N};
    return pie_qmv_fused(y, N, x, K, M, K, &seg, 1, group_size, bits, nullptr, 0.f, PIE_EPI_STORE, nullptr, 0,
                         workspace, workspace_bytes, 0, stream);
  }
  if (group_size != 32 && group_size != 64 && group_size != 128) return PIE_EUNSUPPORTED;
  if (bits != 2 && bits != 4 && bits != 8) return PIE_EUNSUPPORTED;
  if (K % group_size != 0) return PIE_EINVAL_SHAPE;
  return cuda_status(launch_qmv_generic((uint16_t*)y, (const uint16_t*)x, wq, (const uint16_t*)scales,
                                        (const uint16_t*)biases, M, N, K, group_size, bits, (cudaStream_t)stream));
}

// profiling aid: from now on every TMA GEMV / fused-attention launch writes %globaltimer stamps of its CTAs into the next
// 32 KB slot of `buf` (device memory, zeroed by the caller); nullptr switches it off.  See tools/timeline_step.py.
int pie_debug_stamps(void* buf, size_t bytes) {
  debug_stamps(buf, bytes);
  return PIE_OK;
}

}  // extern "C"
#include "pie_api2.inl"
