// TEST INFRASTRUCTURE ONLY -- never linked into, imported by or executed from the product path.
//
// Thin C-ABI harness around the reference's OWN CPU code for the quantized path.  It does not
// copy any reference source: the single translation unit below #includes the vendored MLX file
// where it lies under /root/reference and re-exports its anonymous-namespace templates
//   quantize<T,U>            MLX/mlx/backend/cpu/quantized.cpp:603-661
//   _qmm_t<T,bits,gs>        MLX/mlx/backend/cpu/quantized.cpp:96-153   (the x86 path: simd::max_size==1,
//                            MLX/mlx/backend/cpu/simd/base_simd.h:13-14 -> dispatch at quantized.cpp:240-244)
//   _qmm<T,bits,gs>          MLX/mlx/backend/cpu/quantized.cpp:38-94    (transpose=False: the values product of
//                            quantized_scaled_dot_product_attention, PIE/models/base.py:81-83)
// (MLX = /root/reference/src/pie_core/external/mlx).  Built by oracle/Makefile into
// oracle/_ref/libpie_ref.so (git-ignored, NOT gpurun-ignored).  Unreferenced eval_cpu bodies that need
// libmlx symbols are discarded with -ffunction-sections/--gc-sections.
//
// The row-sliced *_mt variants call the very same reference template on disjoint row ranges from
// several OpenMP threads (the reference itself runs one thread per stream,
// MLX/mlx/scheduler.h:22-35); they exist only so the timed CPU baseline can use the host's cores.

#include "mlx/backend/cpu/quantized.cpp"

#include <cstdint>
#include <cstring>
#ifdef _OPENMP
#include <omp.h>
#endif

using mlx::core::bfloat16_t;

namespace {

template <typename T>
void qmm_t_dispatch(T* y, const T* x, const uint32_t* w, const T* s, const T* b,
                    int M, int N, int K, int bits, int gs) {
  using namespace mlx::core;
#define PIE_CASE(B, G) \
  if (bits == B && gs == G) { _qmm_t<T, B, G>(y, x, w, s, b, M, N, K); return; }
  PIE_CASE(4, 64) PIE_CASE(8, 64) PIE_CASE(4, 32) PIE_CASE(8, 32) PIE_CASE(4, 128) PIE_CASE(8, 128)
  PIE_CASE(2, 64) PIE_CASE(2, 32) PIE_CASE(2, 128)
#undef PIE_CASE
}

template <typename T>
void qmm_n_dispatch(T* y, const T* x, const uint32_t* w, const T* s, const T* b,
                    int M, int N, int K, int bits, int gs) {
  using namespace mlx::core;
#define PIE_CASE(B, G) \
  if (bits == B && gs == G) { _qmm<T, B, G>(y, x, w, s, b, M, N, K); return; }
  PIE_CASE(4, 64) PIE_CASE(8, 64) PIE_CASE(4, 32) PIE_CASE(8, 32) PIE_CASE(4, 128) PIE_CASE(8, 128)
#undef PIE_CASE
}

template <typename T>
void qmm_t_rows_mt(T* y, const T* x, const uint32_t* w, const T* s, const T* b,
                   int M, int N, int K, int bits, int gs, int nthreads) {
  // y is [M, N]; the reference writes rows of length N, so slice N only when M == 1 and otherwise
  // slice M (each call still runs the untouched reference loop nest).
  const size_t wrow = (size_t)K * bits / 32;  // u32 words per weight row
  const size_t grow = (size_t)K / gs;
#ifdef _OPENMP
  if (nthreads > 0) omp_set_num_threads(nthreads);
#endif
  if (M == 1) {
    const int chunk = 64;
    const int nchunks = (N + chunk - 1) / chunk;
#pragma omp parallel for schedule(static)
    for (int c = 0; c < nchunks; ++c) {
      int n0 = c * chunk;
      int n = (N - n0 < chunk) ? (N - n0) : chunk;
      qmm_t_dispatch<T>(y + n0, x, w + n0 * wrow, s + n0 * grow, b + n0 * grow, 1, n, K, bits, gs);
    }
  } else {
#pragma omp parallel for schedule(static)
    for (int m = 0; m < M; ++m) {
      qmm_t_dispatch<T>(y + (size_t)m * N, x + (size_t)m * K, w, s, b, 1, N, K, bits, gs);
    }
  }
}

}  // namespace

#pragma GCC visibility push(default)
extern "C" {

int pie_ref_simd_max_size_f32() { return mlx::core::simd::max_size<float>; }

// ---- quantize<T, uint32_t>  (quantized.cpp:603-661) -------------------------------------------
void pie_ref_quantize_f32(const float* w, uint32_t* out, float* scales, float* biases,
                          int bits, int group_size, size_t w_size) {
  mlx::core::quantize<float, uint32_t>(w, out, scales, biases, bits, group_size, w_size);
}
// bf16 tensors cross the ABI as raw uint16 bit patterns.
void pie_ref_quantize_bf16(const uint16_t* w, uint32_t* out, uint16_t* scales, uint16_t* biases,
                           int bits, int group_size, size_t w_size) {
  static_assert(sizeof(bfloat16_t) == 2, "bf16 layout");
  mlx::core::quantize<bfloat16_t, uint32_t>(
      reinterpret_cast<const bfloat16_t*>(w), out, reinterpret_cast<bfloat16_t*>(scales),
      reinterpret_cast<bfloat16_t*>(biases), bits, group_size, w_size);
}

// ---- _qmm_t<T,bits,gs>  (quantized.cpp:96-153) -------------------------------------------------
void pie_ref_qmm_t_f32(float* y, const float* x, const uint32_t* w, const float* s, const float* b,
                       int M, int N, int K, int bits, int group_size) {
  qmm_t_dispatch<float>(y, x, w, s, b, M, N, K, bits, group_size);
}
void pie_ref_qmm_t_bf16(uint16_t* y, const uint16_t* x, const uint32_t* w, const uint16_t* s,
                        const uint16_t* b, int M, int N, int K, int bits, int group_size) {
  qmm_t_dispatch<bfloat16_t>(
      reinterpret_cast<bfloat16_t*>(y), reinterpret_cast<const bfloat16_t*>(x), w,
      reinterpret_cast<const bfloat16_t*>(s), reinterpret_cast<const bfloat16_t*>(b), M, N, K, bits,
      group_size);
}
void pie_ref_qmm_t_f32_mt(float* y, const float* x, const uint32_t* w, const float* s, const float* b,
                          int M, int N, int K, int bits, int group_size, int nthreads) {
  qmm_t_rows_mt<float>(y, x, w, s, b, M, N, K, bits, group_size, nthreads);
}
void pie_ref_qmm_t_bf16_mt(uint16_t* y, const uint16_t* x, const uint32_t* w, const uint16_t* s,
                           const uint16_t* b, int M, int N, int K, int bits, int group_size,
                           int nthreads) {
  qmm_t_rows_mt<bfloat16_t>(
      reinterpret_cast<bfloat16_t*>(y), reinterpret_cast<const bfloat16_t*>(x), w,
      reinterpret_cast<const bfloat16_t*>(s), reinterpret_cast<const bfloat16_t*>(b), M, N, K, bits,
      group_size, nthreads);
}

// ---- _qmm<T,bits,gs>  (quantized.cpp:38-94): y [M, N] = x [M, K] @ dequant(w [K, N]) -------------------
void pie_ref_qmm_n_f32(float* y, const float* x, const uint32_t* w, const float* s, const float* b,
                       int M, int N, int K, int bits, int group_size) {
  qmm_n_dispatch<float>(y, x, w, s, b, M, N, K, bits, group_size);
}
void pie_ref_qmm_n_bf16(uint16_t* y, const uint16_t* x, const uint32_t* w, const uint16_t* s,
                        const uint16_t* b, int M, int N, int K, int bits, int group_size) {
  qmm_n_dispatch<bfloat16_t>(
      reinterpret_cast<bfloat16_t*>(y), reinterpret_cast<const bfloat16_t*>(x), w,
      reinterpret_cast<const bfloat16_t*>(s), reinterpret_cast<const bfloat16_t*>(b), M, N, K, bits,
      group_size);
}

int pie_ref_max_threads() {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}

}  // extern "C"
#pragma GCC visibility pop
