// common.cuh — shared device helpers for libdroid_b200 (sm_100a).
#pragma once
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <stdint.h>
#include "../../include/droid_b200.h"

// last CUDA error seen by the library (for db200_last_cuda_error()); defined in ba_api.cu
extern "C" void db200_note_cuda_error(int code, const char* file, int line);

// kernel-launch counter (db200_launch_count(); bench.py's "gpu_launches"); defined in ba_api.cu
extern "C" void db200_count_launches(int n);
#define DB200_COUNT(n) db200_count_launches(n)

#define DB200_CHECK_LAUNCH()                                   \
  do {                                                         \
    cudaError_t e__ = cudaGetLastError();                      \
    DB200_COUNT(1);                                            \
    if (e__ != cudaSuccess) {                                  \
      db200_note_cuda_error(int(e__), __FILE__, __LINE__);     \
      return DB200_ERR_CUDA;                                   \
    }                                                          \
  } while (0)
#define DB200_CUDA(call)                                       \
  do {                                                         \
    cudaError_t e__ = (call);                                  \
    if (e__ != cudaSuccess) {                                  \
      db200_note_cuda_error(int(e__), __FILE__, __LINE__);     \
      return DB200_ERR_CUDA;                                   \
    }                                                          \
  } while (0)

namespace db200 {

constexpr int kNumSMs = 148;  // B200

__host__ __device__ inline int cdiv(int a, int b) { return (a + b - 1) / b; }

// ---- element-type helpers: arithmetic follows the reference's `scalar_t` semantics -------------
// c10::Half arithmetic = convert to float, operate, round back to half after every operation.
template <typename T> struct Elem;
template <> struct Elem<float> {
  static __device__ __forceinline__ float from_float(float x) { return x; }
  // acc += s * w as compiled for fp32: one FFMA
  static __device__ __forceinline__ float mad(float s, float w, float acc) { return __fmaf_rn(s, w, acc); }
  static __device__ __forceinline__ float zero() { return 0.0f; }
};
template <> struct Elem<double> {
  static __device__ __forceinline__ double from_float(float x) { return double(x); }
  static __device__ __forceinline__ double mad(double s, double w, double acc) { return fma(s, w, acc); }
  static __device__ __forceinline__ double zero() { return 0.0; }
};
template <> struct Elem<__half> {
  static __device__ __forceinline__ __half from_float(float x) { return __float2half_rn(x); }
  // c10::Half computes Half(float(s) * float(w)) and Half(float(acc) + float(p)).  The float product of two halves
  // is exact, and the float sum of two halves is either exact (exponent gap <= 13) or the small operand lies below
  // a quarter ulp of the large one, so rounding once (native HMUL/HADD, no fusion) gives the same bits.
  static __device__ __forceinline__ __half mad(__half s, __half w, __half acc) {
    return __hadd_rn(acc, __hmul_rn(s, w));
  }
  static __device__ __forceinline__ __half zero() { return __float2half_rn(0.0f); }
};

// ---- SE3 helpers (quaternion x,y,z,w), free-form fp32 (used where the tolerance is 1e-4) --------
struct Rigid {   // relative pose as rotation matrix + translation
  float R[9];
  float t[3];
};

__device__ __forceinline__ void quat_to_R(const float* q, float* R) {
  const float x = q[0], y = q[1], z = q[2], w = q[3];
  R[0] = 1 - 2 * (y * y + z * z); R[1] = 2 * (x * y - z * w);     R[2] = 2 * (x * z + y * w);
  R[3] = 2 * (x * y + z * w);     R[4] = 1 - 2 * (x * x + z * z); R[5] = 2 * (y * z - x * w);
  R[6] = 2 * (x * z - y * w);     R[7] = 2 * (y * z + x * w);     R[8] = 1 - 2 * (x * x + y * y);
}

__device__ __forceinline__ void act_so3(const float* q, const float* X, float* Y) {
  float uv[3];
  uv[0] = 2.0f * (q[1] * X[2] - q[2] * X[1]);
  uv[1] = 2.0f * (q[2] * X[0] - q[0] * X[2]);
  uv[2] = 2.0f * (q[0] * X[1] - q[1] * X[0]);
  Y[0] = X[0] + q[3] * uv[0] + (q[1] * uv[2] - q[2] * uv[1]);
  Y[1] = X[1] + q[3] * uv[1] + (q[2] * uv[0] - q[0] * uv[2]);
  Y[2] = X[2] + q[3] * uv[2] + (q[0] * uv[1] - q[1] * uv[0]);
}

// relative pose Gij = Gj * Gi^-1  (reference relSE3, droid_kernels.cu:97-107)
__device__ __forceinline__ void rel_se3(const float* pi, const float* pj, float* tij, float* qij) {
  const float* ti = pi; const float* qi = pi + 3;
  const float* tj = pj; const float* qj = pj + 3;
  qij[0] = -qj[3] * qi[0] + qj[0] * qi[3] - qj[1] * qi[2] + qj[2] * qi[1];
  qij[1] = -qj[3] * qi[1] + qj[1] * qi[3] - qj[2] * qi[0] + qj[0] * qi[2];
  qij[2] = -qj[3] * qi[2] + qj[2] * qi[3] - qj[0] * qi[1] + qj[1] * qi[0];
  qij[3] = qj[3] * qi[3] + qj[0] * qi[0] + qj[1] * qi[1] + qj[2] * qi[2];
  float Y[3];
  act_so3(qij, ti, Y);
  tij[0] = tj[0] - Y[0]; tij[1] = tj[1] - Y[1]; tij[2] = tj[2] - Y[2];
}

// Tiled slice layout (private to CorrBlock's fused path; include/droid_b200.h, "tiled levels"): the [h2][w2] slice of a source
// pixel is stored as 4x8-element tiles (64 bytes of fp16 = one DRAM burst), tiles in row-major order, the 4 rows of a tile 16
// bytes apart — so the 8x8-sample window of a lookup touches ~5 bursts instead of 8-10 (in the row-major layout every 16-byte
// row segment of a window costs a whole burst).  A level is tiled when its height is a multiple of 4 and its width a multiple
// of 8 and at least 32 (smaller levels already sit in a few bursts).
__host__ __device__ inline bool level_is_tiled(int h2l, int w2l) { return h2l % 4 == 0 && w2l % 8 == 0 && w2l >= 32; }
// position of element (y, x) inside a tiled [h][w] slice
__host__ __device__ inline int tiled_index(int y, int x, int w) { return (((y >> 2) * (w >> 3) + (x >> 3)) << 5) + ((y & 3) << 3) + (x & 7); }

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

}  // namespace db200
