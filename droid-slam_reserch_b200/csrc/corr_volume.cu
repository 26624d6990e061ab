// corr_volume.cu — all-pairs 128-d correlation volume + 4-level pyramid.
// Replaces CorrBlock.corr (torch.matmul) and the three avg_pool2d passes of
// /root/reference/droid_slam/modules/corr.py:30-38,63-71.
//
// corr[n,p1,p2] = sum_c (f1[n,c,p1]/4)(f2[n,c,p2]/4).  Average pooling over (h2,w2) commutes with the
// contraction over c, so level l is computed as f1^T * pool_l(f2): the pooled feature maps are tiny
// (1/4, 1/16, 1/64 of f2) and every level is WRITTEN ONCE instead of being re-read and re-written by three
// pooling passes (the volume is output-write bound: 18.9 MB/edge at level 0 in fp16).
// The 1/16 scale is exact in binary floating point and is applied to the fp32 accumulator.
//
// v1 data path: tensor-core MMA through the portable warp-level API (m16n16k16, fp16 in / fp32 acc),
// fragments read straight from the L2-resident feature maps ([C][HW]: A is "col-major" MxK, B is
// "row-major" KxN, no transposes needed).  The fp32 (training) path uses an SIMT fp32 tile kernel so that
// values stay within 1e-4 of the fp32 reference.  [tcgen05/TMA version: see DESIGN.md roadmap]
#include <mma.h>
#include <cstdlib>
#include "common.cuh"

namespace db200 {

using namespace nvcuda;

// pooled[n,c,y,x] = mean of the 2x2 block of src (floor mode, like F.avg_pool2d(x, 2, stride=2))
template <typename T>
__global__ void pool_fmap_kernel(const T* __restrict__ src, T* __restrict__ dst, int NC, int hs, int ws) {
  const int hd = hs >> 1, wdd = ws >> 1;
  const size_t total = size_t(NC) * hd * wdd;
  for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += size_t(gridDim.x) * blockDim.x) {
    const int x = int(i % wdd);
    const int y = int((i / wdd) % hd);
    const size_t nc = i / (size_t(wdd) * hd);
    const T* s = src + (nc * hs + 2 * y) * ws + 2 * x;
    const float v = (float(s[0]) + float(s[1]) + float(s[ws]) + float(s[ws + 1])) * 0.25f;
    dst[i] = T(v);
  }
}

// fp16: one warp = 32x32 output tile; CTA = 8 warps = 64 (p1) x 128 (p2).  grid (ceil(N2/128), ceil(M/64), n)
__global__ void __launch_bounds__(256)
corr_gemm_f16_kernel(const __half* __restrict__ f1, const __half* __restrict__ f2, __half* __restrict__ out, int C, int M,
                     int N2) {
  __shared__ __align__(32) float stage[8][32 * 32];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n = blockIdx.z;
  const int m0 = blockIdx.y * 64 + (warp >> 2) * 32;
  const int n0 = blockIdx.x * 128 + (warp & 3) * 32;
  if (m0 >= M || n0 >= N2) return;
  const __half* A = f1 + size_t(n) * C * M;     // A(m,k) = A[k*M + m]
  const __half* B = f2 + size_t(n) * C * N2;    // B(k,j) = B[k*N2 + j]
  // wmma loads need 16-byte aligned rows: leading dimensions must be multiples of 8 halfs
  const bool full = (m0 + 32 <= M) && (n0 + 32 <= N2) && (M % 8 == 0) && (N2 % 8 == 0);
  float* st = stage[warp];
  if (full) {
    wmma::fragment<wmma::accumulator, 16, 16, 16, float> acc[2][2];
#pragma unroll
    for (int i = 0; i < 2; i++)
#pragma unroll
      for (int j = 0; j < 2; j++) wmma::fill_fragment(acc[i][j], 0.0f);
    for (int k = 0; k < C; k += 16) {
      wmma::fragment<wmma::matrix_a, 16, 16, 16, __half, wmma::col_major> a[2];
      wmma::fragment<wmma::matrix_b, 16, 16, 16, __half, wmma::row_major> b[2];
#pragma unroll
      for (int i = 0; i < 2; i++) wmma::load_matrix_sync(a[i], A + size_t(k) * M + m0 + 16 * i, M);
#pragma unroll
      for (int j = 0; j < 2; j++) wmma::load_matrix_sync(b[j], B + size_t(k) * N2 + n0 + 16 * j, N2);
#pragma unroll
      for (int i = 0; i < 2; i++)
#pragma unroll
        for (int j = 0; j < 2; j++) wmma::mma_sync(acc[i][j], a[i], b[j], acc[i][j]);
    }
#pragma unroll
    for (int i = 0; i < 2; i++)
#pragma unroll
      for (int j = 0; j < 2; j++) wmma::store_matrix_sync(st + (16 * i) * 32 + 16 * j, acc[i][j], 32, wmma::mem_row_major);
  } else {
    // ragged edge tile: scalar fp32 accumulation
    for (int e = lane; e < 32 * 32; e += 32) {
      const int r = e >> 5, c = e & 31;
      float s = 0.0f;
      if (m0 + r < M && n0 + c < N2)
        for (int k = 0; k < C; k++) s = fmaf(__half2float(A[size_t(k) * M + m0 + r]), __half2float(B[size_t(k) * N2 + n0 + c]), s);
      st[e] = s;
    }
  }
  __syncwarp();
  __half* O = out + size_t(n) * M * N2;
  for (int r = 0; r < 32; r++) {
    if (m0 + r < M && n0 + lane < N2) O[size_t(m0 + r) * N2 + n0 + lane] = __float2half_rn(st[r * 32 + lane] * 0.0625f);
  }
}

// fp32: 64x64 output tile per CTA, 4x4 per thread, K staged through shared memory in slabs of 16.
__global__ void __launch_bounds__(256)
corr_gemm_f32_kernel(const float* __restrict__ f1, const float* __restrict__ f2, float* __restrict__ out, int C, int M, int N2) {
  __shared__ float As[16][64 + 1], Bs[16][64 + 1];
  const int n = blockIdx.z, m0 = blockIdx.y * 64, n0 = blockIdx.x * 64;
  const float* A = f1 + size_t(n) * C * M;
  const float* B = f2 + size_t(n) * C * N2;
  const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
  float acc[4][4] = {};
  for (int k0 = 0; k0 < C; k0 += 16) {
    for (int i = threadIdx.x; i < 16 * 64; i += 256) {
      const int kk = i >> 6, c = i & 63;
      As[kk][c] = (k0 + kk < C && m0 + c < M) ? A[size_t(k0 + kk) * M + m0 + c] * 0.25f : 0.0f;   // fmap/4 (corr.py:66-67)
      Bs[kk][c] = (k0 + kk < C && n0 + c < N2) ? B[size_t(k0 + kk) * N2 + n0 + c] * 0.25f : 0.0f;
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < 16; kk++) {
      float a[4], b[4];
#pragma unroll
      for (int i = 0; i < 4; i++) { a[i] = As[kk][ty + 16 * i]; b[i] = Bs[kk][tx + 16 * i]; }
#pragma unroll
      for (int i = 0; i < 4; i++)
#pragma unroll
        for (int j = 0; j < 4; j++) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    __syncthreads();
  }
  float* O = out + size_t(n) * M * N2;
#pragma unroll
  for (int i = 0; i < 4; i++)
#pragma unroll
    for (int j = 0; j < 4; j++) {
      const int r = m0 + ty + 16 * i, c = n0 + tx + 16 * j;
      if (r < M && c < N2) O[size_t(r) * N2 + c] = acc[i][j];
    }
}

// tcgen05/TMA path (corr_volume_tc.cu)
size_t corr_tc_workspace_bytes(int N, int C, int ht, int wd, int num_levels);
bool corr_tc_supported(int C, int ht, int wd);
int corr_volume_pyramid_tc(const __half* fmap1, const __half* fmap2, void* const* levels_out, int N, int C, int ht, int wd,
                           int num_levels, void* workspace, cudaStream_t st, unsigned tiled_mask);

size_t corr_tc_f32_workspace_bytes(int N, int C, int ht, int wd, int num_levels);
bool corr_tc_f32_supported(int C, int ht, int wd, int num_levels);
int corr_volume_pyramid_tc_f32(const float* fmap1, const float* fmap2, void* const* levels_out, int N, int C, int ht, int wd,
                               int num_levels, void* workspace, cudaStream_t st);

}  // namespace db200

using namespace db200;

static bool use_tc_path(int C, int ht, int wd, int dtype) {
  static const bool off = getenv("DB200_CORR_NO_TCGEN05") != nullptr;     // debugging aid: force the mma.sync path
  return !off && dtype == DB200_F16 && corr_tc_supported(C, ht, wd);
}

static bool use_tc_f32_path(int C, int ht, int wd, int num_levels) {
  static const bool off = getenv("DB200_CORR_NO_TCGEN05") != nullptr;
  return !off && corr_tc_f32_supported(C, ht, wd, num_levels);
}

extern "C" size_t db200_corr_volume_workspace_bytes(int N, int C, int ht, int wd, int num_levels, int dtype) {
  if (N <= 0 || C <= 0 || ht <= 0 || wd <= 0 || num_levels < 1) return 0;
  const size_t es = dtype == DB200_F16 ? 2 : 4;
  size_t tot = 256;
  int h = ht, w = wd;
  for (int l = 1; l < num_levels; l++) { h >>= 1; w >>= 1; tot += ((size_t(N) * C * h * w * es + 255) / 256) * 256; }
  if (dtype == DB200_F16) { const size_t t2 = corr_tc_workspace_bytes(N, C, ht, wd, num_levels); if (t2 > tot) tot = t2; }
  if (dtype == DB200_F32 && use_tc_f32_path(C, ht, wd, num_levels)) {
    const size_t t2 = corr_tc_f32_workspace_bytes(N, C, ht, wd, num_levels);
    if (t2 > tot) tot = t2;
  }
  return tot;
}

extern "C" unsigned db200_corr_tiled_levels(int C, int ht, int wd, int num_levels, int dtype) {
  if (num_levels < 1 || num_levels > 4 || C <= 0 || ht <= 0 || wd <= 0 || !use_tc_path(C, ht, wd, dtype)) return 0u;
  unsigned m = 0;
  for (int l = 0; l < num_levels; l++) {
    if ((wd >> l) % 8 != 0 || (ht >> l) <= 0) return 0u;          // the fused lookup kernel wants whole 16-byte chunks per row
    if (level_is_tiled(ht >> l, wd >> l)) m |= 1u << l;
  }
  return m;
}

static int volume_pyramid(const void* fmap1, const void* fmap2, void* const* levels_out, int N, int C, int ht, int wd,
                          int num_levels, int dtype, void* workspace, size_t workspace_bytes, unsigned tiled_mask,
                          db200_stream_t stream) {
  if (!fmap1 || !fmap2 || !levels_out || N < 0 || C <= 0 || ht <= 0 || wd <= 0 || num_levels < 1 || num_levels > 4)
    return DB200_ERR_ARG;
  if (dtype != DB200_F16 && dtype != DB200_F32) return DB200_ERR_ARG;
  if (N == 0) return DB200_OK;
  for (int l = 0; l < num_levels; l++) if (!levels_out[l]) return DB200_ERR_ARG;
  if (num_levels > 1 && (!workspace || workspace_bytes < db200_corr_volume_workspace_bytes(N, C, ht, wd, num_levels, dtype)))
    return DB200_ERR_WORKSPACE;
  if (dtype == DB200_F16 && C % 16 != 0) return DB200_ERR_ARG;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (use_tc_path(C, ht, wd, dtype)) {
    if (!workspace || workspace_bytes < db200_corr_volume_workspace_bytes(N, C, ht, wd, num_levels, dtype)) return DB200_ERR_WORKSPACE;
    return corr_volume_pyramid_tc(static_cast<const __half*>(fmap1), static_cast<const __half*>(fmap2), levels_out, N, C, ht, wd,
                                  num_levels, workspace, st, tiled_mask);
  }
  if (tiled_mask != 0) return DB200_ERR_ARG;                       // tiled levels exist on the tcgen05 fp16 path only
  if (dtype == DB200_F32 && use_tc_f32_path(C, ht, wd, num_levels)) {
    if (!workspace || workspace_bytes < db200_corr_volume_workspace_bytes(N, C, ht, wd, num_levels, dtype)) return DB200_ERR_WORKSPACE;
    return corr_volume_pyramid_tc_f32(static_cast<const float*>(fmap1), static_cast<const float*>(fmap2), levels_out, N, C, ht, wd,
                                      num_levels, workspace, st);
  }
  const size_t es = dtype == DB200_F16 ? 2 : 4;
  const int M = ht * wd;
  const void* src = fmap2;
  char* wsp = static_cast<char*>(workspace);
  int h = ht, w = wd;
  for (int l = 0; l < num_levels; l++) {
    if (l > 0) {
      if ((h >> 1) == 0 || (w >> 1) == 0) return DB200_ERR_ARG;
      const size_t cnt = size_t(N) * C * (h >> 1) * (w >> 1);
      const int blocks = int((cnt + 255) / 256 < 4096 ? (cnt + 255) / 256 : 4096);
      if (dtype == DB200_F16)
        pool_fmap_kernel<__half><<<blocks, 256, 0, st>>>(static_cast<const __half*>(src), reinterpret_cast<__half*>(wsp), N * C, h, w);
      else
        pool_fmap_kernel<float><<<blocks, 256, 0, st>>>(static_cast<const float*>(src), reinterpret_cast<float*>(wsp), N * C, h, w);
      DB200_COUNT(1);
      src = wsp;
      wsp += ((cnt * es + 255) / 256) * 256;
      h >>= 1; w >>= 1;
    }
    const int N2 = h * w;
    if (dtype == DB200_F16) {
      dim3 grid(cdiv(N2, 128), cdiv(M, 64), N);
      corr_gemm_f16_kernel<<<grid, 256, 0, st>>>(static_cast<const __half*>(fmap1), static_cast<const __half*>(src),
                                                 static_cast<__half*>(levels_out[l]), C, M, N2);
    } else {
      dim3 grid(cdiv(N2, 64), cdiv(M, 64), N);
      corr_gemm_f32_kernel<<<grid, 256, 0, st>>>(static_cast<const float*>(fmap1), static_cast<const float*>(src),
                                                 static_cast<float*>(levels_out[l]), C, M, N2);
    }
  }
  DB200_COUNT(num_levels - 1);
  DB200_CHECK_LAUNCH();
  return DB200_OK;
}

extern "C" int db200_corr_volume_pyramid(const void* fmap1, const void* fmap2, void* const* levels_out, int N, int C, int ht,
                                         int wd, int num_levels, int dtype, void* workspace, size_t workspace_bytes,
                                         db200_stream_t stream) {
  return volume_pyramid(fmap1, fmap2, levels_out, N, C, ht, wd, num_levels, dtype, workspace, workspace_bytes, 0u, stream);
}

extern "C" int db200_corr_volume_pyramid_tiled(const void* fmap1, const void* fmap2, void* const* levels_out, int N, int C, int ht,
                                               int wd, int num_levels, int dtype, void* workspace, size_t workspace_bytes,
                                               unsigned tiled_mask, db200_stream_t stream) {
  if (tiled_mask != 0 && (tiled_mask & ~db200_corr_tiled_levels(C, ht, wd, num_levels, dtype)) != 0) return DB200_ERR_ARG;
  return volume_pyramid(fmap1, fmap2, levels_out, N, C, ht, wd, num_levels, dtype, workspace, workspace_bytes, tiled_mask, stream);
}
