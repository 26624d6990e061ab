// altcorr.cu — low-memory correlation: dot products against gathered fmap2 windows, fwd + bwd.
// Replaces /root/reference/src/altcorr_kernel.cu:27-355 (AltCorrBlock / CorrLayer, corr.py:74-139).
//
// fmap1[B,H1,W1,C], fmap2[B,H2,W2,C] channels-last fp32; coords[B,S,H1,W1,2]; corr[B,S,(2r+1)^2,H1,W1].
// The reference runs ONE warp per 4x8 pixel tile, 64 __syncthreads-separated gathers per 32-channel slab
// and a global read-modify-write per window tap.  Here one warp owns one source pixel: lane l keeps 4
// channels (c = 4l..4l+3 per 128-channel slab) of fmap1 in registers, every window tap is one coalesced
// 512-byte read of fmap2 (L1/L2-resident: neighbouring pixels' windows overlap), the (2r+2)^2 dot
// products are reduced with a transposing shuffle butterfly, bilinear taps are combined in registers
// and a CTA covers 64 consecutive x so every output channel row is stored as one coalesced line.
// L2/compute bound (201 MFLOP per edge for 4 levels), not HBM bound.
#include "common.cuh"

namespace db200 {

constexpr int kAltTX = 64;
constexpr int kAltWarps = 8;

// sum `v` over the warp for NV (<=64, padded) values held as v[0..63] per lane; afterwards lane l holds the
// totals of value indices l and l+32 in v[0], v[1]... implemented as a plain xor-butterfly per value for
// clarity in v1 (5 shuffles per value).
__device__ __forceinline__ float warp_allsum(float v) { return warp_sum(v); }

// grid: (ceil(W1/TX), H1, B*S); block 256
__global__ void __launch_bounds__(kAltWarps * 32)
altcorr_fwd_kernel(const float* __restrict__ fmap1, const float* __restrict__ fmap2, const float* __restrict__ coords,
                   float* __restrict__ corr, int B, int S, int H1, int W1, int H2, int W2, int C, int r) {
  extern __shared__ __align__(16) float smem[];
  const int rd = 2 * r + 1, wn = rd + 1, nch = rd * rd;
  float* sout = smem;                            // [nch][TX]
  float* sdot = smem + size_t(nch) * kAltTX;     // [warps][wn*wn]
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int x_base = blockIdx.x * kAltTX, y = blockIdx.y;
  const int b = blockIdx.z / S, s = blockIdx.z - b * S;
  const int npx = min(kAltTX, W1 - x_base);
  float* dots = sdot + warp * wn * wn;
  for (int px = warp; px < npx; px += kAltWarps) {
    const int x = x_base + px;
    const float* cp = coords + (((size_t(b) * S + s) * H1 + y) * W1 + x) * 2;
    const float x2 = cp[0], y2 = cp[1];
    const float fx = floorf(x2), fy = floorf(y2);
    const float dx = x2 - fx, dy = y2 - fy;
    const int wb = static_cast<int>(fx) - r, hb = static_cast<int>(fy) - r;
    const float* f1 = fmap1 + ((size_t(b) * H1 + y) * W1 + x) * C;
    const float* f2b = fmap2 + size_t(b) * H2 * W2 * C;
    for (int t = 0; t < wn * wn; t++) {
      const int iy = t / wn, ix = t - iy * wn;
      const int h2 = hb + iy, w2 = wb + ix;
      float acc = 0.0f;
      if (h2 >= 0 && h2 < H2 && w2 >= 0 && w2 < W2) {
        const float* f2 = f2b + (size_t(h2) * W2 + w2) * C;
        for (int c = lane * 4; c < C; c += 128) {
          if (c + 3 < C) {
            const float4 a = *reinterpret_cast<const float4*>(f1 + c);
            const float4 q = *reinterpret_cast<const float4*>(f2 + c);
            acc = fmaf(a.x, q.x, acc); acc = fmaf(a.y, q.y, acc); acc = fmaf(a.z, q.z, acc); acc = fmaf(a.w, q.w, acc);
          } else {
            for (int k = c; k < C; k++) acc = fmaf(f1[k], f2[k], acc);
          }
        }
        acc = warp_allsum(acc);
      }
      if (lane == 0) dots[iy * wn + ix] = acc;
    }
    __syncwarp();
    // bilinear taps (altcorr_kernel.cu:107-140): channel o = iy + rd*ix gets
    //   (1-dy)(1-dx) s[iy][ix] + (1-dy)dx s[iy][ix+1] + dy(1-dx) s[iy+1][ix] + dy dx s[iy+1][ix+1]
    const float wse = (1 - dy) * (1 - dx), wsw = (1 - dy) * dx, wne = dy * (1 - dx), wnw = dy * dx;
    for (int o = lane; o < nch; o += 32) {
      const int ix = o / rd, iy = o - ix * rd;
      float v = dots[iy * wn + ix] * wse;
      v += dots[iy * wn + ix + 1] * wsw;
      v += dots[(iy + 1) * wn + ix] * wne;
      v += dots[(iy + 1) * wn + ix + 1] * wnw;
      sout[o * kAltTX + px] = v;
    }
    __syncwarp();
  }
  __syncthreads();
  const size_t plane = size_t(H1) * W1;
  float* outp = corr + (size_t(b) * S + s) * nch * plane + size_t(y) * W1 + x_base;
  for (int idx = threadIdx.x; idx < nch * kAltTX; idx += blockDim.x) {
    const int ch = idx / kAltTX, px = idx - ch * kAltTX;
    if (px < npx) outp[size_t(ch) * plane + px] = sout[idx];
  }
}

// Backward (altcorr_kernel.cu:152-286): for each source pixel and window tap,
//   g = 4-corner combination of corr_grad;  fmap1_grad[pix] += g * fmap2[tap];  fmap2_grad[tap] += g * fmap1[pix]
// fmap1_grad is owned by the pixel's warp (plain store); fmap2_grad uses red.global.add (as the reference does).
// grid: ceil(B*H1*W1 / 8); block 256.  fmap2_grad must be zero on entry (done by the ABI wrapper).
__global__ void __launch_bounds__(256)
altcorr_bwd_kernel(const float* __restrict__ fmap1, const float* __restrict__ fmap2, const float* __restrict__ coords,
                   const float* __restrict__ corr_grad, float* __restrict__ fmap1_grad, float* __restrict__ fmap2_grad,
                   int B, int S, int H1, int W1, int H2, int W2, int C, int r, const unsigned* __restrict__ cond_flags,
                   const int* __restrict__ cond_status) {
  if (cond_flags && !((*cond_flags & 2u) || cond_status[1] != 0)) return;     // only when the tensor-core kernel handed the call over
  extern __shared__ __align__(16) float smem[];
  const int rd = 2 * r + 1, wn = rd + 1, nch = rd * rd;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float* sg = smem + warp * nch;
  const size_t plane = size_t(H1) * W1;
  const size_t pix = size_t(blockIdx.x) * 8 + warp;
  if (pix >= size_t(B) * plane) return;
  const int b = int(pix / plane);
  const size_t yx = pix - size_t(b) * plane;
  const float* f1 = fmap1 + pix * C;
  float* g1 = fmap1_grad + pix * C;
  const float* f2b = fmap2 + size_t(b) * H2 * W2 * C;
  float* g2b = fmap2_grad + size_t(b) * H2 * W2 * C;
  for (int c0 = 0; c0 < C; c0 += 128) {   // 128-channel slabs: 4 channels per lane
    const int c = c0 + lane * 4;
    float a[4] = {0, 0, 0, 0}, acc[4] = {0, 0, 0, 0};
    for (int k = 0; k < 4; k++) if (c + k < C) a[k] = f1[c + k];
    for (int s = 0; s < S; s++) {
      const float* cp = coords + ((size_t(b) * S + s) * plane + yx) * 2;
      const float x2 = cp[0], y2 = cp[1];
      const float fx = floorf(x2), fy = floorf(y2);
      const float dx = x2 - fx, dy = y2 - fy;
      const int wb = static_cast<int>(fx) - r, hb = static_cast<int>(fy) - r;
      __syncwarp();
      for (int o = lane; o < nch; o += 32) sg[o] = corr_grad[((size_t(b) * S + s) * nch + o) * plane + yx];
      __syncwarp();
      for (int t = 0; t < wn * wn; t++) {
        const int iy = t / wn, ix = t - iy * wn;
        const int h2 = hb + iy, w2 = wb + ix;
        if (h2 < 0 || h2 >= H2 || w2 < 0 || w2 >= W2) continue;
        float g = 0.0f;
        if (iy > 0 && ix > 0) g += sg[(iy - 1) + rd * (ix - 1)] * dy * dx;
        if (iy > 0 && ix < rd) g += sg[(iy - 1) + rd * ix] * dy * (1 - dx);
        if (iy < rd && ix > 0) g += sg[iy + rd * (ix - 1)] * (1 - dy) * dx;
        if (iy < rd && ix < rd) g += sg[iy + rd * ix] * (1 - dy) * (1 - dx);
        const float* f2 = f2b + (size_t(h2) * W2 + w2) * C;
        float* g2 = g2b + (size_t(h2) * W2 + w2) * C;
        for (int k = 0; k < 4; k++)
          if (c + k < C) {
            acc[k] = fmaf(g, f2[c + k], acc[k]);
            atomicAdd(g2 + c + k, g * a[k]);
          }
      }
    }
    for (int k = 0; k < 4; k++) if (c + k < C) g1[c + k] = acc[k];
  }
}

__global__ void zero_if_kernel(float4* __restrict__ p, size_t n4, const unsigned* __restrict__ cond_flags,
                               const int* __restrict__ cond_status) {
  if (!((*cond_flags & 2u) || cond_status[1] != 0)) return;
  for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n4; i += size_t(gridDim.x) * blockDim.x)
    p[i] = make_float4(0.f, 0.f, 0.f, 0.f);
}

int altcorr_bwd_simt_conditional(const float* fmap1, const float* fmap2, const float* coords, const float* corr_grad,
                                 float* fmap1_grad, float* fmap2_grad, int B, int S, int H1, int W1, int H2, int W2, int C,
                                 int radius, const unsigned* flags, const int* status, cudaStream_t st) {
  const size_t n4 = size_t(B) * H2 * W2 * C / 4;
  zero_if_kernel<<<kNumSMs * 4, 256, 0, st>>>(reinterpret_cast<float4*>(fmap2_grad), n4, flags, status);
  DB200_CHECK_LAUNCH();
  const int rd = 2 * radius + 1;
  const size_t npix = size_t(B) * H1 * W1;
  altcorr_bwd_kernel<<<unsigned((npix + 7) / 8), 256, size_t(8) * rd * rd * sizeof(float), st>>>(
      fmap1, fmap2, coords, corr_grad, fmap1_grad, fmap2_grad, B, S, H1, W1, H2, W2, C, radius, flags, status);
  DB200_CHECK_LAUNCH();
  return DB200_OK;
}

}  // namespace db200

using namespace db200;

extern "C" int db200_altcorr_forward(const float* fmap1, const float* fmap2, const float* coords, float* corr, int B, int S,
                                     int H1, int W1, int H2, int W2, int C, int radius, db200_stream_t stream) {
  if (!fmap1 || !fmap2 || !coords || !corr || B < 0 || S < 0 || H1 <= 0 || W1 <= 0 || H2 <= 0 || W2 <= 0 || C <= 0 ||
      radius < 0 || radius > 7 || (C % 4) != 0)
    return DB200_ERR_ARG;
  if (B == 0 || S == 0) return DB200_OK;
  const int rd = 2 * radius + 1, wn = rd + 1;
  const size_t smem = (size_t(rd) * rd * kAltTX + size_t(kAltWarps) * wn * wn) * sizeof(float);
  if (smem > 48 * 1024 &&
      cudaFuncSetAttribute(altcorr_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)) != cudaSuccess)
    return DB200_ERR_CUDA;
  dim3 grid(cdiv(W1, kAltTX), H1, B * S);
  altcorr_fwd_kernel<<<grid, kAltWarps * 32, smem, static_cast<cudaStream_t>(stream)>>>(fmap1, fmap2, coords, corr, B, S, H1,
                                                                                     W1, H2, W2, C, radius);
  DB200_CHECK_LAUNCH();
  return DB200_OK;
}

extern "C" int db200_altcorr_backward(const float* fmap1, const float* fmap2, const float* coords, const float* corr_grad,
                                      float* fmap1_grad, float* fmap2_grad, float* coords_grad, int B, int S, int H1, int W1,
                                      int H2, int W2, int C, int radius, db200_stream_t stream) {
  if (!fmap1 || !fmap2 || !coords || !corr_grad || !fmap1_grad || !fmap2_grad || B < 0 || S < 0 || H1 <= 0 || W1 <= 0 ||
      H2 <= 0 || W2 <= 0 || C <= 0 || radius < 0 || radius > 7)
    return DB200_ERR_ARG;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (B == 0) return DB200_OK;
  DB200_CUDA(cudaMemsetAsync(fmap2_grad, 0, sizeof(float) * size_t(B) * H2 * W2 * C, st));
  if (coords_grad && S > 0 &&
      cudaMemsetAsync(coords_grad, 0, sizeof(float) * size_t(B) * S * H1 * W1 * 2, st) != cudaSuccess)
    return DB200_ERR_CUDA;
  const int rd = 2 * radius + 1;
  const size_t npix = size_t(B) * H1 * W1;
  altcorr_bwd_kernel<<<unsigned((npix + 7) / 8), 256, size_t(8) * rd * rd * sizeof(float), st>>>(
      fmap1, fmap2, coords, corr_grad, fmap1_grad, fmap2_grad, B, S, H1, W1, H2, W2, C, radius, nullptr, nullptr);
  DB200_CHECK_LAUNCH();
  return DB200_OK;
}
