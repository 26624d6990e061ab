// geom.cu — frame_distance, iproj, depth_filter.
// Replaces /root/reference/src/droid_kernels.cu:518-657 (frame_distance_kernel), :779-850 (iproj_kernel),
// :661-775 (depth_filter_kernel).
//
// frame_distance feeds thresholds / argsort on the host (factor_graph.py:327-379), so its fp32 result is
// reproduced BIT-EXACTLY: the same per-thread strided accumulation (256 virtual threads), the same
// 256->128->64->32->...->1 reduction tree and the same FMA contraction as the reference binary (pattern
// documented in oracle/oracle.cpp, read from the reference's SASS).  Intrinsics (__fmaf_rn/__fmul_rn/
// __fadd_rn) pin the contraction so that nvcc cannot re-associate.  One CTA per pair, 256 threads; the
// disparity map of frame i (12 KB) is L1/L2 resident across the up-to-t pairs that share it.
#include <cstdlib>
#include "common.cuh"

namespace db200 {

__device__ __forceinline__ float fm(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float fa(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ float ff(float a, float b, float c) { return __fmaf_rn(a, b, c); }

__device__ __forceinline__ void fd_rot(const float* q, const float* X, float* Y) {
  float uv0 = ff(q[1], X[2], -fm(q[2], X[1])); uv0 = fa(uv0, uv0);
  float uv1 = ff(q[2], X[0], -fm(q[0], X[2])); uv1 = fa(uv1, uv1);
  float uv2 = ff(q[0], X[1], -fm(q[1], X[0])); uv2 = fa(uv2, uv2);
  Y[0] = fa(ff(q[3], uv0, X[0]), ff(q[1], uv2, -fm(q[2], uv1)));
  Y[1] = fa(ff(q[3], uv1, X[1]), ff(q[2], uv0, -fm(q[0], uv2)));
  Y[2] = fa(ff(q[3], uv2, X[2]), ff(q[0], uv1, -fm(q[1], uv0)));
}

__device__ __forceinline__ void fd_rel(const float* ti, const float* qi, const float* tj, const float* qj, float* tij,
                                       float* q) {
  q[0] = ff(qj[2], qi[1], ff(-qj[1], qi[2], ff(qj[0], qi[3], -fm(qj[3], qi[0]))));
  q[1] = ff(qj[0], qi[2], ff(-qj[2], qi[0], ff(qj[1], qi[3], -fm(qj[3], qi[1]))));
  q[2] = ff(qj[1], qi[0], ff(-qj[0], qi[1], ff(qj[2], qi[3], -fm(qj[3], qi[2]))));
  q[3] = ff(qj[2], qi[2], ff(qj[1], qi[1], ff(qj[3], qi[3], fm(qj[0], qi[0]))));
  float Y[3];
  fd_rot(q, ti, Y);
  tij[0] = fa(tj[0], -Y[0]); tij[1] = fa(tj[1], -Y[1]); tij[2] = fa(tj[2], -Y[2]);
}

// tree of blockReduce (droid_kernels.cu:36-55) for thread 0's result
__device__ __forceinline__ float fd_block_reduce(float v, float* s) {
  const int t = threadIdx.x;
  s[t] = v;
  __syncthreads();
  if (t < 128) s[t] = fa(s[t], s[t + 128]);
  __syncthreads();
  if (t < 64) s[t] = fa(s[t], s[t + 64]);
  __syncthreads();
  float r = 0.0f;
  if (t < 32) {
    r = fa(s[t], s[t + 32]);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) r = fa(r, __shfl_down_sync(0xffffffffu, r, o));
  }
  __syncthreads();
  return r;  // valid in thread 0
}

__global__ void __launch_bounds__(256)
frame_distance_kernel(const float* __restrict__ poses, const float* __restrict__ disps, const float* __restrict__ intr,
                      const int64_t* __restrict__ ii, const int64_t* __restrict__ jj, int K, int n_poses, int B, int ht,
                      int wd, float beta, float* __restrict__ dist) {
  __shared__ float sred[256];
  const int b = blockIdx.x;
  const int ix = int(ii[b]), jx = int(jj[b]);
  if (ix < 0 || jx < 0 || ix >= n_poses || jx >= n_poses || ix >= B) {   // reference: undefined behaviour
    if (threadIdx.x == 0) dist[b] = __int_as_float(0x7fc00000);
    return;
  }
  const float fx = intr[0], fy = intr[1], cx = intr[2], cy = intr[3];
  float tij[3], q[4];
  fd_rel(poses + 7 * ix, poses + 7 * ix + 3, poses + 7 * jx, poses + 7 * jx + 3, tij, q);
  const float omb = fa(1.0f, -beta);
  const int HW = ht * wd;
  const float* dsp = disps + size_t(ix) * HW;
  float accum = 0.0f, valid = 0.0f, total = 0.0f;
  for (int k = threadIdx.x; k < HW; k += 256) {
    const int i = k / wd, j = k - i * wd;
    const float u = float(j), v = float(i);
    const float X0 = __fdiv_rn(fa(u, -cx), fx), X1 = __fdiv_rn(fa(v, -cy), fy), d = dsp[k];
    float uv0 = ff(-q[2], X1, q[1]); uv0 = fa(uv0, uv0);
    float uv1 = ff(q[2], X0, -q[0]); uv1 = fa(uv1, uv1);
    float uv2 = ff(q[0], X1, -fm(q[1], X0)); uv2 = fa(uv2, uv2);
    float Y0 = fa(ff(q[3], uv0, X0), ff(q[1], uv2, -fm(q[2], uv1)));
    float Y1 = fa(ff(q[3], uv1, X1), ff(q[2], uv0, -fm(q[0], uv2)));
    float Y2 = fa(ff(q[3], uv2, 1.0f), ff(q[0], uv1, -fm(q[1], uv0)));
    const float p0 = fm(tij[0], d), p1 = fm(tij[1], d), p2 = fm(tij[2], d);
    Y0 = fa(Y0, p0); Y1 = fa(Y1, p1); Y2 = fa(Y2, p2);
    float du = fa(ff(fx, __fdiv_rn(Y0, Y2), cx), -u);
    float dv = fa(ff(fy, __fdiv_rn(Y1, Y2), cy), -v);
    float dd = __fsqrt_rn(ff(du, du, fm(dv, dv)));
    total = fa(total, beta);
    if (Y2 > 0.25f) { accum = ff(dd, beta, accum); valid = fa(valid, beta); }
    const float Z0 = fa(p0, X0), Z1 = fa(p1, X1), Z2 = fa(p2, 1.0f);
    du = fa(ff(fx, __fdiv_rn(Z0, Z2), cx), -u);
    dv = fa(ff(fy, __fdiv_rn(Z1, Z2), cy), -v);
    dd = __fsqrt_rn(ff(du, du, fm(dv, dv)));
    total = fa(total, omb);
    if (Z2 > 0.25f) { accum = ff(omb, dd, accum); valid = fa(valid, omb); }
  }
  const float a = fd_block_reduce(accum, sred);
  const float tt = fd_block_reduce(total, sred);
  const float vv = fd_block_reduce(valid, sred);
  if (threadIdx.x == 0) dist[b] = (double(vv) / (double(tt) + 1e-8) < 0.75) ? 1000.0f : __fdiv_rn(a, vv);
}

// Batched variant for maps of at most NPIX*256 pixels: one CTA walks kFdBatch consecutive pairs.  Everything that does
// not depend on the pair is hoisted out of the pair loop — the two IEEE divisions of the back-projection per pixel, the
// pixel coordinates, the `total` sum and its reduction (it is the same constant for every pair), and the disparity
// map while consecutive pairs share their source frame (meshgrid order) — and `accum`/`valid` go through ONE pass of the
// reference's reduction tree instead of three.  Every value is produced by the same operations in the same order as in
// frame_distance_kernel above, so the results are bit-identical.
constexpr int kFdBatch = 8;

template <int NPIX>
__global__ void __launch_bounds__(256, 3)
frame_distance_batch_kernel(const float* __restrict__ poses, const float* __restrict__ disps, const float* __restrict__ intr,
                            const int64_t* __restrict__ ii, const int64_t* __restrict__ jj, int K, int n_poses, int B, int ht,
                            int wd, float beta, float* __restrict__ dist) {
  __shared__ float sa[256], sv[256];
  const int t = threadIdx.x;
  {                                                      // a batch without a single valid pair (callers mark pairs whose
    bool any = false;                                    // result they will discard with a negative index) costs nothing
    for (int g = 0; g < kFdBatch; g++) {
      const int b = blockIdx.x * kFdBatch + g;
      if (b < K) {
        const long long ix = ii[b], jx = jj[b];
        any |= !(ix < 0 || jx < 0 || ix >= n_poses || jx >= n_poses || ix >= B);
      }
    }
    if (!any) {
      if (t < kFdBatch && blockIdx.x * kFdBatch + t < K) dist[blockIdx.x * kFdBatch + t] = __int_as_float(0x7fc00000);
      return;
    }
  }
  const float fx = intr[0], fy = intr[1], cx = intr[2], cy = intr[3];
  const float omb = fa(1.0f, -beta);
  const int HW = ht * wd;
  float X0[NPIX], X1[NPIX], U[NPIX], V[NPIX], D[NPIX];
  float total = 0.0f;
#pragma unroll
  for (int m = 0; m < NPIX; m++) {
    const int k = t + 256 * m;
    const int i = k / wd, j = k - i * wd;
    U[m] = float(j); V[m] = float(i);
    X0[m] = __fdiv_rn(fa(U[m], -cx), fx);
    X1[m] = __fdiv_rn(fa(V[m], -cy), fy);
    D[m] = 0.0f;
    if (k < HW) { total = fa(total, beta); total = fa(total, omb); }
  }
  const float tt = fd_block_reduce(total, sa);          // thread 0 holds the reference's `total` (pair independent)
  int cur_ix = -1;
  const int b0 = blockIdx.x * kFdBatch;
  for (int g = 0; g < kFdBatch; g++) {
    const int b = b0 + g;
    if (b >= K) break;
    const int ix = int(ii[b]), jx = int(jj[b]);
    if (ix < 0 || jx < 0 || ix >= n_poses || jx >= n_poses || ix >= B) {
      if (t == 0) dist[b] = __int_as_float(0x7fc00000);
      continue;
    }
    if (ix != cur_ix) {
      const float* dsp = disps + size_t(ix) * HW;
#pragma unroll
      for (int m = 0; m < NPIX; m++) { const int k = t + 256 * m; D[m] = (k < HW) ? dsp[k] : 0.0f; }
      cur_ix = ix;
    }
    float tij[3], q[4];
    fd_rel(poses + 7 * ix, poses + 7 * ix + 3, poses + 7 * jx, poses + 7 * jx + 3, tij, q);
    float accum = 0.0f, valid = 0.0f;
#pragma unroll
    for (int m = 0; m < NPIX; m++) {
      if (t + 256 * m < HW) {
        const float x0 = X0[m], x1 = X1[m], u = U[m], v = V[m], d = D[m];
        float uv0 = ff(-q[2], x1, q[1]); uv0 = fa(uv0, uv0);
        float uv1 = ff(q[2], x0, -q[0]); uv1 = fa(uv1, uv1);
        float uv2 = ff(q[0], x1, -fm(q[1], x0)); uv2 = fa(uv2, uv2);
        float Y0 = fa(ff(q[3], uv0, x0), ff(q[1], uv2, -fm(q[2], uv1)));
        float Y1 = fa(ff(q[3], uv1, x1), ff(q[2], uv0, -fm(q[0], uv2)));
        float Y2 = fa(ff(q[3], uv2, 1.0f), ff(q[0], uv1, -fm(q[1], uv0)));
        const float p0 = fm(tij[0], d), p1 = fm(tij[1], d), p2 = fm(tij[2], d);
        Y0 = fa(Y0, p0); Y1 = fa(Y1, p1); Y2 = fa(Y2, p2);
        float du = fa(ff(fx, __fdiv_rn(Y0, Y2), cx), -u);
        float dv = fa(ff(fy, __fdiv_rn(Y1, Y2), cy), -v);
        float dd = __fsqrt_rn(ff(du, du, fm(dv, dv)));
        if (Y2 > 0.25f) { accum = ff(dd, beta, accum); valid = fa(valid, beta); }
        const float Z0 = fa(p0, x0), Z1 = fa(p1, x1), Z2 = fa(p2, 1.0f);
        du = fa(ff(fx, __fdiv_rn(Z0, Z2), cx), -u);
        dv = fa(ff(fy, __fdiv_rn(Z1, Z2), cy), -v);
        dd = __fsqrt_rn(ff(du, du, fm(dv, dv)));
        if (Z2 > 0.25f) { accum = ff(omb, dd, accum); valid = fa(valid, omb); }
      }
    }
    // both quantities through the reference's tree (256 -> 128 -> 64 -> 32 -> shuffles) in one pass
    sa[t] = accum; sv[t] = valid;
    __syncthreads();
    if (t < 128) { sa[t] = fa(sa[t], sa[t + 128]); sv[t] = fa(sv[t], sv[t + 128]); }
    __syncthreads();
    if (t < 64) { sa[t] = fa(sa[t], sa[t + 64]); sv[t] = fa(sv[t], sv[t + 64]); }
    __syncthreads();
    if (t < 32) {
      float a = fa(sa[t], sa[t + 32]), vv = fa(sv[t], sv[t + 32]);
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        a = fa(a, __shfl_down_sync(0xffffffffu, a, o));
        vv = fa(vv, __shfl_down_sync(0xffffffffu, vv, o));
      }
      if (t == 0) dist[b] = (double(vv) / (double(tt) + 1e-8) < 0.75) ? 1000.0f : __fdiv_rn(a, vv);
    }
    __syncthreads();
  }
}

// points[n,i,j,:] = (G_n X)_{xyz} / X_w   (droid_kernels.cu:779-850)
__global__ void iproj_kernel(const float* __restrict__ poses, const float* __restrict__ disps,
                             const float* __restrict__ intr, int num, int ht, int wd, float* __restrict__ points) {
  const int n = blockIdx.y;
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  const int HW = ht * wd;
  if (k >= HW) return;
  const float fx = intr[0], fy = intr[1], cx = intr[2], cy = intr[3];
  const float* p = poses + 7 * n;
  const float d = disps[size_t(n) * HW + k];
  float X[3] = {(float(k % wd) - cx) / fx, (float(k / wd) - cy) / fy, 1.0f}, Y[3];
  act_so3(p + 3, X, Y);
  float* o = points + (size_t(n) * HW + k) * 3;
  o[0] = (Y[0] + d * p[0]) / d;
  o[1] = (Y[1] + d * p[1]) / d;
  o[2] = (Y[2] + d * p[2]) / d;
}

// multi-view depth consistency count over the 6 temporal neighbours (droid_kernels.cu:661-775).
// One thread owns one (frame, pixel) and loops over the neighbours: no atomics.
__global__ void depth_filter_kernel(const float* __restrict__ poses, const float* __restrict__ disps,
                                    const float* __restrict__ intr, const int64_t* __restrict__ inds,
                                    const float* __restrict__ thresh, int num, int ht, int wd,
                                    float* __restrict__ counter) {
  const int b = blockIdx.y;
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  const int HW = ht * wd;
  if (k >= HW) return;
  const float fx = intr[0], fy = intr[1], cx = intr[2], cy = intr[3];
  const int ix = int(inds[b]);
  const float t = thresh[b];
  const int i = k / wd, j = k - i * wd;
  const float di = disps[size_t(ix) * HW + k];
  const float X[3] = {(float(j) - cx) / fx, (float(i) - cy) / fy, 1.0f};
  float count = 0.0f;
  for (int nb = 0; nb < 6; nb++) {
    const int jx = (nb < 3) ? ix - nb - 1 : ix + nb;
    if (jx < 0 || jx >= num) continue;
    float tij[3], qij[4], Y[3];
    rel_se3(poses + 7 * ix, poses + 7 * jx, tij, qij);
    act_so3(qij, X, Y);
    Y[0] += di * tij[0]; Y[1] += di * tij[1]; Y[2] += di * tij[2];
    const float uj = fx * (Y[0] / Y[2]) + cx, vj = fy * (Y[1] / Y[2]) + cy;
    const float dj = di / Y[2];
    const int u0 = static_cast<int>(floorf(uj)), v0 = static_cast<int>(floorf(vj));
    if (u0 >= 0 && v0 >= 0 && u0 < wd - 1 && v0 < ht - 1) {
      const float* dn = disps + size_t(jx) * HW;
      const double idj = 1.0 / double(dj), tt = double(t);
      if (fabs(idj - 1.0 / double(dn[v0 * wd + u0])) < tt || fabs(idj - 1.0 / double(dn[v0 * wd + u0 + 1])) < tt ||
          fabs(idj - 1.0 / double(dn[(v0 + 1) * wd + u0])) < tt || fabs(idj - 1.0 / double(dn[(v0 + 1) * wd + u0 + 1])) < tt)
        count += 1.0f;
    }
  }
  counter[size_t(b) * HW + k] = count;
}

// projmap (droid_kernels.cu:427-516, 1463-1490): coords[num,ht,wd,3] (third channel stays zero, as the reference's
// torch::zeros leaves it), valid[num,ht,wd,1].  No stereo special case in the reference kernel.
__global__ void projmap_kernel(const float* __restrict__ poses, const float* __restrict__ disps, const float* __restrict__ intr,
                               const int64_t* __restrict__ ii, const int64_t* __restrict__ jj, int n_poses, int B, int ht, int wd,
                               float* __restrict__ coords, float* __restrict__ valid) {
  const int e = blockIdx.y;
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  const int HW = ht * wd;
  if (k >= HW) return;
  const int ix = int(ii[e]), jx = int(jj[e]);
  float* c = coords + (size_t(e) * HW + k) * 3;
  if (ix < 0 || jx < 0 || ix >= n_poses || jx >= n_poses || ix >= B) {       // reference: undefined behaviour
    c[0] = c[1] = c[2] = valid[size_t(e) * HW + k] = __int_as_float(0x7fc00000);
    return;
  }
  const float fx = intr[0], fy = intr[1], cx = intr[2], cy = intr[3];
  float tij[3], qij[4], Y[3];
  rel_se3(poses + 7 * ix, poses + 7 * jx, tij, qij);
  const int i = k / wd, j = k - i * wd;
  const float u = float(j), v = float(i);
  const float d = disps[size_t(ix) * HW + k];
  const float X[3] = {(u - cx) / fx, (v - cy) / fy, 1.0f};
  act_so3(qij, X, Y);
  Y[0] += d * tij[0]; Y[1] += d * tij[1]; Y[2] += d * tij[2];
  float ou = u, ov = v;
  if (Y[2] > 0.01f) { ou = fx * (Y[0] / Y[2]) + cx; ov = fy * (Y[1] / Y[2]) + cy; }
  c[0] = ou; c[1] = ov; c[2] = 0.0f;
  valid[size_t(e) * HW + k] = (Y[2] > 0.25f) ? 1.0f : 0.0f;
}

// DepthVideo.reproject (depth_video.py:150-158) = projective_ops.projective_transform without Jacobians
// (projective_ops.py:96-125), fused with the motion features of FactorGraph.update (factor_graph.py:203-205):
//   Gij = Gj * Gi^-1 on quaternions normalised first (lietorch's SO3 constructor does, so3.h:31-37), or the fixed stereo
//   baseline when ii == jj (:105);  X1 = Gij * ((x-cx_i)/fx_i, (y-cy_i)/fy_i, 1, d);  Z <- 1 where Z < 0.5*MIN_DEPTH;
//   coords = (fx_j X/Z + cx_j, fy_j Y/Z + cy_j);  valid = X1_z > MIN_DEPTH (= 0.2; X0_z = 1 always passes);
//   motn = clamp(cat(coords - coords0, target - coords), -64, 64) laid out [N,4,ht,wd].
// One thread per (edge, pixel); everything is read once and written once (HBM-bound: 4 B in, 12 B out per pixel, or
// 12 B in, 28 B out with the motion features).
__device__ __forceinline__ void unit_quat(const float* q, float* o) {
  const float n = sqrtf(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
  o[0] = q[0] / n; o[1] = q[1] / n; o[2] = q[2] / n; o[3] = q[3] / n;
}

__global__ void reproject_kernel(const float* __restrict__ poses, const float* __restrict__ disps, const float* __restrict__ intr,
                                 int intr_stride, const int64_t* __restrict__ ii, const int64_t* __restrict__ jj, int n_poses,
                                 int B, int ht, int wd, float* __restrict__ coords, float* __restrict__ valid,
                                 const float* __restrict__ target, float* __restrict__ motn) {
  const int e = blockIdx.y;
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  const int HW = ht * wd;
  if (k >= HW) return;
  const int ix = int(ii[e]), jx = int(jj[e]);
  const size_t o = size_t(e) * HW + k;
  if (ix < 0 || jx < 0 || ix >= n_poses || jx >= n_poses || ix >= B) {
    const float nan = __int_as_float(0x7fc00000);
    if (coords) { coords[2 * o] = nan; coords[2 * o + 1] = nan; }
    if (valid) valid[o] = 0.0f;
    if (motn) for (int c = 0; c < 4; c++) motn[(size_t(e) * 4 + c) * HW + k] = nan;
    return;
  }
  const float* Ki = intr + size_t(ix) * intr_stride;
  const float* Kj = intr + size_t(jx) * intr_stride;
  float tij[3], qij[4];
  if (ix == jx) {
    tij[0] = -0.1f; tij[1] = 0.0f; tij[2] = 0.0f;
    qij[0] = 0.0f; qij[1] = 0.0f; qij[2] = 0.0f; qij[3] = 1.0f;
  } else {
    float pi[7], pj[7];
    for (int c = 0; c < 3; c++) { pi[c] = poses[7 * ix + c]; pj[c] = poses[7 * jx + c]; }
    unit_quat(poses + 7 * ix + 3, pi + 3);
    unit_quat(poses + 7 * jx + 3, pj + 3);
    rel_se3(pi, pj, tij, qij);
  }
  const int i = k / wd, j = k - i * wd;
  const float u = float(j), v = float(i);
  const float d = disps[size_t(ix) * HW + k];
  const float X[3] = {(u - Ki[2]) / Ki[0], (v - Ki[3]) / Ki[1], 1.0f};
  float Y[3];
  act_so3(qij, X, Y);
  Y[0] += d * tij[0]; Y[1] += d * tij[1]; Y[2] += d * tij[2];
  const float Z = (Y[2] < 0.1f) ? 1.0f : Y[2];
  const float iz = 1.0f / Z;
  const float x1 = Kj[0] * (Y[0] * iz) + Kj[2], y1 = Kj[1] * (Y[1] * iz) + Kj[3];
  if (coords) { coords[2 * o] = x1; coords[2 * o + 1] = y1; }
  if (valid) valid[o] = (Y[2] > 0.2f) ? 1.0f : 0.0f;
  if (motn) {
    const float tx = target[2 * o], ty = target[2 * o + 1];
    float* m = motn + size_t(e) * 4 * HW + k;
    m[0] = fminf(fmaxf(x1 - u, -64.0f), 64.0f);
    m[HW] = fminf(fmaxf(y1 - v, -64.0f), 64.0f);
    m[2 * size_t(HW)] = fminf(fmaxf(tx - x1, -64.0f), 64.0f);
    m[3 * size_t(HW)] = fminf(fmaxf(ty - y1, -64.0f), 64.0f);
  }
}

}  // namespace db200

using namespace db200;

extern "C" int db200_frame_distance(const float* poses, const float* disps, const float* intrinsics, const int64_t* ii,
                                    const int64_t* jj, int K, int n_poses, int B, int ht, int wd, float beta, float* dist,
                                    db200_stream_t stream) {
  if (!poses || !disps || !intrinsics || !dist || K < 0 || ht <= 0 || wd <= 0 || n_poses <= 0 || B <= 0) return DB200_ERR_ARG;
  if (K == 0) return DB200_OK;
  if (!ii || !jj) return DB200_ERR_ARG;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  static const bool no_batch = getenv("DB200_FD_NO_BATCH") != nullptr;
  const int npix = cdiv(ht * wd, 256);
  const int nb = cdiv(K, kFdBatch);
  if (!no_batch && npix <= 12 && K >= 4 * kFdBatch * kNumSMs) {       // enough pairs to fill the GPU with batched CTAs
    if (npix <= 10) frame_distance_batch_kernel<10><<<nb, 256, 0, st>>>(poses, disps, intrinsics, ii, jj, K, n_poses, B, ht, wd, beta, dist);
    else frame_distance_batch_kernel<12><<<nb, 256, 0, st>>>(poses, disps, intrinsics, ii, jj, K, n_poses, B, ht, wd, beta, dist);
  } else {
    frame_distance_kernel<<<K, 256, 0, st>>>(poses, disps, intrinsics, ii, jj, K, n_poses, B, ht, wd, beta, dist);
  }
  DB200_CHECK_LAUNCH();
  return DB200_OK;
}

extern "C" int db200_iproj(const float* poses, const float* disps, const float* intrinsics, int num, int ht, int wd,
                           float* points, db200_stream_t stream) {
  if (!poses || !disps || !intrinsics || !points || num < 0 || ht <= 0 || wd <= 0) return DB200_ERR_ARG;
  if (num == 0) return DB200_OK;
  dim3 grid(cdiv(ht * wd, 256), num);
  iproj_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(poses, disps, intrinsics, num, ht, wd, points);
  DB200_CHECK_LAUNCH();
  return DB200_OK;
}

extern "C" int db200_depth_filter(const float* poses, const float* disps, const float* intrinsics, const int64_t* inds,
                                  const float* thresh, int n_inds, int num, int ht, int wd, float* counter,
                                  db200_stream_t stream) {
  if (!poses || !disps || !intrinsics || !counter || n_inds < 0 || num <= 0 || ht <= 0 || wd <= 0) return DB200_ERR_ARG;
  if (n_inds == 0) return DB200_OK;
  if (!inds || !thresh) return DB200_ERR_ARG;
  dim3 grid(cdiv(ht * wd, 256), n_inds);
  depth_filter_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(poses, disps, intrinsics, inds, thresh, num, ht, wd,
                                                                          counter);
  DB200_CHECK_LAUNCH();
  return DB200_OK;
}

extern "C" int db200_projmap(const float* poses, const float* disps, const float* intrinsics, const int64_t* ii,
                             const int64_t* jj, int num, int n_poses, int B, int ht, int wd, float* coords, float* valid,
                             db200_stream_t stream) {
  if (!poses || !disps || !intrinsics || !coords || !valid || num < 0 || ht <= 0 || wd <= 0 || n_poses <= 0 || B <= 0)
    return DB200_ERR_ARG;
  if (num == 0) return DB200_OK;
  if (!ii || !jj) return DB200_ERR_ARG;
  dim3 grid(cdiv(ht * wd, 256), num);
  projmap_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(poses, disps, intrinsics, ii, jj, n_poses, B, ht, wd, coords,
                                                                      valid);
  DB200_CHECK_LAUNCH();
  return DB200_OK;
}

extern "C" int db200_reproject(const float* poses, const float* disps, const float* intrinsics, int per_frame_intrinsics,
                               const int64_t* ii, const int64_t* jj, int N, int n_poses, int B, int ht, int wd, float* coords,
                               float* valid, const float* target, float* motn, db200_stream_t stream) {
  if (!poses || !disps || !intrinsics || N < 0 || ht <= 0 || wd <= 0 || n_poses <= 0 || B <= 0 || (motn && !target) ||
      (!coords && !valid && !motn))
    return DB200_ERR_ARG;
  if (N == 0) return DB200_OK;
  if (!ii || !jj) return DB200_ERR_ARG;
  dim3 grid(cdiv(ht * wd, 256), N);
  reproject_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(poses, disps, intrinsics, per_frame_intrinsics ? 4 : 0, ii,
                                                                        jj, n_poses, B, ht, wd, coords, valid, target, motn);
  DB200_CHECK_LAUNCH();
  return DB200_OK;
}
