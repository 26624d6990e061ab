// ba_solve.cu — on-GPU damped Cholesky solve of the reduced pose system, pose retraction.
// Replaces SparseBlock::solve (host Eigen::SimplicialLLT in fp64 with D2H/H2D copies every iteration,
// /root/reference/src/droid_kernels.cu:1192-1213), pose_retr_kernel/retrSE3/expSE3 (:877-931, :110-175).
//
// Right-looking blocked Cholesky in fp64 (B200 keeps a full-rate FP64 pipe), NB = 32:
//   step k:  panel kernel  — every CTA re-factors the 32x32 diagonal tile in shared memory (one warp,
//                            ~2 us) and solves its own 128 rows of the panel against it (thread = row);
//            update kernel — 128x128 macro tiles, A_ij -= L_ik L_jk^T, 8x8 register tile per thread.
// The right-hand side rides along as one extra row of the matrix ([A b; b^T .] -> its factor row is
// y = L^-1 b), so forward substitution costs nothing; back substitution walks the tile rows upwards.
// Damping diag += ep + lm*diag (:1197) is applied while copying into the work matrix.  A pivot <= 0
// reproduces Eigen's NumericalIssue: the pose update is zero (:1207-1210).
#include "ba.cuh"

namespace db200 {

constexpr int NB = 32;

__global__ void chol_prepare_kernel(BAParams p, float lm, float ep) {
  const int n = p.n, ld = p.ld;
  const size_t total = size_t(ld + NB) * ld;
  const double* H = p.Hb;
  const double* b = p.Hb + size_t(n) * n;
  if (blockIdx.x == 0 && threadIdx.x == 0) p.hdr->chol_fail = 0;
  for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += size_t(gridDim.x) * blockDim.x) {
    const int r = int(i / ld), c = int(i - size_t(r) * ld);
    double v = 0.0;
    if (r < n) {
      if (c <= r) {
        v = H[size_t(r) * n + c];
        if (c == r) v += double(ep) + double(lm) * v;
      }
    } else if (r < ld) {
      v = (c == r) ? 1.0 : 0.0;
    } else if (r == ld) {
      v = (c < n) ? b[c] : 0.0;
    }
    p.Wk[i] = v;
  }
}

// grid.x = ceil(rows_below / 128); 128 threads.  rows_below = (ld + NB) - NB*(k+1)
__global__ void __launch_bounds__(128) chol_panel_kernel(BAParams p, int k) {
  __shared__ double T[NB][NB + 1];
  const int ld = p.ld, n = p.n;
  const int tid = threadIdx.x;
  const double* Akk = p.Wk + size_t(NB * k) * ld + NB * k;
  for (int i = tid; i < NB * NB; i += 128) T[i / NB][i % NB] = Akk[size_t(i / NB) * ld + (i % NB)];
  __syncthreads();
  if (tid < 32) {                         // one warp factors the diagonal tile; lane = row
    const int r = tid;
    for (int c = 0; c < NB; c++) {
      const double dcc = T[c][c];
      __syncwarp();
      if (dcc <= 0.0 && NB * k + c < n && r == c) atomicExch(&p.hdr->chol_fail, 1);   // NaN passes, like Eigen's `d <= 0`
      const double s = sqrt(dcc);
      if (r == c) T[c][c] = s;
      if (r > c) T[r][c] = T[r][c] / s;
      __syncwarp();
      if (r > c) {
        const double lrc = T[r][c];
        for (int m = c + 1; m <= r; m++) T[r][m] -= lrc * T[m][c];
      }
      __syncwarp();
    }
  }
  __syncthreads();
  if (blockIdx.x == 0) {
    double* Lkk = p.Lm + size_t(NB * k) * ld + NB * k;
    for (int i = tid; i < NB * NB; i += 128) {
      const int r = i / NB, c = i % NB;
      Lkk[size_t(r) * ld + c] = (c <= r) ? T[r][c] : 0.0;
    }
  }
  const int gr = NB * (k + 1) + blockIdx.x * 128 + tid;
  if (gr >= ld + NB) return;
  const double* arow = p.Wk + size_t(gr) * ld + NB * k;
  double x[NB];
#pragma unroll
  for (int c = 0; c < NB; c++) x[c] = arow[c];
#pragma unroll
  for (int c = 0; c < NB; c++) {          // x L_kk^T = a  ->  forward substitution along the row
    double s = x[c];
#pragma unroll
    for (int m = 0; m < c; m++) s -= x[m] * T[c][m];
    x[c] = s / T[c][c];
  }
  double* lrow = p.Lm + size_t(gr) * ld + NB * k;
#pragma unroll
  for (int c = 0; c < NB; c++) lrow[c] = x[c];
}

// trailing update A_ij -= L_ik L_jk^T on 128x128 macro tiles; block (16,16); dynamic smem 2*128*33 doubles
__global__ void __launch_bounds__(256) chol_update_kernel(BAParams p, int k) {
  extern __shared__ __align__(16) double sm_d[];
  double (*Pr)[NB + 1] = reinterpret_cast<double (*)[NB + 1]>(sm_d);
  double (*Pc)[NB + 1] = reinterpret_cast<double (*)[NB + 1]>(sm_d + 128 * (NB + 1));
  const int ld = p.ld;
  const int base = NB * (k + 1);
  const int row0 = base + blockIdx.y * 128, col0 = base + blockIdx.x * 128;
  if (col0 > row0 + 127) return;          // entirely above the diagonal
  const int nrows = ld + NB, ncols = ld;
  const int tid = threadIdx.y * 16 + threadIdx.x;
  for (int i = tid; i < 128 * NB; i += 256) {
    const int r = i / NB, c = i % NB;
    Pr[r][c] = (row0 + r < nrows) ? p.Lm[size_t(row0 + r) * ld + NB * k + c] : 0.0;
    Pc[r][c] = (col0 + r < ncols) ? p.Lm[size_t(col0 + r) * ld + NB * k + c] : 0.0;
  }
  __syncthreads();
  double acc[8][8];
#pragma unroll
  for (int i = 0; i < 8; i++)
#pragma unroll
    for (int j = 0; j < 8; j++) acc[i][j] = 0.0;
#pragma unroll 4
  for (int kk = 0; kk < NB; kk++) {
    double a[8], b[8];
#pragma unroll
    for (int i = 0; i < 8; i++) { a[i] = Pr[threadIdx.y + 16 * i][kk]; b[i] = Pc[threadIdx.x + 16 * i][kk]; }
#pragma unroll
    for (int i = 0; i < 8; i++)
#pragma unroll
      for (int j = 0; j < 8; j++) acc[i][j] = fma(a[i], b[j], acc[i][j]);
  }
#pragma unroll
  for (int i = 0; i < 8; i++) {
    const int r = row0 + threadIdx.y + 16 * i;
    if (r >= nrows) continue;
#pragma unroll
    for (int j = 0; j < 8; j++) {
      const int c = col0 + threadIdx.x + 16 * j;
      if (c < ncols && c <= r) p.Wk[size_t(r) * ld + c] -= acc[i][j];
    }
  }
}

__global__ void chol_backsub_init_kernel(BAParams p) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < p.ld) p.xsol[i] = p.Lm[size_t(p.ld) * p.ld + i];   // y = factor row of the appended rhs
}

// step i (descending): grid = i+1 CTAs of one warp.  x_i = L_ii^-T y_i (every CTA, redundantly);
// CTA kk < i: y_kk -= L_{i,kk}^T x_i;  CTA i stores x_i.
__global__ void __launch_bounds__(32) chol_backsub_step_kernel(BAParams p, int i) {
  __shared__ double T[NB][NB + 1];
  const int ld = p.ld, lane = threadIdx.x, kk = blockIdx.x;
  double* y = p.xsol;
  double* x = p.xsol + (ld + NB);
  const double* Lii = p.Lm + size_t(NB * i) * ld + NB * i;
  for (int r = 0; r < NB; r++) T[r][lane] = Lii[size_t(r) * ld + lane];
  __syncwarp();
  double xv = y[NB * i + lane];           // lane m ends up holding x_i[m]
  for (int c = NB - 1; c >= 0; c--) {
    double s = (lane > c) ? T[lane][c] * xv : 0.0;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (lane == c) xv = (xv - s) / T[c][c];
  }
  if (kk == i) { x[NB * i + lane] = xv; return; }
  const double* Lik = p.Lm + size_t(NB * i) * ld + NB * kk;
  double s = 0.0;
  for (int m = 0; m < NB; m++) s += Lik[size_t(m) * ld + lane] * __shfl_sync(0xffffffffu, xv, m);
  y[NB * kk + lane] -= s;
}

// dx (fp32) = failed ? 0 : x ;  also the user-visible dx_out
__global__ void post_dx_kernel(BAParams p, float* __restrict__ dx_out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (p.hdr->status < 0) return;
  const bool fail = p.hdr->chol_fail != 0;
  if (i == 0 && fail) { p.hdr->any_fail += 1; if (p.hdr->status == 0) p.hdr->status = 1; }
  if (i < p.n) {
    const float v = fail ? 0.0f : float(p.xsol[p.ld + NB + i]);
    p.dx[i] = v;
    if (dx_out) dx_out[i] = v;
  }
}

// y_e = [0 < j_e-t0 < P] dx_j - [0 < i-t0 < P] A_e^T dx_i   (EvT6x1_kernel's `idx <= 0` quirk, :1105)
__global__ void post_y_kernel(BAParams p) {
  const int q = blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= p.N || p.hdr->status < 0) return;
  const EdgeRec r = p.erec[q];
  float y[6] = {0, 0, 0, 0, 0, 0};
  if (r.slot >= 0) {
    const int ip = int(p.ii[r.eid]) - p.t0, jpz = r.jpose;
    if (jpz > 0 && jpz < p.P)
      for (int c = 0; c < 6; c++) y[c] = p.dx[6 * jpz + c];
    if (ip > 0 && ip < p.P) {
      const float* v = p.dx + 6 * ip;
      const float tx[9] = {0.0f, -r.t[2], r.t[1], r.t[2], 0.0f, -r.t[0], -r.t[1], r.t[0], 0.0f};
      // A^T v = [R v_tau + [t]x R v_phi ; R v_phi]   (A = [[R^T,0],[-R^T [t]x, R^T]])
      float Rt[3], Rp[3];
      for (int a = 0; a < 3; a++) {
        Rt[a] = r.R[a * 3] * v[0] + r.R[a * 3 + 1] * v[1] + r.R[a * 3 + 2] * v[2];
        Rp[a] = r.R[a * 3] * v[3] + r.R[a * 3 + 1] * v[4] + r.R[a * 3 + 2] * v[5];
      }
      for (int a = 0; a < 3; a++) {
        y[a] -= Rt[a] + (tx[a * 3] * Rp[0] + tx[a * 3 + 1] * Rp[1] + tx[a * 3 + 2] * Rp[2]);
        y[3 + a] -= Rp[a];
      }
    }
  }
  for (int c = 0; c < 6; c++) p.yvec[size_t(q) * 6 + c] = y[c];
}

// poses[k] <- Exp(dx[k-t0]) * poses[k]   (retrSE3/expSE3/expSO3, :110-175, :877-895)
__global__ void pose_retr_kernel(BAParams p) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= p.P || p.hdr->status < 0) return;
  float* pose = p.poses + 7 * size_t(p.t0 + i);
  const float* xi = p.dx + 6 * i;
  const float phi[3] = {xi[3], xi[4], xi[5]};
  float tau[3] = {xi[0], xi[1], xi[2]};
  const float theta_sq = phi[0] * phi[0] + phi[1] * phi[1] + phi[2] * phi[2];
  const float theta_p4 = theta_sq * theta_sq;
  const float theta = sqrtf(theta_sq);
  float imag, real;
  if (double(theta_sq) < 1e-8) {
    imag = float(0.5 - (1.0 / 48.0) * double(theta_sq) + (1.0 / 3840.0) * double(theta_p4));
    real = float(1.0 - (1.0 / 8.0) * double(theta_sq) + (1.0 / 384.0) * double(theta_p4));
  } else {
    imag = sinf(0.5f * theta) / theta;
    real = cosf(0.5f * theta);
  }
  const float dq[4] = {imag * phi[0], imag * phi[1], imag * phi[2], real};
  float dt[3] = {tau[0], tau[1], tau[2]};
  if (double(theta) > 1e-4) {
    const float a = (1.0f - cosf(theta)) / theta_sq;
    float c1[3] = {phi[1] * tau[2] - phi[2] * tau[1], phi[2] * tau[0] - phi[0] * tau[2], phi[0] * tau[1] - phi[1] * tau[0]};
    dt[0] += a * c1[0]; dt[1] += a * c1[1]; dt[2] += a * c1[2];
    const float b = (theta - sinf(theta)) / (theta * theta_sq);
    float c2[3] = {phi[1] * c1[2] - phi[2] * c1[1], phi[2] * c1[0] - phi[0] * c1[2], phi[0] * c1[1] - phi[1] * c1[0]};
    dt[0] += b * c2[0]; dt[1] += b * c2[1]; dt[2] += b * c2[2];
  }
  const float t[3] = {pose[0], pose[1], pose[2]};
  const float q[4] = {pose[3], pose[4], pose[5], pose[6]};
  float q1[4], t1[3];
  q1[0] = dq[3] * q[0] + dq[0] * q[3] + dq[1] * q[2] - dq[2] * q[1];
  q1[1] = dq[3] * q[1] + dq[1] * q[3] + dq[2] * q[0] - dq[0] * q[2];
  q1[2] = dq[3] * q[2] + dq[2] * q[3] + dq[0] * q[1] - dq[1] * q[0];
  q1[3] = dq[3] * q[3] - dq[0] * q[0] - dq[1] * q[1] - dq[2] * q[2];
  act_so3(dq, t, t1);
  pose[0] = t1[0] + dt[0]; pose[1] = t1[1] + dt[1]; pose[2] = t1[2] + dt[2];
  pose[3] = q1[0]; pose[4] = q1[1]; pose[5] = q1[2]; pose[6] = q1[3];
}

int ba_solve(const BAParams& p, float lm, float ep, cudaStream_t st) {
  const int nt = p.nt, ld = p.ld;
  chol_prepare_kernel<<<4 * kNumSMs, 256, 0, st>>>(p, lm, ep);
  const size_t upd_smem = sizeof(double) * 2 * 128 * (NB + 1);
  static bool attr_set = false;
  if (!attr_set) {
    DB200_CUDA(cudaFuncSetAttribute(chol_update_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, int(upd_smem)));
    attr_set = true;
  }
  for (int k = 0; k < nt; k++) {
    const int rows_below = (ld + NB) - NB * (k + 1);
    chol_panel_kernel<<<cdiv(rows_below, 128), 128, 0, st>>>(p, k);
    const int cols_below = ld - NB * (k + 1);
    if (cols_below > 0) {
      dim3 grid(cdiv(cols_below, 128), cdiv(rows_below, 128));
      chol_update_kernel<<<grid, dim3(16, 16), upd_smem, st>>>(p, k);
    }
  }
  chol_backsub_init_kernel<<<cdiv(ld, 256), 256, 0, st>>>(p);
  for (int i = nt - 1; i >= 0; i--) chol_backsub_step_kernel<<<i + 1, 32, 0, st>>>(p, i);
  DB200_COUNT(1 + nt + (nt - 1) + nt);   // prepare, panels, updates, (init counted below), back-substitution steps
  DB200_CHECK_LAUNCH();
  return DB200_OK;
}

int ba_post_solve(const BAParams& p, float* dx_out, cudaStream_t st) {
  post_dx_kernel<<<cdiv(p.n, 256), 256, 0, st>>>(p, dx_out);
  if (p.N > 0 && !p.motion_only) post_y_kernel<<<cdiv(p.N, 256), 256, 0, st>>>(p);
  pose_retr_kernel<<<cdiv(p.P, 128), 128, 0, st>>>(p);
  DB200_COUNT((p.N > 0 && !p.motion_only) ? 2 : 1);
  DB200_CHECK_LAUNCH();
  return DB200_OK;
}

}  // namespace db200
