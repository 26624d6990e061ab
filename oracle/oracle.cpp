// =====================================================================================
// oracle.cpp — CPU restatement of the reference's CUDA hot path.  TEST INFRASTRUCTURE ONLY.
//
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / `--impl reference` legs
// may load this library.  The product path (droid-slam_reserch_b200/) never links or calls it.
//
// What it restates (all citations relative to /root/reference):
//   * src/droid_kernels.cu       :58-175   SE3 helpers (actSO3/actSE3/adjSE3/relSE3/expSO3/expSE3)
//                                 :176-424  projective_transform_kernel (per-edge J, H, E, C, b)
//                                 :518-657  frame_distance_kernel
//                                 :854-998  accum_kernel / accum_cuda (segmented row sums)
//                                 :1001-1115 EEt6x6 / Ev6x1 / EvT6x1 kernels
//                                 :1117-1219 SparseBlock (fp64 assembly, damping, LLT, zero on failure)
//                                 :1222-1311 schur_block
//                                 :1314-1434 ba_cuda driver loop
//                                 :661-850  depth_filter_kernel / iproj_kernel
//   * src/correlation_kernels.cu :20-124   corr_index forward / backward
//   * src/altcorr_kernel.cu      :27-286   altcorr forward / backward
//   * droid_slam/modules/corr.py :23-38,63-71  all-pairs correlation + avg-pool pyramid
//
// Arithmetic mirrors the CUDA source: fp32 everywhere on the "device" side, the 256-thread
// strided accumulation followed by the 256->128->64->warp tree of blockReduce (:36-55), double
// constants where the source writes them (`.001 * w`, `1.0/48.0`, `1e-8`), and an fp64 solve.
// The one deliberate substitution: Eigen::SimplicialLLT (sparse, AMD-ordered) is replaced by a
// dense fp64 Cholesky of the same damped matrix; oracle/ref_eigen_solve.cpp (built into
// oracle/_ref/ from the vendored Eigen) checks that substitution.
//
// Parity pinning: the reference ships no golden vectors for this path.  This oracle
// is pinned against (a) the reference's own PyTorch BA (droid_slam/geom/ba.py) imported unmodified
// under shims — tests/golden/make_pyref_golden.py — and (b) outputs of the reference CUDA extension
// itself run on a B200 — tests/golden/make_ref_cuda_golden.py.  See DESIGN.md §Oracle.
// =====================================================================================
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <vector>
#include <algorithm>
#include <numeric>
#ifdef _OPENMP
#include <omp.h>
#endif

#define ORACLE_API extern "C" __attribute__((visibility("default")))

namespace {

constexpr int THREADS = 256;  // droid_kernels.cu:28

// ---- fp32 multiply-add exactly as nvcc emits it for `acc += a * b` (FFMA contraction) ----
template <typename R> inline R mad(R a, R b, R c);
template <> inline float mad<float>(float a, float b, float c) { return fmaf(a, b, c); }
template <> inline double mad<double>(double a, double b, double c) { return fma(a, b, c); }

// blockReduce (droid_kernels.cu:36-55): tree over 256 per-thread partials.
template <typename R>
R block_reduce(R* s) {
  for (int t = 0; t < 128; t++) s[t] += s[t + 128];
  for (int t = 0; t < 64; t++) s[t] += s[t + 64];
  for (int st = 32; st >= 1; st >>= 1)
    for (int t = 0; t < st; t++) s[t] += s[t + st];
  return s[0];
}

// ---------------- SE3 helpers, droid_kernels.cu:58-175 ----------------
template <typename R>
void actSO3(const R* q, const R* X, R* Y) {
  R uv[3];
  uv[0] = R(2.0 * double(q[1] * X[2] - q[2] * X[1]));
  uv[1] = R(2.0 * double(q[2] * X[0] - q[0] * X[2]));
  uv[2] = R(2.0 * double(q[0] * X[1] - q[1] * X[0]));
  Y[0] = X[0] + q[3] * uv[0] + (q[1] * uv[2] - q[2] * uv[1]);
  Y[1] = X[1] + q[3] * uv[1] + (q[2] * uv[0] - q[0] * uv[2]);
  Y[2] = X[2] + q[3] * uv[2] + (q[0] * uv[1] - q[1] * uv[0]);
}

template <typename R>
void actSE3(const R* t, const R* q, const R* X, R* Y) {
  actSO3(q, X, Y);
  Y[3] = X[3];
  Y[0] += X[3] * t[0];
  Y[1] += X[3] * t[1];
  Y[2] += X[3] * t[2];
}

template <typename R>
void adjSE3(const R* t, const R* q, const R* X, R* Y) {
  R qinv[4] = {-q[0], -q[1], -q[2], q[3]};
  actSO3(qinv, &X[0], &Y[0]);
  actSO3(qinv, &X[3], &Y[3]);
  R u[3], v[3];
  u[0] = t[2] * X[1] - t[1] * X[2];
  u[1] = t[0] * X[2] - t[2] * X[0];
  u[2] = t[1] * X[0] - t[0] * X[1];
  actSO3(qinv, u, v);
  Y[3] += v[0];
  Y[4] += v[1];
  Y[5] += v[2];
}

template <typename R>
void relSE3(const R* ti, const R* qi, const R* tj, const R* qj, R* tij, R* qij) {
  qij[0] = -qj[3] * qi[0] + qj[0] * qi[3] - qj[1] * qi[2] + qj[2] * qi[1];
  qij[1] = -qj[3] * qi[1] + qj[1] * qi[3] - qj[2] * qi[0] + qj[0] * qi[2];
  qij[2] = -qj[3] * qi[2] + qj[2] * qi[3] - qj[0] * qi[1] + qj[1] * qi[0];
  qij[3] = qj[3] * qi[3] + qj[0] * qi[0] + qj[1] * qi[1] + qj[2] * qi[2];
  actSO3(qij, ti, tij);
  tij[0] = tj[0] - tij[0];
  tij[1] = tj[1] - tij[1];
  tij[2] = tj[2] - tij[2];
}

template <typename R> inline R r_sqrt(R x);
template <> inline float r_sqrt<float>(float x) { return sqrtf(x); }
template <> inline double r_sqrt<double>(double x) { return sqrt(x); }
template <typename R> inline R r_sin(R x);
template <> inline float r_sin<float>(float x) { return sinf(x); }
template <> inline double r_sin<double>(double x) { return sin(x); }
template <typename R> inline R r_cos(R x);
template <> inline float r_cos<float>(float x) { return cosf(x); }
template <> inline double r_cos<double>(double x) { return cos(x); }

template <typename R>
void expSO3(const R* phi, R* q) {
  R theta_sq = phi[0] * phi[0] + phi[1] * phi[1] + phi[2] * phi[2];
  R theta_p4 = theta_sq * theta_sq;
  R theta = r_sqrt(theta_sq);
  R imag, real;
  if (double(theta_sq) < 1e-8) {
    imag = R(0.5 - (1.0 / 48.0) * double(theta_sq) + (1.0 / 3840.0) * double(theta_p4));
    real = R(1.0 - (1.0 / 8.0) * double(theta_sq) + (1.0 / 384.0) * double(theta_p4));
  } else {
    imag = r_sin(R(0.5 * double(theta))) / theta;
    real = r_cos(R(0.5 * double(theta)));
  }
  q[0] = imag * phi[0];
  q[1] = imag * phi[1];
  q[2] = imag * phi[2];
  q[3] = real;
}

template <typename R>
void crossInplace(const R* a, R* b) {
  R x[3] = {a[1] * b[2] - a[2] * b[1], a[2] * b[0] - a[0] * b[2], a[0] * b[1] - a[1] * b[0]};
  b[0] = x[0]; b[1] = x[1]; b[2] = x[2];
}

template <typename R>
void expSE3(const R* xi, R* t, R* q) {
  expSO3(xi + 3, q);
  R tau[3] = {xi[0], xi[1], xi[2]};
  R phi[3] = {xi[3], xi[4], xi[5]};
  R theta_sq = phi[0] * phi[0] + phi[1] * phi[1] + phi[2] * phi[2];
  R theta = r_sqrt(theta_sq);
  t[0] = tau[0]; t[1] = tau[1]; t[2] = tau[2];
  if (double(theta) > 1e-4) {
    R a = (1 - r_cos(theta)) / theta_sq;
    crossInplace(phi, tau);
    t[0] += a * tau[0]; t[1] += a * tau[1]; t[2] += a * tau[2];
    R b = (theta - r_sin(theta)) / (theta * theta_sq);
    crossInplace(phi, tau);
    t[0] += b * tau[0]; t[1] += b * tau[1]; t[2] += b * tau[2];
  }
}

// retrSE3, droid_kernels.cu:877-895
template <typename R>
void retrSE3(const R* xi, const R* t, const R* q, R* t1, R* q1) {
  R dt[3] = {0, 0, 0};
  R dq[4] = {0, 0, 0, 1};
  expSE3(xi, dt, dq);
  q1[0] = dq[3] * q[0] + dq[0] * q[3] + dq[1] * q[2] - dq[2] * q[1];
  q1[1] = dq[3] * q[1] + dq[1] * q[3] + dq[2] * q[0] - dq[0] * q[2];
  q1[2] = dq[3] * q[2] + dq[2] * q[3] + dq[0] * q[1] - dq[1] * q[0];
  q1[3] = dq[3] * q[3] - dq[0] * q[0] - dq[1] * q[1] - dq[2] * q[2];
  actSO3(dq, t, t1);
  t1[0] += dt[0]; t1[1] += dt[1]; t1[2] += dt[2];
}

// ---------------- projective_transform_kernel, droid_kernels.cu:176-424 ----------------
// One call = one CUDA block (one edge).  Outputs use the reference layouts:
//   Hs[4][N][6][6], vs[2][N][6], Eii/Eij[N][6][HW], Cii/bz[N][HW].
template <typename R>
void edge_kernel(int e, int N, int ht, int wd, const R* poses, const R* disps, const R* intr,
                 const int64_t* ii, const int64_t* jj, const R* target, const R* weight,
                 R* Hs, R* vs, R* Eii, R* Eij, R* Cii, R* bz) {
  const int HW = ht * wd;
  const int ix = int(ii[e]), jx = int(jj[e]);
  const R fx = intr[0], fy = intr[1], cx = intr[2], cy = intr[3];
  R tij[3], qij[4];
  if (ix == jx) {
    tij[0] = R(-0.1); tij[1] = 0; tij[2] = 0;
    qij[0] = 0; qij[1] = 0; qij[2] = 0; qij[3] = 1;
  } else {
    relSE3(poses + 7 * ix, poses + 7 * ix + 3, poses + 7 * jx, poses + 7 * jx + 3, tij, qij);
  }
  std::vector<R> th(size_t(THREADS) * 90, R(0));  // per-thread hij[78], vi[6], vj[6]
  const R* tgt = target + size_t(e) * 2 * HW;
  const R* wgt = weight + size_t(e) * 2 * HW;
  const R* dsp = disps + size_t(ix) * HW;
  for (int tid = 0; tid < THREADS; tid++) {
    R* hij = &th[size_t(tid) * 90];
    R* vi = hij + 78;
    R* vj = hij + 84;
    for (int k = tid; k < HW; k += THREADS) {
      const int i = k / wd, j = k % wd;
      const R u = R(j), v = R(i);
      R Xi[4], Xj[4], Jx[12], Jz;
      R* Ji = &Jx[0];
      R* Jj = &Jx[6];
      Xi[0] = (u - cx) / fx;
      Xi[1] = (v - cy) / fy;
      Xi[2] = 1;
      Xi[3] = dsp[k];
      actSE3(tij, qij, Xi, Xj);
      const R x = Xj[0], y = Xj[1], h = Xj[3];
      const bool bad = double(Xj[2]) < 0.25;  // MIN_DEPTH
      const R d = bad ? R(0.0) : R(1.0) / Xj[2];
      const R d2 = d * d;
      R wu = bad ? R(0.0) : R(.001 * double(wgt[k]));
      R wv = bad ? R(0.0) : R(.001 * double(wgt[HW + k]));
      const R ru = tgt[k] - (fx * d * x + cx);
      const R rv = tgt[HW + k] - (fy * d * y + cy);

      Jj[0] = fx * (h * d);
      Jj[1] = fx * 0;
      Jj[2] = fx * (-x * h * d2);
      Jj[3] = fx * (-x * y * d2);
      Jj[4] = fx * (1 + x * x * d2);
      Jj[5] = fx * (-y * d);
      Jz = fx * (tij[0] * d - tij[2] * (x * d2));
      Cii[size_t(e) * HW + k] = wu * Jz * Jz;
      bz[size_t(e) * HW + k] = wu * ru * Jz;
      if (ix == jx) wu = 0;
      adjSE3(tij, qij, Jj, Ji);
      for (int n = 0; n < 6; n++) Ji[n] *= -1;
      int l = 0;
      for (int n = 0; n < 12; n++)
        for (int m = 0; m <= n; m++) { hij[l] += wu * Jx[n] * Jx[m]; l++; }
      for (int n = 0; n < 6; n++) {
        vi[n] += wu * ru * Ji[n];
        vj[n] += wu * ru * Jj[n];
        Eii[(size_t(e) * 6 + n) * HW + k] = wu * Jz * Ji[n];
        Eij[(size_t(e) * 6 + n) * HW + k] = wu * Jz * Jj[n];
      }

      Jj[0] = fy * 0;
      Jj[1] = fy * (h * d);
      Jj[2] = fy * (-y * h * d2);
      Jj[3] = fy * (-1 - y * y * d2);
      Jj[4] = fy * (x * y * d2);
      Jj[5] = fy * (x * d);
      Jz = fy * (tij[1] * d - tij[2] * (y * d2));
      Cii[size_t(e) * HW + k] += wv * Jz * Jz;
      bz[size_t(e) * HW + k] += wv * rv * Jz;
      if (ix == jx) wv = 0;
      adjSE3(tij, qij, Jj, Ji);
      for (int n = 0; n < 6; n++) Ji[n] *= -1;
      l = 0;
      for (int n = 0; n < 12; n++)
        for (int m = 0; m <= n; m++) { hij[l] += wv * Jx[n] * Jx[m]; l++; }
      for (int n = 0; n < 6; n++) {
        vi[n] += wv * rv * Ji[n];
        vj[n] += wv * rv * Jj[n];
        Eii[(size_t(e) * 6 + n) * HW + k] += wv * Jz * Ji[n];
        Eij[(size_t(e) * 6 + n) * HW + k] += wv * Jz * Jj[n];
      }
    }
  }
  R s[THREADS];
  for (int n = 0; n < 6; n++) {
    for (int t = 0; t < THREADS; t++) s[t] = th[size_t(t) * 90 + 78 + n];
    vs[(size_t(0) * N + e) * 6 + n] = block_reduce(s);
    for (int t = 0; t < THREADS; t++) s[t] = th[size_t(t) * 90 + 84 + n];
    vs[(size_t(1) * N + e) * 6 + n] = block_reduce(s);
  }
  int l = 0;
  auto H = [&](int blk, int r, int c) -> R& { return Hs[((size_t(blk) * N + e) * 6 + r) * 6 + c]; };
  for (int n = 0; n < 12; n++) {
    for (int m = 0; m <= n; m++) {
      for (int t = 0; t < THREADS; t++) s[t] = th[size_t(t) * 90 + l];
      const R val = block_reduce(s);
      if (n < 6 && m < 6) { H(0, n, m) = val; H(0, m, n) = val; }
      else if (n >= 6 && m < 6) { H(1, m, n - 6) = val; H(2, n - 6, m) = val; }
      else { H(3, n - 6, m - 6) = val; H(3, m - 6, n - 6) = val; }
      l++;
    }
  }
}

// accum_cuda + accum_kernel, droid_kernels.cu:854-874, 948-998: out[j] = sum_{e: ix[e]==jx[j]} data[e]
// (edges visited in ascending-key order; ties in original order = stable argsort).
template <typename R>
void accum(const R* data, int nrows, int D, const int64_t* ix, const int64_t* jx, int nout, R* out) {
  std::vector<int> order(nrows);
  std::iota(order.begin(), order.end(), 0);
  std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return ix[a] < ix[b]; });
  std::vector<std::vector<int>> cols(nout);
  int i = 0;
  for (int j = 0; j < nout; j++) {
    while (i < nrows && ix[order[i]] <= jx[j]) {
      if (ix[order[i]] == jx[j]) cols[j].push_back(order[i]);
      i++;
    }
  }
#pragma omp parallel for schedule(dynamic)
  for (int j = 0; j < nout; j++) {
    for (int k = 0; k < D; k++) {
      R x = 0;
      for (int r : cols[j]) x += data[size_t(r) * D + k];
      out[size_t(j) * D + k] = x;
    }
  }
}

// dense fp64 Cholesky solve of (A + damping) x = b.  Stands in for SparseBlock::solve
// (droid_kernels.cu:1192-1213).  Returns false (x := 0) when a pivot is <= 0, like
// Eigen::SimplicialLLT's NumericalIssue.
bool chol_solve(std::vector<double>& A, std::vector<double>& b, int n, double lm, double ep,
                std::vector<double>& x) {
  for (int i = 0; i < n; i++) A[size_t(i) * n + i] += ep + lm * A[size_t(i) * n + i];
  // in-place lower Cholesky (row-major, lower triangle used)
  for (int j = 0; j < n; j++) {
    double d = A[size_t(j) * n + j];
    for (int k = 0; k < j; k++) d -= A[size_t(j) * n + k] * A[size_t(j) * n + k];
    if (d <= 0.0) { x.assign(n, 0.0); return false; }
    d = sqrt(d);
    A[size_t(j) * n + j] = d;
#pragma omp parallel for schedule(static) if (n - j > 256)
    for (int i = j + 1; i < n; i++) {
      double s = A[size_t(i) * n + j];
      const double* ai = &A[size_t(i) * n];
      const double* aj = &A[size_t(j) * n];
      for (int k = 0; k < j; k++) s -= ai[k] * aj[k];
      A[size_t(i) * n + j] = s / d;
    }
  }
  x = b;
  for (int i = 0; i < n; i++) {
    double s = x[i];
    for (int k = 0; k < i; k++) s -= A[size_t(i) * n + k] * x[k];
    x[i] = s / A[size_t(i) * n + i];
  }
  for (int i = n - 1; i >= 0; i--) {
    double s = x[i];
    for (int k = i + 1; k < n; k++) s -= A[size_t(k) * n + i] * x[k];
    x[i] = s / A[size_t(i) * n + i];
  }
  return true;
}

struct BAOptions {
  int skip_pose0_backsub;  // 1 = reference CUDA behaviour (EvT6x1_kernel:1105 skips pose index <= 0)
  int dump_system;         // 1 = write the undamped reduced system of the LAST iteration
  int damp_before_schur;   // 1 = PyTorch-path damping: H + (ep + lm*H) I, THEN minus S (geom/chol.py:55-63)
  int moba_block_damping;  // 1 = PyTorch MoBA quirk: eye(6) broadcast over EVERY 6x6 block (geom/chol.py:35-37)
};

// ba_cuda, droid_kernels.cu:1314-1434
template <typename R>
int ba_impl(R* poses, R* disps, const R* intr, const R* disps_sens, const R* target,
            const R* weight, const R* eta, const int64_t* ii, const int64_t* jj, int N, int B,
            int ht, int wd, int t0, int t1, int iterations, float lm, float ep, int motion_only,
            int eta_rows, R* dx_out, R* dz_out, const BAOptions& opt, double* Hdump, double* bdump) {
  const int HW = ht * wd;
  const int P = t1 - t0;
  if (P <= 0) return -1;
  for (int e = 0; e < N; e++)
    if (ii[e] < 0 || ii[e] >= B || jj[e] < 0 || jj[e] >= t1 || ii[e] >= t1) return -2;

  // ts, ii_exp, jj_exp, kx, kk_exp  (:1335-1344)
  std::vector<int64_t> ii_exp(P + N), jj_exp(P + N);
  for (int n = 0; n < P; n++) { ii_exp[n] = t0 + n; jj_exp[n] = t0 + n; }
  for (int e = 0; e < N; e++) { ii_exp[P + e] = ii[e]; jj_exp[P + e] = jj[e]; }
  std::vector<int64_t> kx(ii_exp);
  std::sort(kx.begin(), kx.end());
  kx.erase(std::unique(kx.begin(), kx.end()), kx.end());
  const int M = int(kx.size());
  std::vector<int64_t> kk_exp(P + N);
  for (int n = 0; n < P + N; n++)
    kk_exp[n] = std::lower_bound(kx.begin(), kx.end(), ii_exp[n]) - kx.begin();
  if (!motion_only && eta_rows != M) return -3;
  std::vector<int64_t> ts(P);
  for (int n = 0; n < P; n++) ts[n] = t0 + n;

  std::vector<R> Hs(size_t(4) * N * 36), vs(size_t(2) * N * 6);
  std::vector<R> Eii(size_t(N) * 6 * HW), Eij(size_t(N) * 6 * HW), Cii(size_t(N) * HW), wi(size_t(N) * HW);
  std::vector<double> dx(size_t(P) * 6, 0.0);
  std::vector<R> dz;
  int status = 0;
  const int n6 = 6 * P;

  for (int itr = 0; itr < iterations; itr++) {
#pragma omp parallel for schedule(dynamic)
    for (int e = 0; e < N; e++)
      edge_kernel<R>(e, N, ht, wd, poses, disps, intr, ii, jj, target, weight, Hs.data(), vs.data(),
                     Eii.data(), Eij.data(), Cii.data(), wi.data());

    // SparseBlock A: update_lhs / update_rhs (:1376-1383), fp64 accumulation
    std::vector<double> A(size_t(n6) * n6, 0.0), b(n6, 0.0);
    for (int blk = 0; blk < 4; blk++) {
      for (int e = 0; e < N; e++) {
        const int i = int(((blk < 2) ? ii[e] : jj[e]) - t0);
        const int j = int(((blk & 1) ? jj[e] : ii[e]) - t0);
        if (i >= 0 && j >= 0)
          for (int k = 0; k < 6; k++)
            for (int l = 0; l < 6; l++)
              A[size_t(6 * i + k) * n6 + 6 * j + l] += double(Hs[((size_t(blk) * N + e) * 6 + k) * 6 + l]);
      }
    }
    for (int blk = 0; blk < 2; blk++)
      for (int e = 0; e < N; e++) {
        const int i = int((blk == 0 ? ii[e] : jj[e]) - t0);
        if (i >= 0)
          for (int k = 0; k < 6; k++) b[6 * i + k] += double(vs[(size_t(blk) * N + e) * 6 + k]);
      }

    std::vector<double> x;
    if (motion_only) {
      if (opt.dump_system && Hdump) { memcpy(Hdump, A.data(), A.size() * 8); memcpy(bdump, b.data(), b.size() * 8); }
      if (opt.moba_block_damping) {
        for (int i = 0; i < P; i++)
          for (int j = 0; j < P; j++)
            for (int k = 0; k < 6; k++) {
              double& a = A[size_t(6 * i + k) * n6 + 6 * j + k];
              a += double(ep) + double(lm) * a;
            }
      }
      if (!chol_solve(A, b, n6, opt.moba_block_damping ? 0.0 : double(lm), opt.moba_block_damping ? 0.0 : double(ep), x)) status = 1;
      for (int n = 0; n < n6; n++) dx[n] = double(R(x[n]));
    } else {
      const R alpha = R(0.05);
      std::vector<R> m(size_t(M) * HW), C(size_t(M) * HW), w(size_t(M) * HW), Q(size_t(M) * HW);
      accum<R>(Cii.data(), N, HW, ii, kx.data(), M, C.data());
      accum<R>(wi.data(), N, HW, ii, kx.data(), M, w.data());
      for (int k = 0; k < M; k++)
        for (int p = 0; p < HW; p++) {
          const size_t o = size_t(k) * HW + p;
          const R ds = disps_sens[size_t(kx[k]) * HW + p];
          const R mm = ds > 0 ? R(1) : R(0);
          C[o] = C[o] + mm * alpha + (1 - mm) * eta[o];
          w[o] = w[o] - mm * alpha * (disps[size_t(kx[k]) * HW + p] - ds);
          Q[o] = R(1.0) / C[o];
        }
      // E = cat(Ei, Eij)  (:1402-1403)
      std::vector<R> E(size_t(P + N) * 6 * HW);
      accum<R>(Eii.data(), N, 6 * HW, ii, ts.data(), P, E.data());
      memcpy(E.data() + size_t(P) * 6 * HW, Eij.data(), sizeof(R) * size_t(N) * 6 * HW);

      // schur_block (:1222-1311)
      std::vector<std::vector<int>> graph(P), index(P);
      for (int n = 0; n < P + N; n++) {
        const int j = int(jj_exp[n]), k = int(kk_exp[n]);
        if (j >= t0 && j < t1) { graph[j - t0].push_back(k); index[j - t0].push_back(n); }
      }
      struct Pair { int pi, pj, ra, rb, k; };
      std::vector<Pair> pairs;
      for (int i = 0; i < P; i++)
        for (int j = 0; j < P; j++)
          for (size_t k = 0; k < graph[i].size(); k++)
            for (size_t l = 0; l < graph[j].size(); l++)
              if (graph[i][k] == graph[j][l]) pairs.push_back({i, j, index[i][k], index[j][l], graph[i][k]});
      std::vector<R> Sblk(pairs.size() * 36);
#pragma omp parallel for schedule(dynamic, 16)
      for (size_t pidx = 0; pidx < pairs.size(); pidx++) {  // EEt6x6_kernel (:1001-1056)
        const Pair& pr = pairs[pidx];
        std::vector<R> th(size_t(THREADS) * 36, R(0));
        const R* Ea = &E[size_t(pr.ra) * 6 * HW];
        const R* Eb = &E[size_t(pr.rb) * 6 * HW];
        const R* Qk = &Q[size_t(pr.k) * HW];
        for (int tid = 0; tid < THREADS; tid++) {
          R* dS = &th[size_t(tid) * 36];
          for (int k = tid; k < HW; k += THREADS) {
            R ei[6], ej[6];
            for (int n = 0; n < 6; n++) { ei[n] = Ea[size_t(n) * HW + k] * Qk[k]; ej[n] = Eb[size_t(n) * HW + k]; }
            for (int n = 0; n < 6; n++)
              for (int mm = 0; mm < 6; mm++) dS[n * 6 + mm] = mad(ei[n], ej[mm], dS[n * 6 + mm]);
          }
        }
        R s[THREADS];
        for (int v = 0; v < 36; v++) {
          for (int t = 0; t < THREADS; t++) s[t] = th[size_t(t) * 36 + v];
          Sblk[pidx * 36 + v] = block_reduce(s);
        }
      }
      std::vector<R> sv(size_t(P + N) * 6);
#pragma omp parallel for schedule(dynamic, 16)
      for (int row = 0; row < P + N; row++) {  // Ev6x1_kernel (:1059-1093)
        std::vector<R> th(size_t(THREADS) * 6, R(0));
        const int k0 = int(kk_exp[row]);
        for (int tid = 0; tid < THREADS; tid++)
          for (int k = tid; k < HW; k += THREADS) {
            const R q_w = Q[size_t(k0) * HW + k] * w[size_t(k0) * HW + k];
            for (int n = 0; n < 6; n++)
              th[size_t(tid) * 6 + n] = mad(q_w, E[(size_t(row) * 6 + n) * HW + k], th[size_t(tid) * 6 + n]);
          }
        R s[THREADS];
        for (int n = 0; n < 6; n++) {
          for (int t = 0; t < THREADS; t++) s[t] = th[size_t(t) * 6 + n];
          sv[size_t(row) * 6 + n] = block_reduce(s);
        }
      }
      if (opt.damp_before_schur)
        for (int i = 0; i < n6; i++) A[size_t(i) * n6 + i] += double(ep) + double(lm) * A[size_t(i) * n6 + i];
      // (A - S)
      for (size_t pidx = 0; pidx < pairs.size(); pidx++)
        for (int k = 0; k < 6; k++)
          for (int l = 0; l < 6; l++)
            A[size_t(6 * pairs[pidx].pi + k) * n6 + 6 * pairs[pidx].pj + l] -= double(Sblk[pidx * 36 + k * 6 + l]);
      for (int row = 0; row < P + N; row++) {
        const int i = int(jj_exp[row] - t0);
        if (i >= 0)
          for (int k = 0; k < 6; k++) b[6 * i + k] -= double(sv[size_t(row) * 6 + k]);
      }
      if (opt.dump_system && Hdump) { memcpy(Hdump, A.data(), A.size() * 8); memcpy(bdump, b.data(), b.size() * 8); }
      if (!chol_solve(A, b, n6, opt.damp_before_schur ? 0.0 : double(lm), opt.damp_before_schur ? 0.0 : double(ep), x)) status = 1;
      std::vector<R> dxf(n6);
      for (int n = 0; n < n6; n++) { dxf[n] = R(x[n]); dx[n] = double(dxf[n]); }

      // EvT6x1_kernel (:1095-1115) and dz (:1417)
      std::vector<R> dw(size_t(P + N) * HW, R(0));
#pragma omp parallel for schedule(dynamic, 16)
      for (int row = 0; row < P + N; row++) {
        const int ixp = int(jj_exp[row] - t0);
        const bool skip = opt.skip_pose0_backsub ? (ixp <= 0 || ixp >= P) : (ixp < 0 || ixp >= P);
        if (skip) continue;
        for (int k = 0; k < HW; k++) {
          R acc = 0;
          for (int n = 0; n < 6; n++) acc = mad(E[(size_t(row) * 6 + n) * HW + k], dxf[6 * ixp + n], acc);
          dw[size_t(row) * HW + k] = acc;
        }
      }
      std::vector<R> dwsum(size_t(M) * HW);
      accum<R>(dw.data(), P + N, HW, ii_exp.data(), kx.data(), M, dwsum.data());
      dz.resize(size_t(M) * HW);
      for (size_t o = 0; o < dz.size(); o++) dz[o] = Q[o] * (w[o] - dwsum[o]);
      // disp_retr_kernel (:933-946)
      for (int k = 0; k < M; k++)
        for (int p = 0; p < HW; p++) disps[size_t(kx[k]) * HW + p] += dz[size_t(k) * HW + p];
    }
    // pose_retr_kernel (:898-931)
    for (int k = t0; k < t1; k++) {
      R xi[6], t1v[3], q1[4];
      for (int n = 0; n < 6; n++) xi[n] = R(dx[size_t(k - t0) * 6 + n]);
      retrSE3(xi, poses + 7 * k, poses + 7 * k + 3, t1v, q1);
      for (int n = 0; n < 3; n++) poses[7 * k + n] = t1v[n];
      for (int n = 0; n < 4; n++) poses[7 * k + 3 + n] = q1[n];
    }
  }
  if (dx_out) for (int n = 0; n < n6; n++) dx_out[n] = R(dx[n]);
  if (dz_out && !motion_only) memcpy(dz_out, dz.data(), sizeof(R) * dz.size());
  return status;
}

inline bool within_bounds(int h, int w, int H, int W) { return h >= 0 && h < H && w >= 0 && w < W; }

// corr_index_forward_kernel, correlation_kernels.cu:20-70.  T = storage/arith type of the volume.
template <typename T>
void corr_index_fwd(const T* volume, const float* coords, T* corr, int N, int h1, int w1, int h2, int w2, int r) {
  const int rd = 2 * r + 1;
  const size_t plane = size_t(h1) * w1;
  memset(corr, 0, sizeof(T) * size_t(N) * rd * rd * plane);
#pragma omp parallel for collapse(2) schedule(static)
  for (int n = 0; n < N; n++)
    for (int y = 0; y < h1; y++)
      for (int x = 0; x < w1; x++) {
        const float x0 = coords[((size_t(n) * 2 + 0) * h1 + y) * w1 + x];
        const float y0 = coords[((size_t(n) * 2 + 1) * h1 + y) * w1 + x];
        const float dx = x0 - floorf(x0), dy = y0 - floorf(y0);
        const T* vol = volume + ((size_t(n) * h1 + y) * w1 + x) * size_t(h2) * w2;
        auto out = [&](int i, int j) -> T& { return corr[((size_t(n) * rd + i) * rd + j) * plane + size_t(y) * w1 + x]; };
        for (int i = 0; i < rd + 1; i++)
          for (int j = 0; j < rd + 1; j++) {
            const int x1 = int(floorf(x0)) - r + i, y1 = int(floorf(y0)) - r + j;
            if (within_bounds(y1, x1, h2, w2)) {
              const T s = vol[size_t(y1) * w2 + x1];
              if (i > 0 && j > 0) out(i - 1, j - 1) = T(float(out(i - 1, j - 1)) + float(T(float(s) * float(T(dx * dy)))));
              if (i > 0 && j < rd) out(i - 1, j) = T(float(out(i - 1, j)) + float(T(float(s) * float(T(dx * (1.0f - dy))))));
              if (i < rd && j > 0) out(i, j - 1) = T(float(out(i, j - 1)) + float(T(float(s) * float(T((1.0f - dx) * dy)))));
              if (i < rd && j < rd) out(i, j) = T(float(out(i, j)) + float(T(float(s) * float(T((1.0f - dx) * (1.0f - dy))))));
            }
          }
      }
}
// fp32 specialisation: `corr += s * w` is a single contracted FFMA in the compiled kernel.
template <>
void corr_index_fwd<float>(const float* volume, const float* coords, float* corr, int N, int h1, int w1, int h2, int w2, int r) {
  const int rd = 2 * r + 1;
  const size_t plane = size_t(h1) * w1;
  memset(corr, 0, sizeof(float) * size_t(N) * rd * rd * plane);
#pragma omp parallel for collapse(2) schedule(static)
  for (int n = 0; n < N; n++)
    for (int y = 0; y < h1; y++)
      for (int x = 0; x < w1; x++) {
        const float x0 = coords[((size_t(n) * 2 + 0) * h1 + y) * w1 + x];
        const float y0 = coords[((size_t(n) * 2 + 1) * h1 + y) * w1 + x];
        const float dx = x0 - floorf(x0), dy = y0 - floorf(y0);
        const float* vol = volume + ((size_t(n) * h1 + y) * w1 + x) * size_t(h2) * w2;
        auto out = [&](int i, int j) -> float& { return corr[((size_t(n) * rd + i) * rd + j) * plane + size_t(y) * w1 + x]; };
        for (int i = 0; i < rd + 1; i++)
          for (int j = 0; j < rd + 1; j++) {
            const int x1 = int(floorf(x0)) - r + i, y1 = int(floorf(y0)) - r + j;
            if (within_bounds(y1, x1, h2, w2)) {
              const float s = vol[size_t(y1) * w2 + x1];
              if (i > 0 && j > 0) out(i - 1, j - 1) = fmaf(s, dx * dy, out(i - 1, j - 1));
              if (i > 0 && j < rd) out(i - 1, j) = fmaf(s, dx * (1.0f - dy), out(i - 1, j));
              if (i < rd && j > 0) out(i, j - 1) = fmaf(s, (1.0f - dx) * dy, out(i, j - 1));
              if (i < rd && j < rd) out(i, j) = fmaf(s, (1.0f - dx) * (1.0f - dy), out(i, j));
            }
          }
      }
}

// corr_index_backward_kernel, correlation_kernels.cu:73-124 (fp32)
void corr_index_bwd(const float* coords, const float* corr_grad, float* volume_grad, int N, int h1, int w1, int h2, int w2, int r) {
  const int rd = 2 * r + 1;
  const size_t plane = size_t(h1) * w1;
  memset(volume_grad, 0, sizeof(float) * size_t(N) * plane * h2 * w2);
#pragma omp parallel for collapse(2) schedule(static)
  for (int n = 0; n < N; n++)
    for (int y = 0; y < h1; y++)
      for (int x = 0; x < w1; x++) {
        const float x0 = coords[((size_t(n) * 2 + 0) * h1 + y) * w1 + x];
        const float y0 = coords[((size_t(n) * 2 + 1) * h1 + y) * w1 + x];
        const float dx = x0 - floorf(x0), dy = y0 - floorf(y0);
        float* vg = volume_grad + ((size_t(n) * h1 + y) * w1 + x) * size_t(h2) * w2;
        auto gin = [&](int i, int j) -> float { return corr_grad[((size_t(n) * rd + i) * rd + j) * plane + size_t(y) * w1 + x]; };
        for (int i = 0; i < rd + 1; i++)
          for (int j = 0; j < rd + 1; j++) {
            const int x1 = int(floorf(x0)) - r + i, y1 = int(floorf(y0)) - r + j;
            if (within_bounds(y1, x1, h2, w2)) {
              float g = 0.0f;
              if (i > 0 && j > 0) g = fmaf(gin(i - 1, j - 1), dx * dy, g);
              if (i > 0 && j < rd) g = fmaf(gin(i - 1, j), dx * (1.0f - dy), g);
              if (i < rd && j > 0) g = fmaf(gin(i, j - 1), (1.0f - dx) * dy, g);
              if (i < rd && j < rd) g = fmaf(gin(i, j), (1.0f - dx) * (1.0f - dy), g);
              vg[size_t(y1) * w2 + x1] += g;
            }
          }
      }
}

}  // namespace

// =====================================================================================
// exported C API (ctypes)
// =====================================================================================

ORACLE_API int oracle_ba_f32(float* poses, float* disps, const float* intr, const float* disps_sens,
                             const float* target, const float* weight, const float* eta,
                             const int64_t* ii, const int64_t* jj, int N, int B, int ht, int wd,
                             int t0, int t1, int iterations, float lm, float ep, int motion_only,
                             int eta_rows, float* dx_out, float* dz_out, int flags,
                             double* Hdump, double* bdump) {
  BAOptions o{flags & 1, Hdump != nullptr, (flags >> 1) & 1, (flags >> 2) & 1};
  return ba_impl<float>(poses, disps, intr, disps_sens, target, weight, eta, ii, jj, N, B, ht, wd, t0,
                        t1, iterations, lm, ep, motion_only, eta_rows, dx_out, dz_out, o, Hdump, bdump);
}

// same algorithm with every "device" quantity in fp64 — used to measure how much of a
// CUDA-vs-oracle difference is plain fp32 round-off (tests/test_oracle.py).
ORACLE_API int oracle_ba_f64(double* poses, double* disps, const double* intr, const double* disps_sens,
                             const double* target, const double* weight, const double* eta,
                             const int64_t* ii, const int64_t* jj, int N, int B, int ht, int wd,
                             int t0, int t1, int iterations, float lm, float ep, int motion_only,
                             int eta_rows, double* dx_out, double* dz_out, int flags) {
  BAOptions o{flags & 1, 0, (flags >> 1) & 1, (flags >> 2) & 1};
  return ba_impl<double>(poses, disps, intr, disps_sens, target, weight, eta, ii, jj, N, B, ht, wd, t0,
                         t1, iterations, lm, ep, motion_only, eta_rows, dx_out, dz_out, o, nullptr, nullptr);
}

// frame_distance_kernel, droid_kernels.cu:518-657.
// The consumers of this op threshold and argsort its fp32 output (factor_graph.py:327-354), so the
// restatement fixes not only the summation order (thread-strided accumulation + blockReduce tree)
// but also the FMA contraction nvcc applies to the source expressions.  The pattern below was read
// off the SASS of the reference kernel built for sm_100a with the reference's flags (-O3, default
// -fmad=true): `a*b - c*d` -> fma(a,b,-(c*d)); `X + q3*uv + (..)` -> fma(q3,uv,X) + (..);
// `Y += X3*t` -> separate multiply and add (the product is shared with the translation-only branch);
// `fx*(x/z) + cx - u` -> fma(fx, x/z, cx) - u; `du*du + dv*dv` -> fma(du,du,dv*dv);
// `accum += beta*d` -> fma.  Division and sqrt are IEEE (no fast-math).
static inline void fd_rot(const float* q, const float* X, float* Y) {  // actSO3 as compiled
  float uv0 = fmaf(q[1], X[2], -(q[2] * X[1])); uv0 += uv0;
  float uv1 = fmaf(q[2], X[0], -(q[0] * X[2])); uv1 += uv1;
  float uv2 = fmaf(q[0], X[1], -(q[1] * X[0])); uv2 += uv2;
  Y[0] = fmaf(q[3], uv0, X[0]) + fmaf(q[1], uv2, -(q[2] * uv1));
  Y[1] = fmaf(q[3], uv1, X[1]) + fmaf(q[2], uv0, -(q[0] * uv2));
  Y[2] = fmaf(q[3], uv2, X[2]) + fmaf(q[0], uv1, -(q[1] * uv0));
}
static inline void fd_relSE3(const float* ti, const float* qi, const float* tj, const float* qj, float* tij, float* qij) {
  qij[0] = fmaf(qj[2], qi[1], fmaf(-qj[1], qi[2], fmaf(qj[0], qi[3], -(qj[3] * qi[0]))));
  qij[1] = fmaf(qj[0], qi[2], fmaf(-qj[2], qi[0], fmaf(qj[1], qi[3], -(qj[3] * qi[1]))));
  qij[2] = fmaf(qj[1], qi[0], fmaf(-qj[0], qi[1], fmaf(qj[2], qi[3], -(qj[3] * qi[2]))));
  qij[3] = fmaf(qj[2], qi[2], fmaf(qj[1], qi[1], fmaf(qj[3], qi[3], qj[0] * qi[0])));
  float Y[3];
  fd_rot(qij, ti, Y);
  tij[0] = tj[0] - Y[0]; tij[1] = tj[1] - Y[1]; tij[2] = tj[2] - Y[2];
}
ORACLE_API void oracle_frame_distance(const float* poses, const float* disps, const float* intr,
                                      const int64_t* ii, const int64_t* jj, int K, int ht, int wd,
                                      float beta, float* dist) {
  const int HW = ht * wd;
  const float fx = intr[0], fy = intr[1], cx = intr[2], cy = intr[3];
  const float omb = 1 - beta;
#pragma omp parallel for schedule(dynamic, 64)
  for (int b = 0; b < K; b++) {
    const int ix = int(ii[b]), jx = int(jj[b]);
    float tij[3], q[4];
    fd_relSE3(poses + 7 * ix, poses + 7 * ix + 3, poses + 7 * jx, poses + 7 * jx + 3, tij, q);
    float accum[THREADS], valid[THREADS], total[THREADS];
    for (int t = 0; t < THREADS; t++) { accum[t] = 0; valid[t] = 0; total[t] = 0; }
    for (int tid = 0; tid < THREADS; tid++) {
      for (int k = tid; k < HW; k += THREADS) {
        const int i = k / wd, j = k % wd;
        const float u = float(j), v = float(i);
        const float X0 = (u - cx) / fx, X1 = (v - cy) / fy, dsp = disps[size_t(ix) * HW + k];
        // actSO3 with X[2] == 1 folded by the compiler
        float uv0 = fmaf(-q[2], X1, q[1]); uv0 += uv0;
        float uv1 = fmaf(q[2], X0, -q[0]); uv1 += uv1;
        float uv2 = fmaf(q[0], X1, -(q[1] * X0)); uv2 += uv2;
        float Y0 = fmaf(q[3], uv0, X0) + fmaf(q[1], uv2, -(q[2] * uv1));
        float Y1 = fmaf(q[3], uv1, X1) + fmaf(q[2], uv0, -(q[0] * uv2));
        float Y2 = fmaf(q[3], uv2, 1.0f) + fmaf(q[0], uv1, -(q[1] * uv0));
        const float p0 = tij[0] * dsp, p1 = tij[1] * dsp, p2 = tij[2] * dsp;
        Y0 += p0; Y1 += p1; Y2 += p2;
        float du = fmaf(fx, Y0 / Y2, cx) - u;
        float dv = fmaf(fy, Y1 / Y2, cy) - v;
        float d = sqrtf(fmaf(du, du, dv * dv));
        total[tid] += beta;
        if (Y2 > 0.25f) { accum[tid] = fmaf(d, beta, accum[tid]); valid[tid] += beta; }
        const float Z0 = p0 + X0, Z1 = p1 + X1, Z2 = p2 + 1.0f;
        du = fmaf(fx, Z0 / Z2, cx) - u;
        dv = fmaf(fy, Z1 / Z2, cy) - v;
        d = sqrtf(fmaf(du, du, dv * dv));
        total[tid] += omb;
        if (Z2 > 0.25f) { accum[tid] = fmaf(omb, d, accum[tid]); valid[tid] += omb; }
      }
    }
    const float a = block_reduce(accum), tt = block_reduce(total), vv = block_reduce(valid);
    dist[b] = (double(vv) / (double(tt) + 1e-8) < 0.75) ? 1000.0f : a / vv;
  }
}

ORACLE_API void oracle_corr_index_forward_f32(const float* volume, const float* coords, float* corr,
                                              int N, int h1, int w1, int h2, int w2, int r) {
  corr_index_fwd<float>(volume, coords, corr, N, h1, w1, h2, w2, r);
}
// fp16 volumes are passed as raw uint16 bit patterns (numpy float16 storage).
ORACLE_API void oracle_corr_index_forward_f16(const uint16_t* volume, const float* coords, uint16_t* corr,
                                              int N, int h1, int w1, int h2, int w2, int r) {
  corr_index_fwd<_Float16>(reinterpret_cast<const _Float16*>(volume), coords,
                           reinterpret_cast<_Float16*>(corr), N, h1, w1, h2, w2, r);
}
ORACLE_API void oracle_corr_index_backward_f32(const float* coords, const float* corr_grad, float* volume_grad,
                                               int N, int h1, int w1, int h2, int w2, int r) {
  corr_index_bwd(coords, corr_grad, volume_grad, N, h1, w1, h2, w2, r);
}

// altcorr_forward_kernel, altcorr_kernel.cu:27-149 (fp32).  fmap1[B,H1,W1,C], fmap2[B,H2,W2,C],
// coords[B,S,H1,W1,2] -> corr[B,S,(2r+1)^2,H1,W1].  Channels are consumed in slabs of 32
// (CHANNEL_STRIDE) and every slab's partial dot product is splatted separately.
ORACLE_API void oracle_altcorr_forward(const float* fmap1, const float* fmap2, const float* coords, float* corr,
                                       int B, int S, int H1, int W1, int H2, int W2, int C, int r) {
  const int rd = 2 * r + 1;
  const size_t plane = size_t(H1) * W1;
  memset(corr, 0, sizeof(float) * size_t(B) * S * rd * rd * plane);
#pragma omp parallel for collapse(2) schedule(static)
  for (int b = 0; b < B; b++)
    for (int h1 = 0; h1 < H1; h1++)
      for (int w1 = 0; w1 < W1; w1++) {
        const float* f1 = fmap1 + ((size_t(b) * H1 + h1) * W1 + w1) * C;
        for (int c = 0; c < C; c += 32)
          for (int n = 0; n < S; n++) {
            const float* cp = coords + (((size_t(b) * S + n) * H1 + h1) * W1 + w1) * 2;
            const float x2 = cp[0], y2 = cp[1];
            const float dx = x2 - floorf(x2), dy = y2 - floorf(y2);
            float* out = corr + (size_t(b) * S + n) * rd * rd * plane + size_t(h1) * W1 + w1;
            for (int iy = 0; iy < rd + 1; iy++)
              for (int ix = 0; ix < rd + 1; ix++) {
                const int h2 = int(floorf(y2)) - r + iy, w2 = int(floorf(x2)) - r + ix;
                float s = 0.0f;
                if (within_bounds(h2, w2, H2, W2)) {
                  const float* f2 = fmap2 + ((size_t(b) * H2 + h2) * W2 + w2) * C;
                  for (int k = 0; k < 32 && c + k < C; k++) s = fmaf(f1[c + k], f2[c + k], s);
                }
                const float nw = s * (dy * dx), ne = s * (dy * (1 - dx));
                const float sw = s * ((1 - dy) * dx), se = s * ((1 - dy) * (1 - dx));
                if (iy > 0 && ix > 0) out[size_t((iy - 1) + rd * (ix - 1)) * plane] += nw;
                if (iy > 0 && ix < rd) out[size_t((iy - 1) + rd * ix) * plane] += ne;
                if (iy < rd && ix > 0) out[size_t(iy + rd * (ix - 1)) * plane] += sw;
                if (iy < rd && ix < rd) out[size_t(iy + rd * ix) * plane] += se;
              }
          }
      }
}

// altcorr_backward_kernel, altcorr_kernel.cu:152-286.  coords_grad is returned as zeros by the
// reference (:340,355), so it is not an output here.  Summation into fmap2_grad uses atomics in the
// reference (order undefined); here the order is (h1,w1,slab,n,iy,ix) — compare with a tolerance.
ORACLE_API void oracle_altcorr_backward(const float* fmap1, const float* fmap2, const float* coords,
                                        const float* corr_grad, float* fmap1_grad, float* fmap2_grad,
                                        int B, int S, int H1, int W1, int H2, int W2, int C, int r) {
  const int rd = 2 * r + 1;
  const size_t plane = size_t(H1) * W1;
  memset(fmap1_grad, 0, sizeof(float) * size_t(B) * H1 * W1 * C);
  memset(fmap2_grad, 0, sizeof(float) * size_t(B) * H2 * W2 * C);
#pragma omp parallel for schedule(static)
  for (int b = 0; b < B; b++)
    for (int h1 = 0; h1 < H1; h1++)
      for (int w1 = 0; w1 < W1; w1++) {
        const float* f1 = fmap1 + ((size_t(b) * H1 + h1) * W1 + w1) * C;
        float* g1 = fmap1_grad + ((size_t(b) * H1 + h1) * W1 + w1) * C;
        for (int n = 0; n < S; n++) {
          const float* cp = coords + (((size_t(b) * S + n) * H1 + h1) * W1 + w1) * 2;
          const float x2 = cp[0], y2 = cp[1];
          const float dx = x2 - floorf(x2), dy = y2 - floorf(y2);
          const float* gin = corr_grad + (size_t(b) * S + n) * rd * rd * plane + size_t(h1) * W1 + w1;
          for (int iy = 0; iy < rd + 1; iy++)
            for (int ix = 0; ix < rd + 1; ix++) {
              const int h2 = int(floorf(y2)) - r + iy, w2 = int(floorf(x2)) - r + ix;
              float g = 0.0f;
              if (iy > 0 && ix > 0) g += gin[size_t((iy - 1) + rd * (ix - 1)) * plane] * dy * dx;
              if (iy > 0 && ix < rd) g += gin[size_t((iy - 1) + rd * ix) * plane] * dy * (1 - dx);
              if (iy < rd && ix > 0) g += gin[size_t(iy + rd * (ix - 1)) * plane] * (1 - dy) * dx;
              if (iy < rd && ix < rd) g += gin[size_t(iy + rd * ix) * plane] * (1 - dy) * (1 - dx);
              if (!within_bounds(h2, w2, H2, W2)) continue;
              const float* f2 = fmap2 + ((size_t(b) * H2 + h2) * W2 + w2) * C;
              float* g2 = fmap2_grad + ((size_t(b) * H2 + h2) * W2 + w2) * C;
              for (int k = 0; k < C; k++) { g1[k] = fmaf(g, f2[k], g1[k]); g2[k] = fmaf(g, f1[k], g2[k]); }
            }
        }
      }
}

// CorrBlock.corr + pyramid, droid_slam/modules/corr.py:30-38,63-71, fp32 path:
// corr[n,p1,p2] = sum_c (f1[n,c,p1]/4)(f2[n,c,p2]/4); level l+1 = avg_pool2d(level l, 2) over (h2,w2).
ORACLE_API void oracle_corr_volume_f32(const float* fmap1, const float* fmap2, int N, int C, int ht, int wd,
                                       float* l0, float* l1, float* l2, float* l3) {
  const int HW = ht * wd;
#pragma omp parallel for collapse(2) schedule(static)
  for (int n = 0; n < N; n++)
    for (int p1 = 0; p1 < HW; p1++) {
      float* row = l0 + (size_t(n) * HW + p1) * HW;
      for (int p2 = 0; p2 < HW; p2++) row[p2] = 0.0f;
      for (int c = 0; c < C; c++) {
        const float a = fmap1[(size_t(n) * C + c) * HW + p1] / 4.0f;
        const float* f2 = fmap2 + (size_t(n) * C + c) * HW;
        for (int p2 = 0; p2 < HW; p2++) row[p2] = fmaf(a, f2[p2] / 4.0f, row[p2]);
      }
    }
  float* lv[4] = {l0, l1, l2, l3};
  for (int l = 1; l < 4; l++) {
    if (!lv[l]) break;
    const int hs = ht >> (l - 1), ws = wd >> (l - 1), hd = ht >> l, wdd = wd >> l;
#pragma omp parallel for schedule(static)
    for (long np = 0; np < long(N) * HW; np++) {
      const float* src = lv[l - 1] + size_t(np) * hs * ws;
      float* dst = lv[l] + size_t(np) * hd * wdd;
      for (int y = 0; y < hd; y++)
        for (int x = 0; x < wdd; x++)
          dst[y * wdd + x] = (src[(2 * y) * ws + 2 * x] + src[(2 * y) * ws + 2 * x + 1] +
                              src[(2 * y + 1) * ws + 2 * x] + src[(2 * y + 1) * ws + 2 * x + 1]) * 0.25f;
    }
  }
}

// iproj_kernel, droid_kernels.cu:779-850: points[n,i,j,:] = (G_n * X)[0:3] / X[3]
ORACLE_API void oracle_iproj(const float* poses, const float* disps, const float* intr, int num, int ht, int wd, float* points) {
  const int HW = ht * wd;
  const float fx = intr[0], fy = intr[1], cx = intr[2], cy = intr[3];
  for (int n = 0; n < num; n++)
    for (int k = 0; k < HW; k++) {
      float Xi[4] = {(float(k % wd) - cx) / fx, (float(k / wd) - cy) / fy, 1.0f, disps[size_t(n) * HW + k]}, Xj[4];
      actSE3(poses + 7 * n, poses + 7 * n + 3, Xi, Xj);
      for (int c = 0; c < 3; c++) points[(size_t(n) * HW + k) * 3 + c] = Xj[c] / Xj[3];
    }
}

// depth_filter_kernel, droid_kernels.cu:661-775: count of the 6 temporal neighbours whose disparity agrees
ORACLE_API void oracle_depth_filter(const float* poses, const float* disps, const float* intr, const int64_t* inds,
                                    const float* thresh, int nidx, int num, int ht, int wd, float* counter) {
  const int HW = ht * wd;
  const float fx = intr[0], fy = intr[1], cx = intr[2], cy = intr[3];
  memset(counter, 0, sizeof(float) * size_t(nidx) * HW);
  for (int b = 0; b < nidx; b++)
    for (int nb = 0; nb < 6; nb++) {
      const int ix = int(inds[b]);
      const int jx = (nb < 3) ? ix - nb - 1 : ix + nb;
      if (jx < 0 || jx >= num) continue;
      const float t = thresh[b];
      float tij[3], qij[4];
      relSE3(poses + 7 * ix, poses + 7 * ix + 3, poses + 7 * jx, poses + 7 * jx + 3, tij, qij);
      for (int k = 0; k < HW; k++) {
        const int i = k / wd, j = k % wd;
        float Xi[4] = {(float(j) - cx) / fx, (float(i) - cy) / fy, 1.0f, disps[size_t(ix) * HW + k]}, Xj[4];
        actSE3(tij, qij, Xi, Xj);
        const float uj = fmaf(fx, Xj[0] / Xj[2], cx), vj = fmaf(fy, Xj[1] / Xj[2], cy);
        const float dj = Xj[3] / Xj[2];
        const int u0 = int(floorf(uj)), v0 = int(floorf(vj));
        if (u0 >= 0 && v0 >= 0 && u0 < wd - 1 && v0 < ht - 1) {
          const float* dn = disps + size_t(jx) * HW;
          const float d00 = dn[v0 * wd + u0], d01 = dn[v0 * wd + u0 + 1];
          const float d10 = dn[(v0 + 1) * wd + u0], d11 = dn[(v0 + 1) * wd + u0 + 1];
          const double idj = 1.0 / double(dj);
          if (fabs(idj - 1.0 / double(d00)) < double(t) || fabs(idj - 1.0 / double(d01)) < double(t) ||
              fabs(idj - 1.0 / double(d10)) < double(t) || fabs(idj - 1.0 / double(d11)) < double(t))
            counter[size_t(b) * HW + k] += 1.0f;
        }
      }
    }
}

// SE3 retraction exposed for the shim tests (lietorch property checks)
ORACLE_API void oracle_se3_retr(const float* xi, const float* pose, float* out) {
  retrSE3(xi, pose, pose + 3, out, out + 3);
}
ORACLE_API void oracle_se3_rel(const float* pi, const float* pj, float* out) {
  relSE3(pi, pi + 3, pj, pj + 3, out, out + 3);
}
ORACLE_API int oracle_num_threads() {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}
