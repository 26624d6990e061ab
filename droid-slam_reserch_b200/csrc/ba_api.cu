// ba_api.cu — C-ABI of the dense bundle adjustment (see include/droid_b200.h).
// Mirrors droid_backends.ba (/root/reference/src/droid.cpp:88-117, droid_kernels.cu:1314-1434): in-place
// update of poses/disps, `iterations` Gauss-Newton steps, returns dx/dz of the last step.
#include <cstdlib>
#include <atomic>
#include <cstdio>
#include "ba.cuh"

using namespace db200;

// ---- instrumentation: launch counter + optional per-stage CUDA-event timing (bench.py) ----------------
static std::atomic<long long> g_launches{0};
extern "C" void db200_count_launches(int n) { g_launches.fetch_add(n, std::memory_order_relaxed); }
extern "C" long long db200_launch_count(void) { return g_launches.load(); }

namespace {
constexpr int kStages = DB200_NUM_STAGES;
constexpr int kProfCap = 2048;                 // (start, stop) pairs per stage between two collects
struct Prof {
  bool on = false;
  cudaEvent_t ev[kStages][kProfCap][2];
  int used[kStages] = {};
  int made[kStages] = {};
} g_prof;

struct StageTimer {                             // records start/stop around one stage when profiling is on
  int s; cudaStream_t st; bool live;
  StageTimer(int stage, cudaStream_t stream) : s(stage), st(stream), live(false) {
    if (!g_prof.on || g_prof.used[s] >= kProfCap) return;
    const int i = g_prof.used[s];
    if (i >= g_prof.made[s]) {
      if (cudaEventCreate(&g_prof.ev[s][i][0]) != cudaSuccess || cudaEventCreate(&g_prof.ev[s][i][1]) != cudaSuccess) return;
      g_prof.made[s] = i + 1;
    }
    cudaEventRecord(g_prof.ev[s][i][0], st);
    live = true;
  }
  ~StageTimer() {
    if (!live) return;
    cudaEventRecord(g_prof.ev[s][g_prof.used[s]][1], st);
    g_prof.used[s]++;
  }
};
}  // namespace

extern "C" void db200_profile_enable(int on) {
  g_prof.on = on != 0;
  for (int s = 0; s < kStages; s++) g_prof.used[s] = 0;
}

extern "C" int db200_profile_collect(double* ms_out, int* count_out) {
  for (int s = 0; s < kStages; s++) {
    double tot = 0.0;
    for (int i = 0; i < g_prof.used[s]; i++) {
      DB200_CUDA(cudaEventSynchronize(g_prof.ev[s][i][1]));
      float ms = 0.0f;
      DB200_CUDA(cudaEventElapsedTime(&ms, g_prof.ev[s][i][0], g_prof.ev[s][i][1]));
      tot += ms;
    }
    if (ms_out) ms_out[s] = tot;
    if (count_out) count_out[s] = g_prof.used[s];
    g_prof.used[s] = 0;
  }
  return DB200_OK;
}


namespace {

bool args_ok(const float* poses, const float* disps, const float* intr, const float* sens, const float* tg,
             const float* wg, const float* eta, const int64_t* ii, const int64_t* jj, int N, int B, int ht, int wd, int t0,
             int t1, int motion_only, int eta_rows) {
  if (!poses || !disps || !intr || N < 0 || B <= 0 || ht <= 0 || wd <= 0) return false;
  if (t0 < 0 || t1 <= t0 || t1 > B) return false;
  if (N > 0 && (!tg || !wg || !ii || !jj)) return false;
  if (!motion_only && (!sens || !eta || eta_rows <= 0 || eta_rows > B)) return false;
  return true;
}

BAParams make_params(float* poses, float* disps, const float* intr, const float* sens, const float* tg, const float* wg,
                     const float* eta, const int64_t* ii, const int64_t* jj, int N, int B, int ht, int wd, int t0, int t1,
                     int motion_only, int eta_rows, void* ws, const BALayout& L) {
  BAParams p = ba_params(L, ws);
  p.poses = poses; p.disps = disps; p.intr = intr; p.sens = sens; p.target = tg; p.weight = wg; p.eta = eta;
  p.ii = ii; p.jj = jj; p.N = N; p.B = B; p.ht = ht; p.wd = wd; p.HW = ht * wd; p.t0 = t0; p.t1 = t1; p.P = t1 - t0;
  p.eta_rows = eta_rows; p.motion_only = motion_only;
  return p;
}

}  // namespace

extern "C" size_t db200_ba_workspace_bytes(int N, int B, int ht, int wd, int t0, int t1, int eta_rows, int motion_only) {
  if (N < 0 || B <= 0 || ht <= 0 || wd <= 0 || t1 <= t0) return 0;
  return ba_layout(N, B, ht * wd, t0, t1, eta_rows, motion_only).total;
}

extern "C" double* db200_ba_system_ptr(void* workspace, int N, int B, int ht, int wd, int t0, int t1, int eta_rows,
                                       int motion_only) {
  if (!workspace) return nullptr;
  const BALayout L = ba_layout(N, B, ht * wd, t0, t1, eta_rows, motion_only);
  return reinterpret_cast<double*>(static_cast<char*>(workspace) + L.Hb);
}

extern "C" size_t db200_ba_system_elems(int t0, int t1) {
  return sys_elems(6 * (t1 - t0));
}

extern "C" int db200_ba_accumulate(float* poses, float* disps, const float* intrinsics, const float* disps_sens,
                                   const float* targets, const float* weights, const float* eta, const int64_t* ii,
                                   const int64_t* jj, int N, int B, int ht, int wd, int t0, int t1, int motion_only,
                                   int eta_rows, int first_iteration, int own_lo, int own_hi, void* workspace,
                                   size_t workspace_bytes, db200_stream_t stream) {
  if (!workspace || !args_ok(poses, disps, intrinsics, disps_sens, targets, weights, eta, ii, jj, N, B, ht, wd, t0, t1,
                             motion_only, eta_rows))
    return DB200_ERR_ARG;
  const BALayout L = ba_layout(N, B, ht * wd, t0, t1, eta_rows, motion_only);
  if (workspace_bytes < L.total) return DB200_ERR_WORKSPACE;
  BAParams p = make_params(poses, disps, intrinsics, disps_sens, targets, weights, eta, ii, jj, N, B, ht, wd, t0, t1,
                           motion_only, eta_rows, workspace, L);
  if (own_lo < 0 || own_hi > B || own_lo > own_hi) return DB200_ERR_ARG;
  p.own_lo = own_lo; p.own_hi = own_hi;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  int rc = DB200_OK;
  if (first_iteration) { StageTimer t(DB200_STAGE_GRAPH, st); rc = ba_build_graph(p, st); }
  if (rc == DB200_OK) { StageTimer t(DB200_STAGE_PREP, st); rc = ba_edge_prep(p, st); }
  if (rc == DB200_OK) { StageTimer t(DB200_STAGE_ACCUM, st); rc = ba_accumulate_frames(p, st); }
  return rc;
}

extern "C" int db200_ba_solve_update(float* poses, float* disps, const float* intrinsics, const float* targets,
                                     const float* weights, const int64_t* ii, const int64_t* jj, int N, int B, int ht,
                                     int wd, int t0, int t1, float lm, float ep, int motion_only, int eta_rows,
                                     int own_lo, int own_hi, float* dx_out, float* dz_out, void* workspace,
                                     size_t workspace_bytes, db200_stream_t stream) {
  if (!workspace || !poses || !disps || !intrinsics || N < 0 || B <= 0 || ht <= 0 || wd <= 0 || t0 < 0 || t1 <= t0 || t1 > B)
    return DB200_ERR_ARG;
  const BALayout L = ba_layout(N, B, ht * wd, t0, t1, eta_rows, motion_only);
  if (workspace_bytes < L.total) return DB200_ERR_WORKSPACE;
  BAParams p = make_params(poses, disps, intrinsics, nullptr, targets, weights, nullptr, ii, jj, N, B, ht, wd, t0, t1,
                           motion_only, eta_rows, workspace, L);
  if (own_lo < 0 || own_hi > B || own_lo > own_hi) return DB200_ERR_ARG;
  p.own_lo = own_lo; p.own_hi = own_hi;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  int rc;
  { StageTimer t(DB200_STAGE_SOLVE, st); rc = ba_solve(p, lm, ep, st); }
  if (rc == DB200_OK) { StageTimer t(DB200_STAGE_POST, st); rc = ba_post_solve(p, dx_out, st); }
  if (rc == DB200_OK && !motion_only) { StageTimer t(DB200_STAGE_BACKSUB, st); rc = ba_backsub_frames(p, dz_out, st); }
  return rc;
}

extern "C" int db200_ba(float* poses, float* disps, const float* intrinsics, const float* disps_sens, const float* targets,
                        const float* weights, const float* eta, const int64_t* ii, const int64_t* jj, int N, int B, int ht,
                        int wd, int t0, int t1, int iterations, float lm, float ep, int motion_only, int eta_rows,
                        float* dx_out, float* dz_out, void* workspace, size_t workspace_bytes, db200_stream_t stream) {
  if (iterations < 0) return DB200_ERR_ARG;
  for (int it = 0; it < iterations; it++) {
    int rc = db200_ba_accumulate(poses, disps, intrinsics, disps_sens, targets, weights, eta, ii, jj, N, B, ht, wd, t0, t1,
                                 motion_only, eta_rows, it == 0, 0, B, workspace, workspace_bytes, stream);
    if (rc != DB200_OK) return rc;
    const bool last = (it == iterations - 1);
    rc = db200_ba_solve_update(poses, disps, intrinsics, targets, weights, ii, jj, N, B, ht, wd, t0, t1, lm, ep, motion_only,
                               eta_rows, 0, B, last ? dx_out : nullptr, last ? dz_out : nullptr, workspace, workspace_bytes, stream);
    if (rc != DB200_OK) return rc;
  }
  return DB200_OK;
}

extern "C" int db200_ba_status(const void* workspace, db200_stream_t stream) {
  if (!workspace) return DB200_ERR_ARG;
  BAHeader h{};
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  DB200_CUDA(cudaMemcpyAsync(&h, workspace, sizeof(BAHeader), cudaMemcpyDeviceToHost, st));
  DB200_CUDA(cudaStreamSynchronize(st));
  return h.status;
}

static char g_last_err[256] = "none";
extern "C" void db200_note_cuda_error(int code, const char* file, int line) {
  snprintf(g_last_err, sizeof(g_last_err), "%s (%d) at %s:%d", cudaGetErrorString(cudaError_t(code)), code, file, line);
}
extern "C" const char* db200_last_cuda_error(void) { return g_last_err; }

extern "C" int db200_version(void) { return 100; }

extern "C" const char* db200_status_string(int s) {
  switch (s) {
    case DB200_OK: return "ok";
    case 1: return "Cholesky breakdown in at least one iteration: zero pose update applied (reference behaviour)";
    case DB200_ERR_ARG: return "invalid argument (null pointer, bad size, unsupported dtype/radius)";
    case DB200_ERR_CUDA: return "CUDA runtime error";
    case DB200_ERR_WORKSPACE: return "workspace too small";
    case DB200_ERR_INDEX: return "edge index out of range [0, t1)";
    case DB200_ERR_STATIC_LIMIT: return "problem exceeds the built-in static scratch of this no-allocation entry point (see db200_*_static_max_bytes)";
    case DB200_ERR_ETA: return "eta rows != number of unique frames in cat(arange(t0,t1), ii)";
  }
  return "unknown status";
}
