// graph_edges.cu — factor-graph edge selection on the GPU.
// Replaces the host loops of /root/reference/droid_slam/factor_graph.py:315-379 (`add_proximity_factors`: pairwise
// frame distances, masking, non-maximum suppression around existing and newly picked edges, greedy pick in order of
// distance) and :44-55 (`__filter_repeated_edges`).  The result is an EDGE LIST — integer work whose expected output is
// exact: the order of `es` and every (i, j) in it equal the reference's.
//
// Pipeline (all on the caller's stream, no host round trip):
//   1  pair indices of meshgrid(arange(t0,t), arange(t1,t)) and both directions of frame_distance (bit-exact kernel of
//      geom.cu);  d = 0.5f * (d1 + d2)                                     (depth_video.py:171-181)
//   2  d[ii - rad < jj] = inf; d[d > 100] = inf                            (factor_graph.py:328-329)
//   3  suppression around the existing edges (active + bad + inactive)      (:331-341)
//   4  neighbourhood edges (i,i) [stereo], (i,j),(j,i) for j in [max(i-rad-1,0), i): written straight into `es` at
//      closed-form offsets; their cells set to inf with Python's negative-index wrap   (:344-353)
//   5  stable radix sort of (d, cell)                                        (:355; ties — which torch.argsort leaves
//      unspecified — resolve to the lower cell index)
//   6  greedy pick by ONE warp: 32 sorted candidates per step are resolved in order with ballots/shuffles (a pick
//      suppresses later lanes of the same batch in registers and later batches through `d`)   (:356-378)
#include <cub/device/device_radix_sort.cuh>

#include "common.cuh"

namespace db200 {

struct ProxParams {
  int t0, t1, t, rad, nms, stereo, max_factors;
  double thresh;    // compared in double like Python's `d[k].item() > thresh`
  int cols;          // t - t1
  int K;             // (t - t0) * (t - t1)
  int n_init;        // neighbourhood edges
};

__device__ __forceinline__ float pos_inf() { return __int_as_float(0x7f800000); }

__global__ void prox_pairs_kernel(ProxParams p, int64_t* __restrict__ ii, int64_t* __restrict__ jj) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= p.K) return;
  const int i = p.t0 + k / p.cols, j = p.t1 + k % p.cols;
  // cells that factor_graph.py:328 overwrites with inf (ii - rad < jj) are never looked at: mark them invalid so that
  // frame_distance skips them (it returns NaN there; prox_mask_kernel writes the inf)
  const bool masked = (i - p.rad < j);
  ii[k] = masked ? -1 : i;
  jj[k] = masked ? -1 : j;
}

// steps 1 (average) + 2 + cell index values for the sort
__global__ void prox_mask_kernel(ProxParams p, const float* __restrict__ d1, const float* __restrict__ d2,
                                 float* __restrict__ d, int* __restrict__ cell) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= p.K) return;
  const int i = p.t0 + k / p.cols, j = p.t1 + k % p.cols;
  float v = __fmul_rn(0.5f, __fadd_rn(d1[k], d2[k]));
  if (i - p.rad < j) v = pos_inf();
  if (v > 100.0f) v = pos_inf();
  d[k] = v;
  cell[k] = k;
}

__device__ __forceinline__ int nms_reach(int i, int j, int nms) {
  int a = i - j; a = a < 0 ? -a : a;
  return max(min(a - 2, nms), 0);
}

// every (di,dj) with |di|+|dj| <= reach inside the candidate rectangle -> inf       (factor_graph.py:333-341, 370-378)
__device__ __forceinline__ void suppress_cell(const ProxParams& p, float* d, int i, int j, int di, int dj, int reach) {
  const int adi = di < 0 ? -di : di, adj = dj < 0 ? -dj : dj;
  if (adi + adj > reach) return;
  const int i1 = i + di, j1 = j + dj;
  if (i1 >= p.t0 && i1 < p.t && j1 >= p.t1 && j1 < p.t) d[(i1 - p.t0) * p.cols + (j1 - p.t1)] = pos_inf();
}

__global__ void prox_suppress_old_kernel(ProxParams p, const int64_t* __restrict__ io, const int64_t* __restrict__ jo, int n_old,
                                         float* __restrict__ d) {
  const int side = 2 * p.nms + 1, per = side * side;
  const long long gid = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (gid >= static_cast<long long>(n_old) * per) return;
  const int e = int(gid / per), c = int(gid % per);
  const int i = int(io[e]), j = int(jo[e]);
  suppress_cell(p, d, i, j, c / side - p.nms, c % side - p.nms, nms_reach(i, j, p.nms));
}

// edges before frame x in the neighbourhood list: sum_{i' in [t0, x)} (stereo + 2 * min(i', rad + 1))
__host__ __device__ inline long long prox_init_offset(int x, int t0, int r1, int stereo) {
  if (r1 < 0) r1 = 0;
  auto S = [r1](long long v) { return v <= r1 ? v * (v - 1) / 2 : static_cast<long long>(r1) * (r1 - 1) / 2 + (v - r1) * r1; };
  return static_cast<long long>(stereo) * (x - t0) + 2 * (S(x) - S(t0));
}

__global__ void prox_init_kernel(ProxParams p, float* __restrict__ d, int64_t* __restrict__ es) {
  const int i = p.t0 + blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= p.t) return;
  const int r1 = p.rad + 1 < 0 ? 0 : p.rad + 1;
  long long o = prox_init_offset(i, p.t0, r1, p.stereo);
  auto kill = [&](int j) {
    long long idx = static_cast<long long>(i - p.t0) * p.cols + (j - p.t1);
    if (idx < 0) idx += p.K;                                  // Python: d[negative] wraps once
    if (idx >= 0 && idx < p.K) d[idx] = pos_inf();
  };
  if (p.stereo) {
    es[2 * o] = i; es[2 * o + 1] = i; o++;
    kill(i);
  }
  for (int j = max(i - p.rad - 1, 0); j < i; j++) {
    es[2 * o] = i; es[2 * o + 1] = j; o++;
    es[2 * o] = j; es[2 * o + 1] = i; o++;
    kill(j);
  }
}

// one warp; sorted (key, cell); d is the live distance array
__global__ void __launch_bounds__(32)
prox_greedy_kernel(ProxParams p, const float* __restrict__ skey, const int* __restrict__ scell, float* __restrict__ d,
                   int64_t* __restrict__ es, int* __restrict__ n_out) {
  const int lane = threadIdx.x;
  int n = p.n_init;
  const int side = 2 * p.nms + 1, per = side * side;
  bool done = false;
  for (int base = 0; base < p.K && !done; base += 32) {
    const int idx = base + lane;
    int c = -1; float key = pos_inf(), cur = pos_inf();
    if (idx < p.K) { key = skey[idx]; c = scell[idx]; cur = d[c]; }
    // sorted ascending: once the first key of a batch exceeds the threshold nothing later can be picked
    const float key0 = __shfl_sync(0xffffffffu, key, 0);
    if (double(key0) > p.thresh || key0 != key0) break;
    const int i = c >= 0 ? p.t0 + c / p.cols : 0, j = c >= 0 ? p.t1 + c % p.cols : 0;
    bool alive = (c >= 0) && (double(cur) <= p.thresh);
    unsigned m = __ballot_sync(0xffffffffu, alive);
    while (m) {
      if (n > p.max_factors) { done = true; break; }            // factor_graph.py:360-361
      const int src = __ffs(m) - 1;
      const int pi = __shfl_sync(0xffffffffu, i, src), pj = __shfl_sync(0xffffffffu, j, src);
      if (lane == 0) {
        es[2 * n] = pi; es[2 * n + 1] = pj;
        es[2 * n + 2] = pj; es[2 * n + 3] = pi;
      }
      n += 2;
      const int reach = nms_reach(pi, pj, p.nms);
      for (int q = lane; q < per; q += 32) suppress_cell(p, d, pi, pj, q / side - p.nms, q % side - p.nms, reach);
      const int adi = i > pi ? i - pi : pi - i, adj = j > pj ? j - pj : pj - j;
      if (adi + adj <= reach) alive = false;                    // includes the picked cell itself (reach >= 0)
      m = __ballot_sync(0xffffffffu, alive);
    }
    __syncwarp();                                               // suppression writes ordered before the next batch's reads
  }
  if (lane == 0) *n_out = n;
}

__global__ void filter_repeated_kernel(const int64_t* __restrict__ in_i, const int64_t* __restrict__ in_j, int n_new,
                                       const int64_t* __restrict__ oi, const int64_t* __restrict__ oj, int n_old,
                                       unsigned char* __restrict__ keep) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n_new) return;
  const int64_t a = in_i[k], b = in_j[k];
  bool found = false;
  for (int e = 0; e < n_old; e++) found |= (oi[e] == a) & (oj[e] == b);
  keep[k] = found ? 0 : 1;
}

struct ProxLayout {
  size_t ii, jj, d1, d2, d, cell, skey, scell, cub, total, cub_bytes;
};

static ProxLayout prox_layout(int K) {
  ProxLayout L{};
  size_t o = 0;
  auto take = [&o](size_t bytes) { size_t r = o; o += (bytes + 255) / 256 * 256; return r; };
  const size_t k = size_t(K > 0 ? K : 1);
  L.ii = take(k * 8); L.jj = take(k * 8);
  L.d1 = take(k * 4); L.d2 = take(k * 4); L.d = take(k * 4); L.cell = take(k * 4);
  L.skey = take(k * 4); L.scell = take(k * 4);
  // radix sort temporaries: histograms + one alternate buffer per array in the worst case; sized generously so that the
  // byte count can be computed without a device (cub's own query needs one)
  L.cub_bytes = k * 16 + (size_t(1) << 20);
  L.cub = take(L.cub_bytes);
  L.total = o;
  return L;
}

static bool prox_params(ProxParams& p, int t0, int t1, int t, int rad, int nms, double thresh, int max_factors, int stereo) {
  if (t0 < 0 || t1 < 0 || t < 0 || nms < 0 || nms > 16) return false;
  const long long rows = t > t0 ? t - t0 : 0, cols = t > t1 ? t - t1 : 0;
  if (rows * cols > 0x3fffffff) return false;
  p.t0 = t0; p.t1 = t1; p.t = t; p.rad = rad; p.nms = nms; p.stereo = stereo ? 1 : 0; p.max_factors = max_factors;
  p.thresh = thresh; p.cols = int(cols); p.K = int(rows * cols);
  const int r1 = rad + 1 < 0 ? 0 : rad + 1;
  p.n_init = t > t0 ? int(prox_init_offset(t, t0, r1, p.stereo)) : 0;
  return true;
}

}  // namespace db200

using namespace db200;

extern "C" size_t db200_proximity_workspace_bytes(int t0, int t1, int t) {
  ProxParams p{};
  if (!prox_params(p, t0, t1, t, 0, 0, 0.0, 0, 0)) return 0;
  return prox_layout(p.K).total;
}

extern "C" int db200_proximity_max_edges(int t0, int t, int rad, int stereo, int max_factors) {
  const int r1 = rad + 1 < 0 ? 0 : rad + 1;
  const long long n_init = t > t0 ? prox_init_offset(t, t0, r1, stereo ? 1 : 0) : 0;
  const long long cap = (n_init > max_factors ? n_init : max_factors) + 2;
  return cap > 0x7fffffff ? -1 : int(cap);
}

extern "C" int db200_proximity_edges(const float* poses, const float* disps, const float* intrinsics, int B, int ht, int wd,
                                     int t0, int t1, int t, int rad, int nms, float beta, double thresh, int max_factors,
                                     int stereo, const int64_t* ii_old, const int64_t* jj_old, int n_old, int64_t* es_out,
                                     int es_capacity, int* n_out, void* workspace, size_t workspace_bytes,
                                     db200_stream_t stream) {
  ProxParams p{};
  if (!poses || !disps || !intrinsics || !es_out || !n_out || !workspace || n_old < 0 || B <= 0 || ht <= 0 || wd <= 0 ||
      !prox_params(p, t0, t1, t, rad, nms, thresh, max_factors, stereo) || t > B || (n_old > 0 && (!ii_old || !jj_old)))
    return DB200_ERR_ARG;
  if (es_capacity < db200_proximity_max_edges(t0, t, rad, stereo, max_factors)) return DB200_ERR_ARG;
  const ProxLayout L = prox_layout(p.K);
  if (workspace_bytes < L.total) return DB200_ERR_WORKSPACE;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  unsigned char* ws = static_cast<unsigned char*>(workspace);
  int64_t* ii = reinterpret_cast<int64_t*>(ws + L.ii); int64_t* jj = reinterpret_cast<int64_t*>(ws + L.jj);
  float* d1 = reinterpret_cast<float*>(ws + L.d1); float* d2 = reinterpret_cast<float*>(ws + L.d2);
  float* d = reinterpret_cast<float*>(ws + L.d); int* cell = reinterpret_cast<int*>(ws + L.cell);
  float* skey = reinterpret_cast<float*>(ws + L.skey); int* scell = reinterpret_cast<int*>(ws + L.scell);
  if (p.K > 0) {
    const int nb = cdiv(p.K, 256);
    prox_pairs_kernel<<<nb, 256, 0, st>>>(p, ii, jj);
    DB200_CHECK_LAUNCH();
    int rc = db200_frame_distance(poses, disps, intrinsics, ii, jj, p.K, t, B, ht, wd, beta, d1, stream);
    if (rc != DB200_OK) return rc;
    rc = db200_frame_distance(poses, disps, intrinsics, jj, ii, p.K, t, B, ht, wd, beta, d2, stream);
    if (rc != DB200_OK) return rc;
    prox_mask_kernel<<<nb, 256, 0, st>>>(p, d1, d2, d, cell);
    DB200_CHECK_LAUNCH();
    if (n_old > 0) {
      const long long work = static_cast<long long>(n_old) * (2 * nms + 1) * (2 * nms + 1);
      prox_suppress_old_kernel<<<unsigned((work + 255) / 256), 256, 0, st>>>(p, ii_old, jj_old, n_old, d);
      DB200_CHECK_LAUNCH();
    }
  }
  if (t > t0) {
    prox_init_kernel<<<cdiv(t - t0, 128), 128, 0, st>>>(p, d, es_out);
    DB200_CHECK_LAUNCH();
  }
  if (p.K > 0) {
    size_t need = 0;
    DB200_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, need, d, skey, cell, scell, p.K, 0, 32, st));
    if (need > L.cub_bytes) return DB200_ERR_WORKSPACE;
    size_t have = L.cub_bytes;
    DB200_CUDA(cub::DeviceRadixSort::SortPairs(ws + L.cub, have, d, skey, cell, scell, p.K, 0, 32, st));
    DB200_COUNT(4);
  }
  prox_greedy_kernel<<<1, 32, 0, st>>>(p, skey, scell, d, es_out, n_out);
  DB200_CHECK_LAUNCH();
  return DB200_OK;
}

extern "C" int db200_filter_repeated_edges(const int64_t* ii_new, const int64_t* jj_new, int n_new, const int64_t* ii_old,
                                           const int64_t* jj_old, int n_old, unsigned char* keep, db200_stream_t stream) {
  if (n_new < 0 || n_old < 0 || (n_new > 0 && (!ii_new || !jj_new || !keep)) || (n_old > 0 && (!ii_old || !jj_old)))
    return DB200_ERR_ARG;
  if (n_new == 0) return DB200_OK;
  filter_repeated_kernel<<<cdiv(n_new, 128), 128, 0, static_cast<cudaStream_t>(stream)>>>(ii_new, jj_new, n_new, ii_old, jj_old,
                                                                                         n_old, keep);
  DB200_CHECK_LAUNCH();
  return DB200_OK;
}
