// ba_graph.cu — device-side factor-graph bookkeeping for the dense BA (no host round trips).
// Replaces the host work of ba_cuda/accum_cuda/schur_block (/root/reference/src/droid_kernels.cu:
// 1335-1344 torch::_unique on cat(ts,ii); :948-998 four argsort+CSR builds on the CPU per iteration;
// :1244-1272 the O((N+P)^2) pair enumeration).  Built ONCE per ba call:
//   frame_k[f]  -> row k of the depth system (kx = sorted unique(cat(arange(t0,t1), ii))), or -1
//   deg_off[k]  -> CSR offsets of the edges whose source frame is kx[k], edges kept in original order
//   slot        -> row of a pose edge inside its source frame (stereo ii==jj edges get -1)
// and per iteration the per-edge relative pose records (relSE3, :97-107; stereo baseline :219-229).
#include "ba.cuh"

namespace db200 {

__global__ void graph_flags_kernel(BAParams p) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < p.B) {
    p.frame_k[i] = (i >= p.t0 && i < p.t1) ? 1 : 0;
    p.nd[i] = 0;
    p.fill[i] = 0;
    p.deg_off[i] = 0;
  }
  if (i == 0) {
    p.deg_off[p.B] = 0;
    p.hdr->status = 0; p.hdr->M = 0; p.hdr->chol_fail = 0; p.hdr->any_fail = 0; p.hdr->max_deg = 0; p.hdr->hd_items = 0;
  }
}

__global__ void graph_mark_kernel(BAParams p) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= p.N) return;
  const int64_t i = p.ii[e], j = p.jj[e];
  // the reference indexes poses/disps/Eigen blocks with these unchecked (UB when out of range)
  if (i < 0 || j < 0 || i >= p.B || j >= p.B || i >= p.t1 || j >= p.t1) { atomicMin(&p.hdr->status, DB200_ERR_INDEX); return; }
  p.frame_k[i] = 1;
}

// single CTA: exclusive scan of the frame flags -> frame_k, kx, M
__global__ void graph_scan_kernel(BAParams p) {
  __shared__ int s_part[1024];
  __shared__ int s_base;
  if (threadIdx.x == 0) s_base = 0;
  __syncthreads();
  for (int base = 0; base < p.B; base += 1024) {
    const int f = base + threadIdx.x;
    const int flag = (f < p.B) ? p.frame_k[f] : 0;
    s_part[threadIdx.x] = flag;
    __syncthreads();
    for (int o = 1; o < 1024; o <<= 1) {   // Hillis-Steele inclusive scan
      const int v = (threadIdx.x >= o) ? s_part[threadIdx.x - o] : 0;
      __syncthreads();
      s_part[threadIdx.x] += v;
      __syncthreads();
    }
    const int incl = s_part[threadIdx.x];
    if (f < p.B) {
      const int k = s_base + incl - flag;
      p.frame_k[f] = flag ? k : -1;
      if (flag && k < p.B) p.kx[k] = f;
    }
    __syncthreads();
    if (threadIdx.x == 1023) s_base += incl;
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    p.hdr->M = s_base;
    if (!p.motion_only && s_base != p.eta_rows) atomicMin(&p.hdr->status, DB200_ERR_ETA);
    if (s_base > p.Mcap) atomicMin(&p.hdr->status, DB200_ERR_ETA);
  }
}

__global__ void graph_degree_kernel(BAParams p) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= p.N || p.hdr->status < 0) return;
  const int k = p.frame_k[p.ii[e]];
  atomicAdd(&p.deg_off[k + 1], 1);      // counts; scanned below (integer atomics: deterministic result)
  if (p.ii[e] != p.jj[e]) atomicAdd(&p.nd[k], 1);
}

__global__ void graph_offsets_kernel(BAParams p) {   // single thread block, serial scan over M <= B entries
  if (p.hdr->status < 0) return;
  if (threadIdx.x == 0) {
    int run = 0, mx = 0, items = 0;
    const int M = p.hdr->M;
    for (int k = 0; k < M; k++) {
      const int c = p.deg_off[k + 1];
      p.deg_off[k] = run;
      run += c;
      const int d = p.nd[k];
      mx = max(mx, d);
      // frames beyond the fused path's staging limits (the reference has none: droid_kernels.cu:1244-1272 enumerates
      // every pair of rows of a depth map): items (chunk a >= chunk b) of the high-degree path, ba_frame.cu
      p.hd_off[k] = items;
      if (!p.motion_only && frame_is_hd(d, c)) { const int nc = hd_chunks(d); items += nc * (nc + 1) / 2; }
    }
    p.deg_off[M] = run;
    p.hd_off[M] = items;
    p.hdr->max_deg = mx;
    p.hdr->hd_items = items;
  }
}

// O(N): every edge takes the next free position of its source frame's segment (arrival order, not yet the final order)
__global__ void graph_scatter_kernel(BAParams p) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= p.N || p.hdr->status < 0) return;
  const int k = p.frame_k[p.ii[e]];
  const int pos = atomicAdd(&p.fill[k], 1);
  p.edge_order[p.deg_off[k] + pos] = e;
}

// one CTA per depth row k: its edges in the order of their edge ids (the reference's stable argsort, droid_kernels.cu:948-998
// — and the order the fp32 sums of a frame are formed in, so results do not depend on the arrival order above), and the
// pose-edge slots.  Rank of an edge = number of edges of the segment with a smaller id: O(sum_k ne_k^2) instead of the
// M x N scan of the first version (ne ~ 12; 0.5 GB of index reads per call at 1536 frames x 20 000 edges).
constexpr int kFillStage = 1024;           // segment entries staged in shared memory; longer segments rank from global memory
__global__ void graph_fill_kernel(BAParams p) {
  if (p.hdr->status < 0) return;
  const int k = blockIdx.x;
  if (k >= p.hdr->M) return;
  const int f = p.kx[k];
  const int base = p.deg_off[k], ne = p.deg_off[k + 1] - base;
  const int* seg = p.edge_order + base;
  __shared__ int s_key[kFillStage];          // (edge id << 1) | is-pose-edge: ordered like the edge ids
  auto key = [&](int b) -> int { const int e2 = seg[b]; return (e2 << 1) | int(p.jj[e2] != f); };
  const bool staged = ne <= kFillStage;
  if (staged) {
    for (int a = threadIdx.x; a < ne; a += blockDim.x) s_key[a] = key(a);
    __syncthreads();
  }
  for (int a = threadIdx.x; a < ne; a += blockDim.x) {
    const int v = staged ? s_key[a] : key(a);
    int rank = 0, slot = 0;
    for (int b = 0; b < ne; b++) {
      const int w = staged ? s_key[b] : key(b);
      if (w < v) { rank++; slot += w & 1; }
    }
    const int e = v >> 1;
    EdgeRec& r = p.erec[base + rank];
    r.eid = e;
    r.jpose = int(p.jj[e]) - p.t0;
    r.slot = (v & 1) ? slot : -1;
    r.pad = 0;
  }
}

// per iteration: relative poses of all edges from the CURRENT poses (sorted order)
__global__ void edge_prep_kernel(BAParams p) {
  const int q = blockIdx.x * blockDim.x + threadIdx.x;
  if (q >= p.N || p.hdr->status < 0) return;
  EdgeRec& r = p.erec[q];
  const int e = r.eid;
  const int i = int(p.ii[e]), j = int(p.jj[e]);
  float t[3], qq[4];
  if (i == j) {
    t[0] = -0.1f; t[1] = 0.0f; t[2] = 0.0f; qq[0] = 0.0f; qq[1] = 0.0f; qq[2] = 0.0f; qq[3] = 1.0f;
  } else {
    rel_se3(p.poses + 7 * i, p.poses + 7 * j, t, qq);
  }
  r.t[0] = t[0]; r.t[1] = t[1]; r.t[2] = t[2];
  // R columns = images of the basis vectors under the SAME quaternion action the reference applies
  // (actSO3, :58-68 — no normalisation of q)
  const float ex[3] = {1, 0, 0}, ey[3] = {0, 1, 0}, ez[3] = {0, 0, 1};
  float c0[3], c1[3], c2[3];
  act_so3(qq, ex, c0); act_so3(qq, ey, c1); act_so3(qq, ez, c2);
  r.R[0] = c0[0]; r.R[1] = c1[0]; r.R[2] = c2[0];
  r.R[3] = c0[1]; r.R[4] = c1[1]; r.R[5] = c2[1];
  r.R[6] = c0[2]; r.R[7] = c1[2]; r.R[8] = c2[2];
}

int ba_build_graph(const BAParams& p, cudaStream_t st) {
  const int nb = cdiv(p.B, 256);
  graph_flags_kernel<<<nb, 256, 0, st>>>(p);
  if (p.N > 0) graph_mark_kernel<<<cdiv(p.N, 256), 256, 0, st>>>(p);
  graph_scan_kernel<<<1, 1024, 0, st>>>(p);
  if (p.N > 0) graph_degree_kernel<<<cdiv(p.N, 256), 256, 0, st>>>(p);
  graph_offsets_kernel<<<1, 32, 0, st>>>(p);
  if (p.N > 0) graph_scatter_kernel<<<cdiv(p.N, 256), 256, 0, st>>>(p);
  if (p.N > 0) graph_fill_kernel<<<p.Mcap, 128, 0, st>>>(p);
  DB200_COUNT(p.N > 0 ? 6 : 2);
  DB200_CHECK_LAUNCH();
  return DB200_OK;
}

int ba_edge_prep(const BAParams& p, cudaStream_t st) {
  if (p.N > 0) edge_prep_kernel<<<cdiv(p.N, 256), 256, 0, st>>>(p);
  else DB200_COUNT(-1);
  DB200_CHECK_LAUNCH();
  return DB200_OK;
}

}  // namespace db200
