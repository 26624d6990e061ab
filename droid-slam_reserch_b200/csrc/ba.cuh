// ba.cuh — shared declarations of the dense bundle adjustment (libdroid_b200, sm_100a).
//
// Replaces ba_cuda and its kernels (/root/reference/src/droid_kernels.cu:176-424, 854-1434).
//
// Formulation.  Edges are grouped by SOURCE frame i (= depth map k).  For one source
// frame with pose edges e = 1..d to frames j_e the reference forms, per pixel, E_ii/E_ij rows, C, w and
// per-edge 12x12 Hessians, writes them to HBM (172 KB/edge) and re-reads them in accum/EEt6x6/Ev6x1/EvT6x1.
// Because J_i = -A_e J_j with A_e = [[R^T,0],[-R^T [t]x, R^T]] constant per edge (adjSE3, :79-94), only
//   Hjj_e = sum_p w J_j J_j^T (21), vj_e = sum_p w r J_j (6), E_e[p] = w Jz J_j (6 per pixel, kept in SMEM),
//   G_ef = sum_p Q E_e E_f^T, g_e = sum_p Q w E_e
// are accumulated; every block of the reduced system H - E Q E^T follows from D_ef = [e==f] Hjj_e - G_ef,
// c_e = vj_e - g_e by 6x6 transforms with A_e in fp64 (ba_frame.cu epilogue).  Nothing per-pixel but
// Q and w (8 B/pixel/keyframe) goes back to HBM.
#pragma once
#include "common.cuh"

namespace db200 {

constexpr int kBAThreads = 256;       // threads per CTA of the frame kernels = pixels per sub-tile
constexpr int kMaxDeg = 28;           // max pose edges per source frame on the fused one-CTA-per-frame path
constexpr int kMaxStereo = 4;         // ... and max stereo (ii == jj) edges beside them; frames beyond either limit take the
constexpr int kHdChunk = 14;          // high-degree path: work items (frame, pair of kHdChunk-slot chunks), ba_frame.cu
constexpr int kRec = 16;              // floats per edge record
constexpr int kCholTile = 64;         // tile edge of the reduced-system Cholesky (ba_solve.cu)

// dataflow flags of the solver: one per tile of the lower block triangle (+ the rhs block row), one counter per
// block column, one x-ready flag per block column, a few control words
// per diagonal tile: the 64x64 factor, streamed out panel by panel, + the 64 reciprocal pivots
__host__ __device__ inline size_t chol_stage_doubles() { return size_t(kCholTile) * kCholTile + kCholTile; }
__host__ __device__ inline size_t chol_flag_words(int nt) { return size_t(nt) * (nt + 3) / 2 + 2 * size_t(nt) + 16; }

// per-edge record (recomputed every iteration from the current poses), indexed by SORTED position
struct EdgeRec {
  float t[3];     // t_ij
  float R[9];     // R(q_ij), row-major
  int eid;        // original edge index (addresses target/weight)
  int jpose;      // jj[e] - t0   (<0: fixed)
  int slot;       // row among the pose edges of its source frame, -1 for stereo (ii==jj) edges
  int pad;
};
static_assert(sizeof(EdgeRec) == kRec * 4, "EdgeRec size");

// workspace header (device)
struct BAHeader {
  int status;      // 0 ok, <0 error (DB200_ERR_INDEX/ETA/...), sticky
  int M;           // |kx| found on the device
  int chol_fail;   // >0: a pivot <= 0 was met in the current solve
  int any_fail;    // number of iterations whose solve failed
  int max_deg;     // max pose edges per source frame
  int hd_items;    // work items of the high-degree path (sum over its frames of nc (nc + 1) / 2, nc = ceil(d / kHdChunk))
  int pad[2];
};

struct BALayout {
  size_t header, frame_k, kx, deg_off, nd, hd_off, edge_order, fill, erec, yvec, Q, w, Hb, Wk, cflags, xsol, dx, Lstage, Xstage, total;
  int n, nt, ld, Mcap;
};

// Reduced system Hb = [H | b] in fp64: H is symmetric and only its lower triangle is ever read (ba_solve.cu), so only
// that is stored — packed row by row, element (r, c <= r) at r (r+1)/2 + c.  Half the bytes for the all-reduce of the
// edge-sharded multi-GPU runs (the only exchange of an iteration) and half the atomics of the scatter.
__host__ __device__ inline size_t sys_index(int r, int c) { return size_t(r) * size_t(r + 1) / 2 + size_t(c); }
__host__ __device__ inline size_t sys_h_elems(int n) { return size_t(n) * size_t(n + 1) / 2; }
__host__ __device__ inline size_t sys_elems(int n) { return sys_h_elems(n) + size_t(n); }

// a source frame leaves the one-CTA-per-frame path when its pose edges or its stereo edges exceed what that CTA stages
__host__ __device__ inline bool frame_is_hd(int pose_edges, int all_edges) {
  return pose_edges > kMaxDeg || all_edges - pose_edges > kMaxStereo;
}
__host__ __device__ inline int hd_chunks(int pose_edges) { return (pose_edges + kHdChunk - 1) / kHdChunk; }

inline size_t align_up(size_t x, size_t a = 256) { return (x + a - 1) / a * a; }

// Host-side layout of the caller-provided workspace.
inline BALayout ba_layout(int N, int B, int HW, int t0, int t1, int eta_rows, int motion_only) {
  BALayout L{};
  const int P = t1 - t0;
  L.n = 6 * P;
  L.nt = (L.n + kCholTile - 1) / kCholTile;     // 64x64 tiles of the solver (ba_solve.cu)
  L.ld = L.nt * kCholTile;
  L.Mcap = motion_only ? (B < P + N ? B : P + N) : eta_rows;
  if (L.Mcap < 1) L.Mcap = 1;
  size_t o = 0;
  auto take = [&](size_t bytes) { size_t r = o; o = align_up(o + bytes); return r; };
  L.header = take(sizeof(BAHeader));
  L.frame_k = take(sizeof(int) * size_t(B));
  L.kx = take(sizeof(int) * size_t(B));
  L.deg_off = take(sizeof(int) * size_t(B + 1));
  L.nd = take(sizeof(int) * size_t(B));
  L.hd_off = take(sizeof(int) * size_t(B + 1));
  L.edge_order = take(sizeof(int) * size_t(N > 0 ? N : 1));
  L.fill = take(sizeof(int) * size_t(B));
  L.erec = take(sizeof(EdgeRec) * size_t(N > 0 ? N : 1));
  L.yvec = take(sizeof(float) * 6 * size_t(N > 0 ? N : 1));
  L.Q = take(motion_only ? 16 : sizeof(float) * size_t(L.Mcap) * HW);
  L.w = take(motion_only ? 16 : sizeof(float) * size_t(L.Mcap) * HW);
  L.Hb = take(sizeof(double) * sys_elems(L.n));
  // [A; b^T] -> [L; y^T] in place: lower block triangle + the rhs block row, tile-major, tiles of [64][64+4] (ba_solve.cu)
  L.Wk = take(sizeof(double) * (size_t(L.nt) * (L.nt + 3) / 2) * kCholTile * (kCholTile + 4));
  L.cflags = take(sizeof(unsigned) * chol_flag_words(L.nt));
  L.xsol = take(sizeof(double) * size_t(L.ld));
  L.dx = take(sizeof(float) * size_t(L.n));
  L.Lstage = take(sizeof(double) * size_t(L.nt) * chol_stage_doubles());   // streamed diagonal factors (ba_solve.cu)
  L.Xstage = take(sizeof(double) * size_t(L.nt) * kCholTile * kCholTile);   // streamed sub-diagonal tiles of the chain
  L.total = o;
  return L;
}

struct BAParams {
  // caller tensors
  float* poses; float* disps; const float* intr; const float* sens; const float* target; const float* weight;
  const float* eta; const int64_t* ii; const int64_t* jj;
  int N, B, ht, wd, HW, t0, t1, P, eta_rows, motion_only;
  int own_lo, own_hi;   // owned source frames [own_lo, own_hi) (edge-sharded multi-GPU runs; 0, B otherwise)
  // workspace views
  BAHeader* hdr; int* frame_k; int* kx; int* deg_off; int* nd; int* hd_off; int* edge_order; int* fill; EdgeRec* erec; float* yvec;
  float* Q; float* w; double* Hb; double* Wk; unsigned* cflags; double* xsol; float* dx; double* Lstage; double* Xstage;
  int n, nt, ld, Mcap;
  int chol_nd;          // solver: number of diagonal CTAs, -1 = automatic (ba_solve.cu)
};

inline BAParams ba_params(const BALayout& L, void* ws) {
  BAParams p{};
  char* b = static_cast<char*>(ws);
  p.hdr = reinterpret_cast<BAHeader*>(b + L.header);
  p.frame_k = reinterpret_cast<int*>(b + L.frame_k);
  p.kx = reinterpret_cast<int*>(b + L.kx);
  p.deg_off = reinterpret_cast<int*>(b + L.deg_off);
  p.nd = reinterpret_cast<int*>(b + L.nd);
  p.hd_off = reinterpret_cast<int*>(b + L.hd_off);
  p.edge_order = reinterpret_cast<int*>(b + L.edge_order);
  p.fill = reinterpret_cast<int*>(b + L.fill);
  p.erec = reinterpret_cast<EdgeRec*>(b + L.erec);
  p.yvec = reinterpret_cast<float*>(b + L.yvec);
  p.Q = reinterpret_cast<float*>(b + L.Q);
  p.w = reinterpret_cast<float*>(b + L.w);
  p.Hb = reinterpret_cast<double*>(b + L.Hb);
  p.Wk = reinterpret_cast<double*>(b + L.Wk);
  p.cflags = reinterpret_cast<unsigned*>(b + L.cflags);
  p.xsol = reinterpret_cast<double*>(b + L.xsol);
  p.dx = reinterpret_cast<float*>(b + L.dx);
  p.Lstage = reinterpret_cast<double*>(b + L.Lstage);
  p.Xstage = reinterpret_cast<double*>(b + L.Xstage);
  p.n = L.n; p.nt = L.nt; p.ld = L.ld; p.Mcap = L.Mcap; p.chol_nd = -1;
  return p;
}

// stage entry points (each enqueues kernels on `st`; returns DB200_OK / DB200_ERR_CUDA)
int ba_build_graph(const BAParams& p, cudaStream_t st);                 // ba_graph.cu
int ba_edge_prep(const BAParams& p, cudaStream_t st);                   // ba_graph.cu
int ba_accumulate_frames(const BAParams& p, cudaStream_t st);           // ba_frame.cu
int ba_backsub_frames(const BAParams& p, float* dz_out, cudaStream_t st);  // ba_frame.cu
int ba_solve(const BAParams& p, float lm, float ep, cudaStream_t st);   // ba_solve.cu
int ba_post_solve(const BAParams& p, float* dx_out, cudaStream_t st);   // ba_solve.cu (dx, y_e, pose retraction)

}  // namespace db200
