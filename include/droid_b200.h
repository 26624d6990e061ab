/* droid_b200.h — C-ABI of the B200-native DROID-SLAM hot path (libdroid_b200.so).
 *
 * Every entry point replaces one operator of the reference's pybind module `droid_backends`
 * (/root/reference/src/droid.cpp:237-250) or one PyTorch step of droid_slam/modules/corr.py.
 * Conventions: plain device pointers + sizes, an explicit cudaStream_t (passed as void*), no
 * allocation, no exceptions, no torch types.  All tensors are contiguous (droid.cpp:84-85).
 * Return value: 0 = ok; <0 = invalid argument / CUDA error (nothing useful was computed);
 * Launches are asynchronous on `stream`.
 */
#ifndef DROID_B200_H
#define DROID_B200_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DB200_OK 0
#define DB200_ERR_ARG (-1)      /* null pointer / negative size / unsupported dtype or radius */
#define DB200_ERR_CUDA (-2)     /* a CUDA runtime call or launch failed */
#define DB200_ERR_WORKSPACE (-3)/* workspace too small */
#define DB200_ERR_INDEX (-4)    /* (device status) edge index out of range */
#define DB200_ERR_ETA (-5)      /* (device status) eta rows != |unique(cat(arange(t0,t1), ii))| */
#define DB200_ERR_STATIC_LIMIT (-6) /* problem needs more scratch than the built-in static array of a *_static entry point holds */

#define DB200_F32 0
#define DB200_F16 1
#define DB200_F64 2

typedef void* db200_stream_t; /* cudaStream_t */

int db200_version(void);
const char* db200_status_string(int status);
const char* db200_last_cuda_error(void); /* text of the last CUDA error the library saw (diagnostics) */

/* ---- instrumentation (no reference counterpart; used by bench.py) ----------------------------
 * db200_launch_count: kernels launched by this library since load.
 * db200_profile_enable(1): record CUDA events around the BA stages on the caller's stream;
 * db200_profile_collect: synchronise those events, return summed milliseconds and number of
 * recorded intervals per stage (arrays of DB200_NUM_STAGES), and reset. */
#define DB200_STAGE_GRAPH 0   /* device-side graph build (once per ba call) */
#define DB200_STAGE_PREP 1    /* per-edge relative poses */
#define DB200_STAGE_ACCUM 2   /* fused reprojection/Jacobian/Hessian/Schur accumulation (the HBM-bound kernel) */
#define DB200_STAGE_SOLVE 3   /* damped Cholesky of the reduced pose system */
#define DB200_STAGE_POST 4    /* dx, pose retraction, per-edge back-substitution vectors */
#define DB200_STAGE_BACKSUB 5 /* depth back-substitution + disparity update */
#define DB200_NUM_STAGES 6
long long db200_launch_count(void);
void db200_profile_enable(int on);
int db200_profile_collect(double* ms_out, int* count_out);
/* FP64 rate (GFLOP/s) of mma.sync.m8n8k4.f64 (DMMA), register-only, measured now on the current device: the denominator
 * bench.py reports the solver against.  Synchronises; allocates a small sink (debug/bench use only). */
double db200_debug_fp64_peak_gflops(db200_stream_t stream);
/* debug timeline of the Cholesky panel chain: device buffer of 8*(2*nt) int64 (globaltimer ns), NULL = off */
int db200_debug_chol_trace(long long* device_buffer);

/* ---- correlation lookup ------------------------------------------------------------------
 * corr_index_forward  (droid.cpp:170-178, correlation_kernels.cu:20-70,126-155)
 *   volume[N,h1,w1,h2,w2] (dtype), coords[N,2,h1,w1] f32 (x then y) -> corr[N,2r+1,2r+1,h1,w1] (dtype),
 *   first output axis = x offset.  corr is fully overwritten. */
int db200_corr_index_forward(const void* volume, const float* coords, void* corr, int N, int h1, int w1,
                             int h2, int w2, int radius, int dtype, db200_stream_t stream);
/* corr_index_backward (droid.cpp:180-191, correlation_kernels.cu:73-124,157-185)
 *   corr_grad[N,2r+1,2r+1,h1,w1] -> volume_grad[N,h1,w1,h2,w2], fully overwritten (zeros outside windows). */
int db200_corr_index_backward(const float* coords, const void* corr_grad, void* volume_grad, int N, int h1,
                              int w1, int h2, int w2, int radius, int dtype, db200_stream_t stream);
/* CorrBlock.__call__ fused (corr.py:40-50): all `num_levels` lookups + the channel concat in one launch.
 *   levels[l] = device pointer of volume level l [N,h1,w1,h2>>l,w2>>l]; coords are LEVEL-0 coords
 *   (the kernel divides by 2^l);  out[N, num_levels*(2r+1)^2, h1, w1]. */
int db200_corr_pyramid_lookup(const void* const* levels, const float* coords, void* out, int N, int h1, int w1,
                              int h2, int w2, int num_levels, int radius, int dtype, db200_stream_t stream);

/* ---- all-pairs correlation volume + pyramid (corr.py:30-38,63-71) ---------------------------
 *   fmap1,fmap2[N,C,ht,wd] (dtype f16 or f32) -> levels_out[l][N,ht,wd,ht>>l,wd>>l] (same dtype),
 *   corr = (f1/4)^T (f2/4), level l+1 = avg_pool2d(level l, 2) — computed as f1^T pool(f2) (linear);
 *   `workspace` (db200_corr_volume_workspace_bytes) holds the pooled copies of fmap2. */
size_t db200_corr_volume_workspace_bytes(int N, int C, int ht, int wd, int num_levels, int dtype);
int db200_corr_volume_pyramid(const void* fmap1, const void* fmap2, void* const* levels_out, int N, int C,
                              int ht, int wd, int num_levels, int dtype, void* workspace, size_t workspace_bytes,
                              db200_stream_t stream);

/* ---- tiled levels: CorrBlock's private pyramid layout (no counterpart in the reference; corr.py:23-50 keeps the volumes
 * behind the class, so their layout is the implementation's choice as long as corr_pyramid[i] can be materialised) ----
 * The [h2][w2] slice of every source pixel is stored as 4x8-element tiles (64 bytes of fp16 = one DRAM burst; tiles row-major,
 * the 4 rows of a tile 16 bytes apart): element (y, x) sits at ((y>>2)*(w2>>3) + (x>>3))*32 + (y&3)*8 + (x&7).  The 8x8-sample
 * window of a radius-3 lookup then touches ~5 DRAM bursts instead of 8-10 (row-major: one burst per 16-byte window row).
 * db200_corr_tiled_levels: bit l set = level l of a (C, ht, wd, num_levels, dtype) pyramid may be tiled (fp16, C = 128 on the
 *   tcgen05 path; level height % 4 == 0, width % 8 == 0 and >= 32); 0 = not supported for this shape.
 * db200_corr_volume_pyramid_tiled: db200_corr_volume_pyramid with the levels of `tiled_mask` (a subset of the above) written
 *   in tiled order.  Level 0 holds the same bits as the row-major call (the GEMM's B operand rows are permuted, nothing else
 *   changes).  When wd == 64, ht % 8 == 0 and the mask is the full db200_corr_tiled_levels mask, every level comes from ONE
 *   launch: fmap1 is read in place, and levels 1..3 are the 2x2 / 4x4 / 8x8 means of the level-0 fp32 accumulators rounded to
 *   fp16 once (within half an fp16 spacing of the exact mean; the row-major call pools fp16 operands first, the reference
 *   pools the rounded level below: all within fp16 rounding, not the same bits — DESIGN.md §5).
 * db200_corr_pyramid_lookup_tiled: db200_corr_pyramid_lookup reading such a pyramid (fp16, radius 3); bit-identical output. */
unsigned db200_corr_tiled_levels(int C, int ht, int wd, int num_levels, int dtype);
int db200_corr_volume_pyramid_tiled(const void* fmap1, const void* fmap2, void* const* levels_out, int N, int C,
                                    int ht, int wd, int num_levels, int dtype, void* workspace, size_t workspace_bytes,
                                    unsigned tiled_mask, db200_stream_t stream);
int db200_corr_pyramid_lookup_tiled(const void* const* levels, const float* coords, void* out, int N, int h1, int w1,
                                    int h2, int w2, int num_levels, int radius, int dtype, unsigned tiled_mask,
                                    db200_stream_t stream);

/* ---- low-memory correlation (altcorr) ------------------------------------------------------
 * altcorr_forward (droid.cpp:193-203, altcorr_kernel.cu:27-149,290-319), fp32:
 *   fmap1[B,H1,W1,C], fmap2[B,H2,W2,C], coords[B,S,H1,W1,2] -> corr[B,S,(2r+1)^2,H1,W1], channel = iy + (2r+1)*ix */
int db200_altcorr_forward(const float* fmap1, const float* fmap2, const float* coords, float* corr, int B, int S,
                          int H1, int W1, int H2, int W2, int C, int radius, db200_stream_t stream);
/* altcorr_backward (droid.cpp:205-217, altcorr_kernel.cu:152-286,321-355): fmap1_grad, fmap2_grad fully
 * overwritten; coords_grad[B,S,H1,W1,2] is zero-filled like the reference (:340). */
int db200_altcorr_backward(const float* fmap1, const float* fmap2, const float* coords, const float* corr_grad,
                           float* fmap1_grad, float* fmap2_grad, float* coords_grad, int B, int S, int H1, int W1,
                           int H2, int W2, int C, int radius, db200_stream_t stream);

/* altcorr_forward on the tensor cores (csrc/altcorr_tc.cu; C = 128, radius = 3; other shapes and a NULL workspace take
 * db200_altcorr_forward): same arguments + a workspace of db200_altcorr_workspace_bytes(...) that holds the fp16 hi/lo
 * split of the operands and the status words (int[0] error, int[1] tiles handed to the SIMT kernel) at offset 256,
 * readable with db200_altcorr_status (synchronises). */
size_t db200_altcorr_workspace_bytes(int B, int S, int H1, int W1, int H2, int W2, int C);
int db200_altcorr_forward_ws(const float* fmap1, const float* fmap2, const float* coords, float* corr, int B, int S,
                             int H1, int W1, int H2, int W2, int C, int radius, void* workspace,
                             size_t workspace_bytes, db200_stream_t stream);
int db200_altcorr_status(const void* status_words, db200_stream_t stream);
/* instrumentation: number of (pixel tile x target cell) GEMMs the last forward launch issued (synchronises).  It depends on
 * the coordinates: ~10 per tile for smooth flow over 4 levels, up to 64 per level for scattered coordinates. */
long long db200_altcorr_cells(const void* status_words, db200_stream_t stream);
/* debug (env DB200_ALT_DBG & 16): cycles summed over the CTAs that the producer / MMA / one epilogue thread of the forward
 * kernel spent in each of their waits: out12 = {producer: plan slot, A stage, B stage, total | MMA: plan, A landed, accumulator
 * free, B landed, total | epilogue: plan, accumulator ready, total}. */
int db200_debug_altcorr_waits(const void* status_words, long long* out12, db200_stream_t stream);
/* altcorr_backward on the tensor cores (same workspace size as the forward; C = 128, radius = 3, else the SIMT kernel):
 * d fmap1 = G . fmap2[cell], d fmap2[cell] = G^T . fmap1[tile] per pixel tile and target cell, G = the 4-corner combination
 * of corr_grad; fmap2_grad is accumulated with one vector red per 4 channels and cell position. */
int db200_altcorr_backward_ws(const float* fmap1, const float* fmap2, const float* coords, const float* corr_grad,
                              float* fmap1_grad, float* fmap2_grad, float* coords_grad, int B, int S, int H1, int W1,
                              int H2, int W2, int C, int radius, void* workspace, size_t workspace_bytes,
                              db200_stream_t stream);
/* AltCorrBlock.corr_fn fused (corr.py:109-131 as driven by factor_graph.py:261-277): all `num_levels` levels, the
 * frame gathers `pyramid[0][:, ii]`, `pyramid[l][:, jj]`, the casts and the channel concat in ONE launch.
 *   fmap1_hi, fmap2_hi[l] : fp16 feature pyramid in the CHANNEL-BLOCKED operand layout [frames, 2, H>>l, W>>l, 64] (channel c of a
 *                           position in block c / 64: a TMA box of 64 channels x 16 x 8 positions is then 8 runs of 2 KB instead of
 *                           128 segments of 128 bytes; db200_split_hilo_f16 with hw = (H>>l)*(W>>l) writes it from channels-last
 *                           fp32) (fmap1 = level 0 of the same or another pyramid); for fp32 features pass the fp16 split x = hi + lo (db200_split_hilo_f16)
 *                           with its flag word and the fp32 arrays themselves (used when a value leaves the fp16 range);
 *                           lo / flags / f32 pointers are NULL for fp16 features.
 *   idx1, idx2 [B] int64  : frame of the source / target image of every edge;
 *   coords                : level-0 coordinates (x, y) at coords[b*c_sb + s*c_ss + pixel*c_sp (+0,1)] (strides in floats);
 *   out                   : out[b*o_sb + s*o_ss + (49*l + iy + 7*ix)*o_sc + pixel*o_sp], every element written.
 * workspace: db200_altcorr_pyramid_workspace_bytes(B, S, H, W) bytes (status words first). */
size_t db200_altcorr_pyramid_workspace_bytes(int B, int S, int H1, int W1);
int db200_altcorr_pyramid_forward(const void* fmap1_hi, const void* fmap1_lo, const void* const* fmap2_hi,
                                  const void* const* fmap2_lo, const unsigned* flags, const float* fmap1_f32,
                                  const float* const* fmap2_f32, long long frames, int H, int W, int C, int num_levels,
                                  const int64_t* idx1, const int64_t* idx2, const float* coords, long long c_sb,
                                  long long c_ss, long long c_sp, float* out, long long o_sb, long long o_ss,
                                  long long o_sc, long long o_sp, int B, int S, int radius, void* workspace,
                                  size_t workspace_bytes, db200_stream_t stream);
/* altcorr_forward WITHOUT any allocation and without a caller workspace.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops,
 * harnesses that forbid it).  Same arguments, kernels and results as db200_altcorr_forward_ws (C = 128, radius = 3 only;
 * other shapes -> DB200_ERR_ARG, use db200_altcorr_forward, which needs no scratch).  The scratch is a static device array
 * of db200_altcorr_static_max_bytes() = 8 MiB per GPU that loads the library; a problem whose
 * db200_altcorr_workspace_bytes(...) exceeds it returns DB200_ERR_STATIC_LIMIT and computes nothing.  The call makes no
 * cudaMalloc/cudaFree and calls no allocating library routine.  Every call shares the array: calls that overlap on the
 * GPU (other streams, parallel graph branches) race; calls ordered on one stream are safe (the words that must start at
 * zero are cleared on `stream` at the start of each call). */
size_t db200_altcorr_static_max_bytes(void);
int db200_altcorr_forward_static(const float* fmap1, const float* fmap2, const float* coords, float* corr, int B, int S,
                                 int H1, int W1, int H2, int W2, int C, int radius, db200_stream_t stream);
/* x[n] fp32 (channels-last [F, hw, 128] when hw != 0) -> hi[n], lo[n] fp16 with x ~= hi + lo, written channel-blocked
 * [F, 2, hw, 64] when hw != 0 (hw = 0: same element order as x); *flags |= 1 if any lo != 0, |= 2 if any value is non-finite or
 * outside the fp16 range (flags must be zeroed by the caller). n % 4 == 0. */
int db200_split_hilo_f16(const float* x, void* hi, void* lo, size_t n, size_t hw, unsigned* flags, db200_stream_t stream);

/* ---- frame_distance (droid.cpp:120-136, droid_kernels.cu:518-657,1438-1460) ----------------- */
int db200_frame_distance(const float* poses, const float* disps, const float* intrinsics, const int64_t* ii,
                         const int64_t* jj, int K, int n_poses, int B, int ht, int wd, float beta, float* dist,
                         db200_stream_t stream);

/* ---- factor-graph edge selection (droid_slam/factor_graph.py:315-379 `add_proximity_factors`, :44-55
 * `__filter_repeated_edges`) ---------------------------------------------------------------
 * db200_proximity_edges writes the reference's `es` list, in its order, to es_out[n][2] (int64, device) and the
 * count to *n_out (device int): first the neighbourhood edges ((i,i) when `stereo`, then (i,j),(j,i) for
 * j in [max(i-rad-1,0), i), i = t0..t-1), then the greedy distance-ordered picks with (2*nms+1)^2 suppression, both
 * directions per pick, stopping once more than `max_factors` entries exist.  `t` = video.counter.value;
 * (ii_old, jj_old) = cat(ii, ii_bad, ii_inac) / cat(jj, ...) of the graph.  Distances are the mean of both
 * directions of frame_distance with `beta` over meshgrid(arange(t0,t), arange(t1,t)) (depth_video.py:171-181).
 * Ties in distance (unspecified order in torch.argsort) resolve to the lower flat pair index; NaN distances
 * (non-finite poses) are never picked.  es_capacity >= db200_proximity_max_edges(...). */
size_t db200_proximity_workspace_bytes(int t0, int t1, int t);
int db200_proximity_max_edges(int t0, int t, int rad, int stereo, int max_factors);
int db200_proximity_edges(const float* poses, const float* disps, const float* intrinsics, int B, int ht, int wd,
                          int t0, int t1, int t, int rad, int nms, float beta, double thresh, int max_factors,
                          int stereo, const int64_t* ii_old, const int64_t* jj_old, int n_old, int64_t* es_out,
                          int es_capacity, int* n_out, void* workspace, size_t workspace_bytes,
                          db200_stream_t stream);
/* keep[k] = 1 iff (ii_new[k], jj_new[k]) is not among the old edges (factor_graph.py:44-55; duplicates inside the
 * new list are kept, as there). */
int db200_filter_repeated_edges(const int64_t* ii_new, const int64_t* jj_new, int n_new, const int64_t* ii_old,
                                const int64_t* jj_old, int n_old, unsigned char* keep, db200_stream_t stream);

/* ---- projmap (droid.cpp:139-154, droid_kernels.cu:427-516,1463-1490): coords[num,ht,wd,3] (channel 2 = 0),
 * valid[num,ht,wd,1] ------------------------------------------------------------------------ */
int db200_projmap(const float* poses, const float* disps, const float* intrinsics, const int64_t* ii,
                  const int64_t* jj, int num, int n_poses, int B, int ht, int wd, float* coords, float* valid,
                  db200_stream_t stream);

/* ---- reprojection + motion features: DepthVideo.reproject (depth_video.py:150-158 ->
 * geom/projective_ops.py:96-125 without Jacobians) fused with FactorGraph.update's `motn` (factor_graph.py:203-205).
 * intrinsics: [4] shared, or [B,4] per frame when per_frame_intrinsics != 0 (ii's for the back-projection, jj's
 * for the projection).  Outputs (each may be NULL): coords[N,ht,wd,2], valid[N,ht,wd,1],
 * motn[N,4,ht,wd] = clamp(cat(coords - grid, target - coords), -64, 64) with target[N,ht,wd,2]. */
int db200_reproject(const float* poses, const float* disps, const float* intrinsics, int per_frame_intrinsics,
                    const int64_t* ii, const int64_t* jj, int N, int n_poses, int B, int ht, int wd, float* coords,
                    float* valid, const float* target, float* motn, db200_stream_t stream);

/* ---- iproj / depth_filter (droid.cpp:157-167,220-234, droid_kernels.cu:661-850) ------------- */
int db200_iproj(const float* poses, const float* disps, const float* intrinsics, int num, int ht, int wd,
                float* points, db200_stream_t stream);
int db200_depth_filter(const float* poses, const float* disps, const float* intrinsics, const int64_t* inds,
                       const float* thresh, int n_inds, int num, int ht, int wd, float* counter,
                       db200_stream_t stream);

/* ---- dense bundle adjustment (droid.cpp:88-117, droid_kernels.cu:1314-1434) -------------------
 * poses[B,7] and disps[B,ht,wd] are updated IN PLACE.  dx_out[P,6], dz_out[M,ht*wd] receive the last
 * iteration's updates (dz_out may be NULL when motion_only).  `eta_rows` = eta.shape[0] = M.
 * `workspace` must hold db200_ba_workspace_bytes(...) bytes; its first int is a device status word
 * (0 ok, DB200_ERR_INDEX, DB200_ERR_ETA, 1 = Cholesky broke down in some iteration -> zero pose update,
 * like droid_kernels.cu:1207-1210) which db200_ba_status() copies back (synchronises the stream). */
size_t db200_ba_workspace_bytes(int N, int B, int ht, int wd, int t0, int t1, int eta_rows, int motion_only);
int db200_ba(float* poses, float* disps, const float* intrinsics, const float* disps_sens, const float* targets,
             const float* weights, const float* eta, const int64_t* ii, const int64_t* jj, int N, int B, int ht,
             int wd, int t0, int t1, int iterations, float lm, float ep, int motion_only, int eta_rows,
             float* dx_out, float* dz_out, void* workspace, size_t workspace_bytes, db200_stream_t stream);
int db200_ba_status(const void* workspace, db200_stream_t stream);

/* Phase-split BA for edge-sharded multi-GPU runs.  One iteration =
 *   db200_ba_accumulate  : local edges -> this rank's partial reduced system Hb = [H (lower triangle packed row by
 *                          row: element (r, c <= r) at r(r+1)/2 + c), b (n)] in fp64, n = 6*(t1-t0);
 *                          db200_ba_system_elems() = n(n+1)/2 + n doubles;   (first call builds the graph)
 *   <allreduce(sum) of Hb over ranks — NCCL, done by the caller>
 *   db200_ba_solve_update: damped Cholesky solve, pose retraction (all poses), depth back-substitution
 *                          and disparity update for the LOCAL source frames.
 * `Hb` points into the workspace: db200_ba_system_ptr().
 * [own_lo, own_hi) = the source frames (depth maps) this rank owns: depth work (C, w, Q, dz, the disparity
 * update) is done for those frames only; every local edge must have its source frame in that range.
 * Single-GPU callers pass 0, B. */
int db200_ba_accumulate(float* poses, float* disps, const float* intrinsics, const float* disps_sens,
                        const float* targets, const float* weights, const float* eta, const int64_t* ii,
                        const int64_t* jj, int N, int B, int ht, int wd, int t0, int t1, int motion_only,
                        int eta_rows, int first_iteration, int own_lo, int own_hi, void* workspace,
                        size_t workspace_bytes, db200_stream_t stream);
int db200_ba_solve_update(float* poses, float* disps, const float* intrinsics, const float* targets,
                          const float* weights, const int64_t* ii, const int64_t* jj, int N, int B, int ht, int wd,
                          int t0, int t1, float lm, float ep, int motion_only, int eta_rows, int own_lo, int own_hi,
                          float* dx_out, float* dz_out, void* workspace, size_t workspace_bytes,
                          db200_stream_t stream);
double* db200_ba_system_ptr(void* workspace, int N, int B, int ht, int wd, int t0, int t1, int eta_rows,
                            int motion_only);
size_t db200_ba_system_elems(int t0, int t1);

#ifdef __cplusplus
}
#endif
#endif /* DROID_B200_H */
