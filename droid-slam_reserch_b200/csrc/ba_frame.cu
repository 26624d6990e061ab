// ba_frame.cu — fused per-source-frame accumulation of the dense BA and the depth back-substitution.
// Replaces projective_transform_kernel, accum_kernel, EEt6x6_kernel, Ev6x1_kernel, EvT6x1_kernel,
// disp_retr_kernel and the host-side block assembly of /root/reference/src/droid_kernels.cu
// (:176-424, :854-874, :1001-1115, :933-946, :1117-1190, :1222-1311).
//
// ba_schur_kernel (full BA).  One CTA = (source frame k, pixel split s), 256 threads, 64-pixel chunks:
//   phase 1 (thread = (edge, pixel lane)): every thread owns ONE edge of the frame and walks the chunk's pixels
//           with a stride: reprojection, residual, J_j, J_z; the 27 per-edge sums (Hjj lower triangle + vj)
//           accumulate in REGISTERS over all chunks (one cross-lane reduction per CTA, not per pixel);
//           E_e[6] and the edge's C/w contributions go to shared memory.
//   phase 1b (thread = pixel): C, w over the frame's edges in edge order -> Q = 1/C, w to HBM; the stash is
//           scaled by sqrt(Q) and split into TF32 hi/lo parts; one extra row holds sqrt(Q) w.
//   phase 2 (warp = list of 16x8 output tiles): G = Es Es^T (lower) and g = Es (sqrt(Q) w) on the tensor
//           cores, mma.sync m16n8k8 TF32 with the 3-product split (hi*hi + hi*lo + lo*hi: fp32-faithful),
//           accumulators persistent in registers over the CTA's chunks.
//   epilogue: fp64 6x6 transforms with A_e and scatter of this CTA's contribution into the reduced
//           system with fp64 red.global.add (the only atomics; they carry ~d^2*36 values per CTA).
// ba_motion_kernel (motion-only BA): thread = pixel, shuffle-butterfly reduction of the 27 sums.
// HBM traffic per edge-pixel: target 8 B + weight 8 B (+ disp 4 B per keyframe-pixel) — the algorithmic
// 20*HW*N + 16*HW*M of SURVEY.md §8d; the back-substitution re-reads weight (8 B) to recompute E.
#include <cstdlib>
#include <mutex>
#include "ba.cuh"

namespace db200 {

// ------------------------------------------------------------------------------------------------
// per pixel-edge geometry shared by the accumulate and back-substitution kernels
struct PixEdge {
  float Ju[6], Jv[6];   // rows of J_j (u and v residual); Ju[1] == 0, Jv[0] == 0
  float Jzu, Jzv;       // d r / d disparity
  float ru, rv, wu, wv; // residuals and weights (0.001 * weight, 0 when behind the MIN_DEPTH plane)
};

__device__ __forceinline__ void pix_edge(const EdgeRec& r, float X0, float X1, float h, float fx, float fy, float cx,
                                         float cy, float tu, float tv, float wur, float wvr, PixEdge& o) {
  // Y = R (X0, X1, 1) + h t     (actSE3, droid_kernels.cu:70-77)
  const float x = r.R[0] * X0 + r.R[1] * X1 + r.R[2] + h * r.t[0];
  const float y = r.R[3] * X0 + r.R[4] * X1 + r.R[5] + h * r.t[1];
  const float z = r.R[6] * X0 + r.R[7] * X1 + r.R[8] + h * r.t[2];
  const bool bad = z < 0.25f;                       // MIN_DEPTH (:26, :302)
  const float d = bad ? 0.0f : 1.0f / z;
  const float d2 = d * d;
  o.wu = bad ? 0.0f : 0.001f * wur;   // `.001 * weight` (:305) — the source's double product differs by <= 1 ulp
  o.wv = bad ? 0.0f : 0.001f * wvr;
  o.ru = tu - (fx * d * x + cx);
  o.rv = tv - (fy * d * y + cy);
  o.Ju[0] = fx * (h * d);
  o.Ju[1] = 0.0f;
  o.Ju[2] = fx * (-x * h * d2);
  o.Ju[3] = fx * (-x * y * d2);
  o.Ju[4] = fx * (1.0f + x * x * d2);
  o.Ju[5] = fx * (-y * d);
  o.Jv[0] = 0.0f;
  o.Jv[1] = fy * (h * d);
  o.Jv[2] = fy * (-y * h * d2);
  o.Jv[3] = fy * (-1.0f - y * y * d2);
  o.Jv[4] = fy * (x * y * d2);
  o.Jv[5] = fy * (x * d);
  o.Jzu = fx * (r.t[0] * d - r.t[2] * (x * d2));
  o.Jzv = fy * (r.t[1] * d - r.t[2] * (y * d2));
}

__device__ __forceinline__ EdgeRec load_rec(const EdgeRec* g) {
  EdgeRec r;
  const float4* s = reinterpret_cast<const float4*>(g);
  float4* d = reinterpret_cast<float4*>(&r);
  d[0] = __ldg(s); d[1] = __ldg(s + 1); d[2] = __ldg(s + 2); d[3] = __ldg(s + 3);
  return r;
}

// Transposing butterfly: every lane passes 32 values; on return lane L holds sum over lanes of v[L].
__device__ __forceinline__ float warp_transpose_sum32(float (&v)[32], int lane) {
#pragma unroll
  for (int off = 16, n = 16; off >= 1; off >>= 1, n >>= 1) {
    const bool hi = (lane & off) != 0;
#pragma unroll
    for (int i = 0; i < n; i++) {
      const float send = hi ? v[i] : v[i + n];
      const float keep = hi ? v[i + n] : v[i];
      v[i] = keep + __shfl_xor_sync(0xffffffffu, send, off);
    }
  }
  return v[0];
}

// A_e = [[R^T, 0], [-R^T [t]x, R^T]]  (J_i = -A_e J_j; adjSE3, droid_kernels.cu:79-94)
__device__ __forceinline__ void build_A(const EdgeRec& r, float* A) {
  const float tx[9] = {0.0f, -r.t[2], r.t[1], r.t[2], 0.0f, -r.t[0], -r.t[1], r.t[0], 0.0f};
  for (int a = 0; a < 3; a++)
    for (int b = 0; b < 3; b++) {
      const float rt = r.R[b * 3 + a];   // R^T[a][b]
      A[a * 6 + b] = rt;
      A[a * 6 + 3 + b] = 0.0f;
      A[(3 + a) * 6 + 3 + b] = rt;
      float s = 0.0f;
      for (int m = 0; m < 3; m++) s += r.R[m * 3 + a] * tx[m * 3 + b];
      A[(3 + a) * 6 + b] = -s;
    }
}

__device__ __forceinline__ int hs_index(int a, int b) {  // symmetric access into the 21-entry lower triangle
  const int n = a > b ? a : b, m = a > b ? b : a;
  return n * (n + 1) / 2 + m;
}

// ------------------------------------------------------------------------------------------------
// motion-only BA (droid_kernels.cu:1336-1346 with motion_only = true: the depth terms are dropped, the pose system is
// H = sum_e [A_e; I]^T Hjj_e [A_e; I]).  Shared-memory carve-up (sizes derive from `dcap` = pose edges handled per pass)
struct FrameSmem {
  double* Ws;     // [dcap][36]
  double* cs;     // [dcap][6]
  float* wacc;    // [8][dcap][27]
  float* Am;      // [dcap][36]
  float* Hs;      // [dcap][27]
  int* jp;        // [dcap]
  int* slotq;     // [dcap]
};

__host__ __device__ inline size_t frame_smem_bytes(int dcap) {
  size_t b = sizeof(double) * size_t(dcap) * 42;                  // Ws, cs (first: 8-byte aligned)
  b += sizeof(float) * (size_t(8) * dcap * 27 + size_t(dcap) * 63);
  b += sizeof(int) * size_t(dcap) * 2;
  return (b + 15) / 16 * 16;
}

__device__ __forceinline__ FrameSmem carve(unsigned char* base, int dcap) {
  FrameSmem s;
  s.Ws = reinterpret_cast<double*>(base);
  s.cs = s.Ws + size_t(dcap) * 36;
  s.wacc = reinterpret_cast<float*>(s.cs + size_t(dcap) * 6);
  s.Am = s.wacc + size_t(8) * dcap * 27;
  s.Hs = s.Am + size_t(dcap) * 36;
  s.jp = reinterpret_cast<int*>(s.Hs + size_t(dcap) * 27);
  s.slotq = s.jp + dcap;
  return s;
}

// grid = Mcap * S CTAs; CTA (k, s) handles sub-tiles s, s+S, ... of source frame kx[k], thread = pixel.  The pose edges of
// the frame are handled in passes of `dcap` slots (no cross-edge coupling without the depth terms: any degree).
__global__ void __launch_bounds__(kBAThreads)
ba_motion_kernel(BAParams p, int S, int dcap) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  if (p.hdr->status < 0) return;
  const int k = blockIdx.x / S, split = blockIdx.x - k * S;
  if (k >= p.hdr->M) return;
  const int dtot = p.nd[k];
  const FrameSmem sm = carve(smem_raw, dcap);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int f = p.kx[k];
  if (f < p.own_lo || f >= p.own_hi) return;       // depth map owned by another rank
  const int q_beg = p.deg_off[k], ne = p.deg_off[k + 1] - q_beg;
  const int HW = p.HW, wd = p.wd;
  const float fx = p.intr[0], fy = p.intr[1], cx = p.intr[2], cy = p.intr[3];
  const int ip = f - p.t0;                          // pose row of the source frame (<0: fixed)
  const int nsub = cdiv(HW, kBAThreads);
  const float* dsp = p.disps + size_t(f) * HW;

  for (int c0 = 0; c0 < dtot; c0 += dcap) {
    const int d = min(dcap, dtot - c0);             // pose edges in this pass
    __syncthreads();
    for (int i = tid; i < 8 * dcap * 27; i += kBAThreads) sm.wacc[i] = 0.0f;
    for (int qq = tid; qq < ne; qq += kBAThreads) {    // slot -> sorted edge position
      const int sl = p.erec[q_beg + qq].slot - c0;
      if (sl >= 0 && sl < d) sm.slotq[sl] = q_beg + qq;
    }
    __syncthreads();

    for (int sub = split; sub < nsub; sub += S) {
      const int pix = sub * kBAThreads + tid;
      const bool inb = pix < HW;
      const int pi_ = inb ? pix : 0;
      const float h = inb ? dsp[pi_] : 1.0f;
      const float X0 = (float(pi_ % wd) - cx) / fx, X1 = (float(pi_ / wd) - cy) / fy;
      for (int qq = 0; qq < ne; qq++) {
        const EdgeRec r = load_rec(p.erec + q_beg + qq);
        const int sl = (r.slot < 0) ? -1 : r.slot - c0;
        if (sl < 0 || sl >= d) continue;            // stereo edges carry depth terms only: nothing to do here
        const float* tg = p.target + size_t(r.eid) * 2 * HW;
        const float* wg = p.weight + size_t(r.eid) * 2 * HW;
        const float tu = inb ? __ldg(tg + pi_) : 0.0f, tv = inb ? __ldg(tg + HW + pi_) : 0.0f;
        const float wur = inb ? __ldg(wg + pi_) : 0.0f, wvr = inb ? __ldg(wg + HW + pi_) : 0.0f;
        PixEdge pe_;
        pix_edge(r, X0, X1, h, fx, fy, cx, cy, tu, tv, wur, wvr, pe_);
        float hv[32];
        {
          float wJu[6], wJv[6];
#pragma unroll
          for (int n = 0; n < 6; n++) { wJu[n] = pe_.wu * pe_.Ju[n]; wJv[n] = pe_.wv * pe_.Jv[n]; }
          int l = 0;
#pragma unroll
          for (int n = 0; n < 6; n++)
#pragma unroll
            for (int m = 0; m <= n; m++) {
              float a = 0.0f;
              if (n != 1 && m != 1) a = wJu[n] * pe_.Ju[m];
              if (n != 0 && m != 0) a += wJv[n] * pe_.Jv[m];
              hv[l++] = a;
            }
#pragma unroll
          for (int n = 0; n < 6; n++) hv[21 + n] = wJu[n] * pe_.ru + wJv[n] * pe_.rv;       // vj (:335,368)
#pragma unroll
          for (int n = 27; n < 32; n++) hv[n] = 0.0f;
        }
        const float tot = warp_transpose_sum32(hv, lane);
        if (lane < 27) sm.wacc[(size_t(warp) * dcap + sl) * 27 + lane] += tot;
      }
    }
    __syncthreads();

    // ---------------- epilogue: reduce, transform with A_e in fp64, scatter into the pose system
    for (int i = tid; i < d * 27; i += kBAThreads) {  // per-edge sums over the 8 warps, fixed order
      float s = 0.0f;
      for (int w = 0; w < 8; w++) s += sm.wacc[size_t(w) * dcap * 27 + i];
      sm.Hs[i] = s;
    }
    for (int e = tid; e < d; e += kBAThreads) {
      const EdgeRec r = load_rec(p.erec + sm.slotq[e]);
      sm.jp[e] = r.jpose;
      build_A(r, sm.Am + e * 36);
    }
    __syncthreads();
    for (int i = tid; i < d * 6; i += kBAThreads) sm.cs[i] = double(sm.Hs[(i / 6) * 27 + 21 + (i % 6)]);    // c_e = vj_e
    for (int i = tid; i < d * 36; i += kBAThreads) {   // W_e = A_e Hjj_e
      const int e = i / 36, rc = i - e * 36, r = rc / 6, c = rc - r * 6;
      double s = 0.0;
      for (int kk = 0; kk < 6; kk++) s += double(sm.Am[e * 36 + r * 6 + kk]) * double(sm.Hs[e * 27 + hs_index(kk, c)]);
      sm.Ws[i] = s;
    }
    __syncthreads();
    const int n = p.n;
    double* H = p.Hb;
    double* bvec = p.Hb + sys_h_elems(n);
    for (int i = tid; i < d * 36; i += kBAThreads) {
      const int e = i / 36, rc = i - e * 36, r = rc / 6, c = rc - r * 6;
      const int je = sm.jp[e];
      if (je < 0) continue;
      if (r >= c) atomicAdd(&H[sys_index(6 * je + r, 6 * je + c)], double(sm.Hs[e * 27 + hs_index(r, c)]));   // (j_e, j_e) += Hjj_e
      if (ip >= 0) {                                       // (i, j_e) += -W_e ; (j_e, i) += -W_e^T: one of the two is in the lower triangle
        const int ra = 6 * ip + r, ca = 6 * je + c;
        atomicAdd(&H[ra >= ca ? sys_index(ra, ca) : sys_index(ca, ra)], -sm.Ws[i]);
      }
    }
    if (ip >= 0) {
      if (tid < 36) {                                      // (i, i) += sum_e W_e A_e^T
        const int r = tid / 6, c = tid - r * 6;
        double s = 0.0;
        for (int e = 0; e < d; e++)
          for (int kk = 0; kk < 6; kk++) s += sm.Ws[e * 36 + r * 6 + kk] * double(sm.Am[e * 36 + c * 6 + kk]);
        if (r >= c) atomicAdd(&H[sys_index(6 * ip + r, 6 * ip + c)], s);
      } else if (tid >= 64 && tid < 70) {                  // b_i += -sum_e A_e c_e
        const int r = tid - 64;
        double s = 0.0;
        for (int e = 0; e < d; e++)
          for (int kk = 0; kk < 6; kk++) s += double(sm.Am[e * 36 + r * 6 + kk]) * sm.cs[e * 6 + kk];
        atomicAdd(&bvec[6 * ip + r], -s);
      }
    }
    for (int i = tid; i < d * 6; i += kBAThreads) {        // b_{j_e} += c_e
      const int e = i / 6, r = i - e * 6;
      if (sm.jp[e] >= 0) atomicAdd(&bvec[6 * sm.jp[e] + r], sm.cs[i]);
    }
  }
}

// ------------------------------------------------------------------------------------------------
// fused Schur accumulation on the tensor cores (non-motion BA)
constexpr int kChunk = 64;                 // pixels per chunk
constexpr int kStride = kChunk + 4;        // stash row stride (floats): 4*row + col -> conflict-free MMA fragment loads
constexpr int kCwStride = kChunk + 4;      // Cw row stride: the 4 lanes of a pixel read edges q, q+1, q+2, q+3 -> banks 8 apart

__host__ __device__ inline int schur_rows(int d) { return 6 * d + 1; }                    // E rows + the sqrt(Q) w row
__host__ __device__ inline int schur_rpad(int d) { return (schur_rows(d) + 15) / 16 * 16; }
__host__ __device__ inline int schur_gn(int d) { return (6 * d + 7) / 8 * 8; }
constexpr int kSegTiles = 4;               // 16x8 output tiles per segment (one M-row, consecutive N-tiles)
__host__ __device__ inline int schur_segs(int d) {                                        // segments covering the lower triangle
  if (d <= 0) return 0;
  const int nM = (schur_rows(d) + 15) / 16, nN = (6 * d + 7) / 8;
  int t = 0;
  for (int i = 0; i < nM; i++) { const int c = 2 * i + 2; t += ((c < nN ? c : nN) + kSegTiles - 1) / kSegTiles; }
  return t;
}

struct SchurSmem {
  double* Ws; double* cs;                  // [dcap][36], [dcap][6]
  float* shi; float* slo;                  // [rpad][kStride] stash, hi / lo TF32 parts     (aliased by G, red in the epilogue)
  float* G; float* red;                    // [rpad][gn] (aliases the stash), [256][27] (aliases the staging buffers)
  float* stage;                            // [2][necap][4][kChunk] target/weight of the current and the next chunk (cp.async)
  int* eid;                                // [necap] original edge index
  float* Cw;                               // [necap][2][kCwStride]
  float* px;                               // [2][7][kChunk]: X0, X1, disp, sqrt(Q), sqrt(Q) w, disps_sens, eta of this chunk / the next
  float* Am; float* Hs;                    // [dcap][36], [dcap][27]
  int* jp; int* slotq; int* eslot;         // [dcap], [dcap], [necap]
  unsigned char* segs;                     // [nsegs][4]: M-tile, first N-tile, count
};

__host__ __device__ inline size_t schur_alias_floats(int dcap) {
  const size_t stash = size_t(2) * schur_rpad(dcap) * kStride;
  const size_t epi = size_t(schur_rpad(dcap)) * schur_gn(dcap);
  return stash > epi ? stash : epi;
}
__host__ __device__ inline size_t schur_stage_floats(int necap) {
  const size_t st = size_t(2) * necap * 4 * kChunk, red = size_t(256) * 27;
  return st > red ? st : red;
}
__host__ __device__ inline size_t schur_smem_bytes(int dcap, int necap) {
  size_t b = sizeof(double) * size_t(dcap > 0 ? dcap : 1) * 42;
  b += sizeof(float) * (schur_alias_floats(dcap) + schur_stage_floats(necap) + size_t(necap) * 2 * kCwStride + 14 * kChunk + size_t(dcap > 0 ? dcap : 1) * 63);
  b += sizeof(int) * (size_t(2) * (dcap > 0 ? dcap : 1) + 2 * size_t(necap));
  b += 4 * size_t(schur_segs(dcap) + 8);
  return (b + 15) / 16 * 16;
}
__device__ __forceinline__ SchurSmem schur_carve(unsigned char* base, int dcap, int necap) {
  SchurSmem s;
  const int dc = dcap > 0 ? dcap : 1;
  s.Ws = reinterpret_cast<double*>(base);
  s.cs = s.Ws + size_t(dc) * 36;
  s.shi = reinterpret_cast<float*>(s.cs + size_t(dc) * 6);
  s.slo = s.shi + size_t(schur_rpad(dcap)) * kStride;
  s.G = s.shi;
  s.stage = s.shi + schur_alias_floats(dcap);
  s.red = s.stage;
  s.Cw = s.stage + schur_stage_floats(necap);
  s.px = s.Cw + size_t(necap) * 2 * kCwStride;
  s.Am = s.px + 14 * kChunk;
  s.Hs = s.Am + size_t(dc) * 36;
  s.jp = reinterpret_cast<int*>(s.Hs + size_t(dc) * 27);
  s.slotq = s.jp + dc;
  s.eslot = s.slotq + dc;
  s.eid = s.eslot + necap;
  s.segs = reinterpret_cast<unsigned char*>(s.eid + necap);
  return s;
}

// round to TF32 (10 mantissa bits), nearest, ties away from zero: what cvt.rna.tf32.f32 gives for every finite value.  ptxas
// expands that instruction to this add + mask PLUS an |x| < Inf test and a select (4 instructions per value, 13.7 % of the
// kernel's instructions in the r2 profile); Inf stays Inf here too and a NaN stays a NaN.
__device__ __forceinline__ float to_tf32(float x) { return __uint_as_float((__float_as_uint(x) + 0x1000u) & 0xffffe000u); }
// D += A(16x8, row) * B(8x8, col), TF32 inputs, fp32 accumulate
__device__ __forceinline__ void mma_tf32(float (&d)[4], const float (&a)[4], const float (&b)[2]) {
  asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
               : "r"(__float_as_uint(a[0])), "r"(__float_as_uint(a[1])), "r"(__float_as_uint(a[2])), "r"(__float_as_uint(a[3])),
                 "r"(__float_as_uint(b[0])), "r"(__float_as_uint(b[1])));
}

// 16-byte / 4-byte asynchronous global->shared copies; src_bytes = 0 zero-fills (pixels past the end of the map)
__device__ __forceinline__ void cp_async16(unsigned dst, const void* src, int src_bytes) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void cp_async4(unsigned dst, const void* src, int src_bytes) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"(dst), "l"(src), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }

__device__ __forceinline__ void ldsm_x4(float (&r)[4], unsigned addr) {
  unsigned a, b, c, d;
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];" : "=r"(a), "=r"(b), "=r"(c), "=r"(d) : "r"(addr));
  r[0] = __uint_as_float(a); r[1] = __uint_as_float(b); r[2] = __uint_as_float(c); r[3] = __uint_as_float(d);
}

// one segment (CNT consecutive 16x8 tiles of one M-row) over one chunk: straight-line code, no predicates.
// The chunk's 24 MMAs per tile accumulate into a ZEROED temporary that is then added to the running sums with an ordinary
// (round-to-nearest) fp32 add: the tensor core's accumulator addition truncates, and over the 24 chunks x 8 k-steps a CTA
// walks at backend sizes that bias — amplified by the cancellation in D = Hjj - G — showed up as 2e-3 in dx.
template <int CNT>
__device__ __forceinline__ void schur_segment(float (&acc)[kSegTiles][4], unsigned aA, unsigned aB, unsigned lo_off) {
  float tmp[CNT][4];
#pragma unroll
  for (int q = 0; q < CNT; q++) { tmp[q][0] = 0.0f; tmp[q][1] = 0.0f; tmp[q][2] = 0.0f; tmp[q][3] = 0.0f; }
#pragma unroll 1
  for (int k0 = 0; k0 < kChunk; k0 += 8) {
    float ah[4], al[4], bf[CNT][4];
    ldsm_x4(ah, aA);
    ldsm_x4(al, aA + lo_off);
#pragma unroll
    for (int q = 0; q < CNT; q++) ldsm_x4(bf[q], aB + 4u * unsigned(8 * q * kStride));
    // three passes over the segment's tiles: dependent MMAs on one accumulator are CNT apart
#pragma unroll
    for (int q = 0; q < CNT; q++) { const float bh[2] = {bf[q][0], bf[q][1]}; mma_tf32(tmp[q], al, bh); }
#pragma unroll
    for (int q = 0; q < CNT; q++) { const float bl[2] = {bf[q][2], bf[q][3]}; mma_tf32(tmp[q], ah, bl); }
#pragma unroll
    for (int q = 0; q < CNT; q++) { const float bh[2] = {bf[q][0], bf[q][1]}; mma_tf32(tmp[q], ah, bh); }
    aA += 32u; aB += 32u;
  }
#pragma unroll
  for (int q = 0; q < CNT; q++) { acc[q][0] += tmp[q][0]; acc[q][1] += tmp[q][1]; acc[q][2] += tmp[q][2]; acc[q][3] += tmp[q][3]; }
}

// One work item of the Schur accumulation: chunks split, split+S, ... of source frame kx[k].
//   HD = false: all edges of the frame (d <= kMaxDeg pose edges, <= kMaxStereo stereo edges): C, w, Q are formed here.
//   HD = true : the pose edges of slot chunks `ca` >= `cb` of a frame beyond those limits; Q, w were written by
//               ba_hd_depth_kernel.  Local slots: chunk cb first (kHdChunk of them when ca > cb), then chunk ca.  The item
//               owns the blocks (e, f) inside the chunk when ca == cb, and the cross blocks chunk ca x chunk cb (both
//               orders) otherwise: every pair of rows of the depth map is covered exactly once over the frame's items,
//               which is the reference's unbounded pair enumeration (droid_kernels.cu:1244-1272).
template <int NSEG, bool HD>
__device__ __forceinline__ void schur_item(const BAParams& p, unsigned char* smem_raw, int k, int split, int S, int dcap,
                                           int necap, int ca, int cb) {
  const int f = p.kx[k];
  const int q_beg = p.deg_off[k], ne_all = p.deg_off[k + 1] - q_beg;
  int d = p.nd[k], ne = ne_all;
  if (HD) {
    const int na = min(kHdChunk, d - ca * kHdChunk);
    d = (ca == cb) ? na : kHdChunk + na;
    ne = d;
  }
  const bool diag_item = !HD || ca == cb;
  auto own = [&](int e1, int f2) -> bool { return !HD || ca == cb || ((e1 < kHdChunk) != (f2 < kHdChunk)); };
  const SchurSmem sm = schur_carve(smem_raw, dcap, necap);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int HW = p.HW, wd = p.wd;
  const float fx = p.intr[0], fy = p.intr[1], cx = p.intr[2], cy = p.intr[3];
  const int ip = f - p.t0;                          // pose row of the source frame (<0: fixed)
  const int nchunks = cdiv(HW, kChunk);
  const float* dsp = p.disps + size_t(f) * HW;
  const int R = schur_rows(d), nM = (R + 15) / 16, nN = (6 * d + 7) / 8;
  const int rpad = schur_rpad(dcap), gn = schur_gn(dcap);

  // ---- set-up: segment list, stash cleared (padding rows must read as zero), thread -> (edge, lane) map
  int nsegs = 0;
  for (int i = 0; i < nM && d > 0; i++) { const int c = 2 * i + 2; nsegs += ((c < nN ? c : nN) + kSegTiles - 1) / kSegTiles; }
  if (tid == 0) {
    int t = 0;
    for (int i = 0; i < nM && d > 0; i++) {
      const int c = 2 * i + 2, cnt = c < nN ? c : nN;
      for (int j0 = 0; j0 < cnt; j0 += kSegTiles) {
        sm.segs[4 * t] = (unsigned char)i; sm.segs[4 * t + 1] = (unsigned char)j0;
        sm.segs[4 * t + 2] = (unsigned char)(cnt - j0 < kSegTiles ? cnt - j0 : kSegTiles); sm.segs[4 * t + 3] = 0;
        t++;
      }
    }
  }
  for (int i = tid; i < 2 * rpad * kStride; i += kBAThreads) sm.shi[i] = 0.0f;
  if (HD) {
    for (int qq = tid; qq < ne_all; qq += kBAThreads) {
      const int sl = p.erec[q_beg + qq].slot;
      if (sl < 0) continue;                           // stereo edges: depth terms only, already in Q, w
      const int c = sl / kHdChunk;
      int ls = -1;
      if (c == cb) ls = sl - cb * kHdChunk;
      else if (c == ca) ls = kHdChunk + sl - ca * kHdChunk;
      if (ls >= 0) { sm.eslot[ls] = ls; sm.eid[ls] = p.erec[q_beg + qq].eid; sm.slotq[ls] = q_beg + qq; }
    }
  } else {
    for (int qq = tid; qq < ne; qq += kBAThreads) {
      const int sl = p.erec[q_beg + qq].slot;
      sm.eslot[qq] = sl;
      sm.eid[qq] = p.erec[q_beg + qq].eid;
      if (sl >= 0) sm.slotq[sl] = q_beg + qq;
    }
  }
  // pixel lanes per edge: minimise (warps touched) x (pixel iterations)
  int L = 1;
  if (ne > 0) {
    int best = 1 << 30;
    for (int l = 1; l <= kBAThreads / ne && l <= kChunk; l++) {
      const int cost = cdiv(ne * l, 32) * cdiv(kChunk, l);
      if (cost <= best) { best = cost; L = l; }
    }
  }
  const int e = tid / L, g = tid - e * L;
  const bool active = ne > 0 && e < ne;
  EdgeRec rec{};
  int sl = -1;
  if (!HD && active) { rec = load_rec(p.erec + q_beg + e); sl = rec.slot; }
  float hacc[27];
#pragma unroll
  for (int i = 0; i < 27; i++) hacc[i] = 0.0f;
  float accT[NSEG][kSegTiles][4];
#pragma unroll
  for (int u = 0; u < NSEG; u++)
#pragma unroll
    for (int q = 0; q < kSegTiles; q++) { accT[u][q][0] = 0.0f; accT[u][q][1] = 0.0f; accT[u][q][2] = 0.0f; accT[u][q][3] = 0.0f; }
  const int gid = lane >> 2, tig = lane & 3;
  __syncthreads();
  if (HD && active) { rec = load_rec(p.erec + sm.slotq[e]); sl = e; }      // local edge e = local slot e
  // this warp's segments u = 0..NSEG-1 -> segment warp + 8u; ldmatrix row addresses of this lane
  //   A (x4): matrices {rows 0-7,k 0-3}, {rows 8-15,k 0-3}, {rows 0-7,k 4-7}, {rows 8-15,k 4-7}  -> a0..a3
  //   B (x4): matrices {hi,k 0-3}, {hi,k 4-7}, {lo,k 0-3}, {lo,k 4-7} of the 8 rows of one N-tile   -> b0h,b1h,b0l,b1l
  const unsigned shi_base = unsigned(__cvta_generic_to_shared(sm.shi));
  const unsigned lo_off = unsigned(rpad) * kStride * 4u;
  const int lm = lane >> 3, lr = lane & 7;
  unsigned a_addr[NSEG], b_addr[NSEG];
  int seg_cnt[NSEG], seg_i[NSEG], seg_j0[NSEG];
#pragma unroll
  for (int u = 0; u < NSEG; u++) {
    const int sgi = warp + 8 * u;
    seg_cnt[u] = 0; seg_i[u] = 0; seg_j0[u] = 0;
    if (sgi < nsegs) { seg_i[u] = sm.segs[4 * sgi]; seg_j0[u] = sm.segs[4 * sgi + 1]; seg_cnt[u] = sm.segs[4 * sgi + 2]; }
    a_addr[u] = shi_base + 4u * unsigned((16 * seg_i[u] + lr + 8 * (lm & 1)) * kStride + 4 * (lm >> 1));
    b_addr[u] = shi_base + (lm >> 1) * lo_off + 4u * unsigned((8 * seg_j0[u] + lr) * kStride + 4 * (lm & 1));
  }
  // target/weight of one chunk -> staging buffer `buf` (row (q, r) = edge q, r in {tu, tv, wu, wv})
  const unsigned stage_base = unsigned(__cvta_generic_to_shared(sm.stage));
  const bool vec16 = (HW % 4 == 0) && (((reinterpret_cast<uintptr_t>(p.target) | reinterpret_cast<uintptr_t>(p.weight)) & 15) == 0);
  auto stage_chunk = [&](int chunk, int buf) {
    const int cbase = chunk * kChunk;
    const unsigned dst0 = stage_base + unsigned(buf) * unsigned(necap) * 4u * kChunk * 4u;
    if (vec16) {
      for (int i = tid; i < ne * 4 * (kChunk / 4); i += kBAThreads) {
        const int q = i / (4 * (kChunk / 4)), rem = i - q * (4 * (kChunk / 4)), r = rem / (kChunk / 4), sg = rem - r * (kChunk / 4);
        const int pix = cbase + 4 * sg;
        const float* src = (r < 2 ? p.target : p.weight) + size_t(sm.eid[q]) * 2 * HW + size_t(r & 1) * HW + (pix < HW ? pix : 0);
        cp_async16(dst0 + 4u * unsigned((q * 4 + r) * kChunk + 4 * sg), src, pix < HW ? 16 : 0);
      }
    } else {
      for (int i = tid; i < ne * 4 * kChunk; i += kBAThreads) {
        const int q = i / (4 * kChunk), rem = i - q * (4 * kChunk), r = rem / kChunk, px = rem - r * kChunk;
        const int pix = cbase + px;
        const float* src = (r < 2 ? p.target : p.weight) + size_t(sm.eid[q]) * 2 * HW + size_t(r & 1) * HW + (pix < HW ? pix : 0);
        cp_async4(dst0 + 4u * unsigned((q * 4 + r) * kChunk + px), src, pix < HW ? 4 : 0);
      }
    }
    cp_async_commit();
  };
  if (split < nchunks) stage_chunk(split, 0);
  int sbuf = 0;
  // per-pixel rows of one chunk (X0, X1, disparity, sensor disparity, eta) -> px buffer `b`
  auto px_fill = [&](float* pb, int chunk, float h, float sv, float ev) {
    const int pix = chunk * kChunk + tid;
    const int pi_ = pix < HW ? pix : 0;
    pb[tid] = (float(pi_ % wd) - cx) / fx;
    pb[kChunk + tid] = (float(pi_ / wd) - cy) / fy;
    pb[2 * kChunk + tid] = h;
    pb[5 * kChunk + tid] = sv;
    pb[6 * kChunk + tid] = ev;
  };
  auto px_load = [&](int chunk, float& h, float& sv, float& ev) {
    const int pix = chunk * kChunk + tid;
    h = 1.0f; sv = 0.0f; ev = 0.0f;
    if (chunk < nchunks && pix < HW) { h = dsp[pix]; sv = p.sens[size_t(f) * HW + pix]; ev = p.eta[size_t(k) * HW + pix]; }
  };
  float nx_h = 1.0f, nx_s = 0.0f, nx_e = 0.0f;
  if (tid < kChunk) { px_load(split, nx_h, nx_s, nx_e); px_fill(sm.px, split, nx_h, nx_s, nx_e); }
  cp_async_wait_all();
  __syncthreads();

  // Four block-wide barriers per chunk (the kernel's first stall reason at 8 warps per CTA): the chunk's target/weight and
  // per-pixel rows are complete BEFORE the last barrier of the previous chunk — copies are issued at the top of a chunk into
  // the other staging buffer and waited for at its end; the per-pixel rows are loaded at the top and stored at the end.
  int pbuf = 0;
  for (int c = split; c < nchunks; c += S) {
    const int base = c * kChunk;
    if (c + S < nchunks) stage_chunk(c + S, sbuf ^ 1);   // next chunk's copies fly during this chunk's work
    const float* stg = sm.stage + size_t(sbuf) * necap * 4 * kChunk;
    sbuf ^= 1;
    float* pxc = sm.px + pbuf * 7 * kChunk;
    float* pxn = sm.px + (pbuf ^ 1) * 7 * kChunk;
    pbuf ^= 1;
    if (tid < kChunk) px_load(c + S, nx_h, nx_s, nx_e);  // consumed after phase 2
    // ---------------- phase 1
    if (active) {
      const float* se = stg + size_t(e) * 4 * kChunk;
      for (int pp = g; pp < kChunk; pp += L) {
        const float tu = se[pp], tv = se[kChunk + pp], wur = se[2 * kChunk + pp], wvr = se[3 * kChunk + pp];
        PixEdge pe_;
        pix_edge(rec, pxc[pp], pxc[kChunk + pp], pxc[2 * kChunk + pp], fx, fy, cx, cy, tu, tv, wur, wvr, pe_);
        if (!HD) {
          sm.Cw[(size_t(e) * 2 + 0) * kCwStride + pp] = pe_.wu * pe_.Jzu * pe_.Jzu + pe_.wv * pe_.Jzv * pe_.Jzv;   // Cii (:320,353)
          sm.Cw[(size_t(e) * 2 + 1) * kCwStride + pp] = pe_.wu * pe_.ru * pe_.Jzu + pe_.wv * pe_.rv * pe_.Jzv;     // bz  (:321,354)
        }
        if (sl >= 0) {                                                         // stereo edges: depth terms only
          const float au = pe_.wu * pe_.Jzu, av = pe_.wv * pe_.Jzv;
          float* st = sm.shi + size_t(6 * sl) * kStride + pp;
#pragma unroll
          for (int n = 0; n < 6; n++) st[n * kStride] = au * pe_.Ju[n] + av * pe_.Jv[n];   // Eij (:339,372), unscaled for now
          float wJu[6], wJv[6];
#pragma unroll
          for (int n = 0; n < 6; n++) { wJu[n] = pe_.wu * pe_.Ju[n]; wJv[n] = pe_.wv * pe_.Jv[n]; }
          int l = 0;
#pragma unroll
          for (int n = 0; n < 6; n++)
#pragma unroll
            for (int m = 0; m <= n; m++) {
              float a = hacc[l];
              if (n != 1 && m != 1) a = fmaf(wJu[n], pe_.Ju[m], a);
              if (n != 0 && m != 0) a = fmaf(wJv[n], pe_.Jv[m], a);
              hacc[l++] = a;
            }
#pragma unroll
          for (int n = 0; n < 6; n++) hacc[21 + n] = fmaf(wJu[n], pe_.ru, fmaf(wJv[n], pe_.rv, hacc[21 + n]));   // vj (:335,368)
        }
      }
    }
    __syncthreads();
    // ---------------- phase 1b: depth system of the chunk (:1396-1400), scaling by sqrt(Q), TF32 split
    {
      // 4 threads per pixel (lanes 4p..4p+3) sum the edges q = part, part+4, ...; combined in a fixed order by shuffles
      const int pp = tid >> 2, part = tid & 3;
      const int pix = base + pp;
      float Csum = 0.0f, wsum = 0.0f;
      if (!HD) {
        for (int q = part; q < ne; q += 4) { Csum += sm.Cw[(size_t(q) * 2) * kCwStride + pp]; wsum += sm.Cw[(size_t(q) * 2 + 1) * kCwStride + pp]; }
        Csum += __shfl_xor_sync(0xffffffffu, Csum, 1); wsum += __shfl_xor_sync(0xffffffffu, wsum, 1);
        Csum += __shfl_xor_sync(0xffffffffu, Csum, 2); wsum += __shfl_xor_sync(0xffffffffu, wsum, 2);
      }
      if (part == 0) {
        float Q = 0.0f, wv_ = 0.0f;
        if (HD) {
          if (pix < HW) { Q = p.Q[size_t(k) * HW + pix]; wv_ = p.w[size_t(k) * HW + pix]; }   // ba_hd_depth_kernel
        } else if (pix < HW) {
          const float h = pxc[2 * kChunk + pp];
          const float ds = pxc[5 * kChunk + pp];
          const float m = ds > 0.0f ? 1.0f : 0.0f;
          const float C = Csum + m * 0.05f + (1.0f - m) * pxc[6 * kChunk + pp];
          wv_ = wsum - m * 0.05f * (h - ds);
          Q = 1.0f / C;
          p.Q[size_t(k) * HW + pix] = Q;
          p.w[size_t(k) * HW + pix] = wv_;
        }
        const float sq = sqrtf(Q);
        pxc[3 * kChunk + pp] = sq;
        const float vq = sq * wv_;
        const float hi = to_tf32(vq);
        sm.shi[size_t(6 * d) * kStride + pp] = hi;
        sm.slo[size_t(6 * d) * kStride + pp] = to_tf32(vq - hi);
      }
    }
    __syncthreads();
    for (int idx = tid; idx < 6 * d * (kChunk / 4); idx += kBAThreads) {   // 4 pixels per thread and step
      const int m = idx / (kChunk / 4), pp = 4 * (idx - m * (kChunk / 4));
      float4* ph = reinterpret_cast<float4*>(sm.shi + size_t(m) * kStride + pp);
      float4* pl = reinterpret_cast<float4*>(sm.slo + size_t(m) * kStride + pp);
      const float4 sq4 = *reinterpret_cast<const float4*>(pxc + 3 * kChunk + pp);
      float4 x = *ph, hi, lo;
      x.x *= sq4.x; x.y *= sq4.y; x.z *= sq4.z; x.w *= sq4.w;
      hi.x = to_tf32(x.x); hi.y = to_tf32(x.y); hi.z = to_tf32(x.z); hi.w = to_tf32(x.w);
      lo.x = to_tf32(x.x - hi.x); lo.y = to_tf32(x.y - hi.y); lo.z = to_tf32(x.z - hi.z); lo.w = to_tf32(x.w - hi.w);
      *ph = hi; *pl = lo;
    }
    __syncthreads();
    // ---------------- phase 2: G += Es Es^T on the tensor cores (EEt6x6/Ev6x1, :1001-1093)
#pragma unroll
    for (int u = 0; u < NSEG; u++) {
      switch (seg_cnt[u]) {                           // warp-uniform
        case 4: schur_segment<4>(accT[u], a_addr[u], b_addr[u], lo_off); break;
        case 3: schur_segment<3>(accT[u], a_addr[u], b_addr[u], lo_off); break;
        case 2: schur_segment<2>(accT[u], a_addr[u], b_addr[u], lo_off); break;
        case 1: schur_segment<1>(accT[u], a_addr[u], b_addr[u], lo_off); break;
        default: break;
      }
    }
    if (tid < kChunk && c + S < nchunks) px_fill(pxn, c + S, nx_h, nx_s, nx_e);
    cp_async_wait_all();                              // the next chunk's target/weight (issued at the top) have landed
    __syncthreads();
  }

  // ---------------- epilogue: per-edge sums over the pixel lanes, G tiles -> shared, fp64 transforms, scatter
  if (active && sl >= 0) {
#pragma unroll
    for (int i = 0; i < 27; i++) sm.red[size_t(tid) * 27 + i] = hacc[i];
  }
#pragma unroll
  for (int u = 0; u < NSEG; u++)
#pragma unroll
    for (int q = 0; q < kSegTiles; q++)
      if (q < seg_cnt[u]) {
        float* gp = sm.G + size_t(16 * seg_i[u] + gid) * gn + 8 * (seg_j0[u] + q) + 2 * tig;
        gp[0] = accT[u][q][0]; gp[1] = accT[u][q][1]; gp[8 * gn] = accT[u][q][2]; gp[8 * gn + 1] = accT[u][q][3];
      }
  __syncthreads();
  for (int i = tid; i < ne * 27; i += kBAThreads) {       // fixed-order sums over the L lanes of each edge
    const int q = i / 27, l = i - q * 27, s2 = sm.eslot[q];
    if (s2 >= 0) {
      float s = 0.0f;
      for (int gg = 0; gg < L; gg++) s += sm.red[size_t(q * L + gg) * 27 + l];
      sm.Hs[s2 * 27 + l] = s;
    }
  }
  for (int q = tid; q < d; q += kBAThreads) {
    const EdgeRec r = load_rec(p.erec + sm.slotq[q]);
    sm.jp[q] = r.jpose;
    build_A(r, sm.Am + q * 36);
  }
  // the tensor-core tiles cover the lower triangle of G: mirror it so that the products below are plain inner products
  const int n6 = 6 * d;
  for (int i = tid; i < n6 * n6; i += kBAThreads) {
    const int row = i / n6, col = i - row * n6;
    if (col > row) sm.G[size_t(row) * gn + col] = sm.G[size_t(col) * gn + row];       // reads lower entries only, writes upper only
  }
  __syncthreads();
  // D_ef = [e==f] Hjj_e - G_ef;  c_e = vj_e - g_e
  for (int i = tid; i < d * 6; i += kBAThreads) {
    const int e1 = i / 6, r = i - e1 * 6;
    sm.cs[i] = diag_item ? double(sm.Hs[e1 * 27 + 21 + r]) - double(sm.G[size_t(6 * d) * gn + i]) : 0.0;
  }
  for (int i = tid; i < d * 36; i += kBAThreads) {   // W_f = sum_e A_e D_ef over the blocks this item owns
    const int f2 = i / 36, rc = i - f2 * 36, r = rc / 6, c2 = rc - r * 6;
    // owned rows e of column block f: all (fused path, diagonal items) or the other slot chunk (cross items)
    int elo = 0, ehi = d;
    if (HD && !diag_item) { if (f2 < kHdChunk) elo = kHdChunk; else ehi = kHdChunk; }
    double s = 0.0;
    const float* gcol = sm.G + 6 * f2 + c2;
    for (int e1 = elo; e1 < ehi; e1++) {
      const float* a = sm.Am + e1 * 36 + r * 6;
      const float* gp = gcol + size_t(6 * e1) * gn;
#pragma unroll
      for (int kk = 0; kk < 6; kk++) s -= double(a[kk]) * double(gp[size_t(kk) * gn]);
    }
    if (diag_item) {
      const float* a = sm.Am + f2 * 36 + r * 6;
#pragma unroll
      for (int kk = 0; kk < 6; kk++) s += double(a[kk]) * double(sm.Hs[f2 * 27 + hs_index(kk, c2)]);
    }
    sm.Ws[i] = s;
  }
  __syncthreads();
  const int n = p.n;
  double* H = p.Hb;
  double* bvec = p.Hb + sys_h_elems(n);
  for (int i = tid; i < d * d * 36; i += kBAThreads) {   // (j_e, j_f) += D_ef
    const int ef = i / 36, rc = i - ef * 36, e1 = ef / d, f2 = ef - e1 * d, r = rc / 6, c2 = rc - r * 6;
    const int je = sm.jp[e1], jf = sm.jp[f2];
    if (own(e1, f2) && je >= 0 && jf >= 0 && 6 * je + r >= 6 * jf + c2) {
      double v = -double(sm.G[size_t(6 * e1 + r) * gn + 6 * f2 + c2]);
      if (e1 == f2) v += double(sm.Hs[e1 * 27 + hs_index(r, c2)]);
      atomicAdd(&H[sys_index(6 * je + r, 6 * jf + c2)], v);
    }
  }
  if (ip >= 0) {
    for (int i = tid; i < d * 36; i += kBAThreads) {     // (i, j_f) += -W_f ; (j_f, i) += -W_f^T
      const int f2 = i / 36, rc = i - f2 * 36, r = rc / 6, c2 = rc - r * 6;
      const int jf = sm.jp[f2];
      if (jf >= 0) {
        const double v = -sm.Ws[i];
        const int ra = 6 * ip + r, ca = 6 * jf + c2;     // the element and its mirror: one of them is in the lower triangle
        atomicAdd(&H[ra >= ca ? sys_index(ra, ca) : sys_index(ca, ra)], v);
      }
    }
    if (tid < 36) {                                      // (i, i) += sum_f W_f A_f^T
      const int r = tid / 6, c2 = tid - r * 6;
      double s = 0.0;
      for (int f2 = 0; f2 < d; f2++)
        for (int kk = 0; kk < 6; kk++) s += sm.Ws[f2 * 36 + r * 6 + kk] * double(sm.Am[f2 * 36 + c2 * 6 + kk]);
      if (r >= c2) atomicAdd(&H[sys_index(6 * ip + r, 6 * ip + c2)], s);
    } else if (diag_item && tid >= 64 && tid < 70) {     // b_i += -sum_e A_e c_e
      const int r = tid - 64;
      double s = 0.0;
      for (int e1 = 0; e1 < d; e1++)
        for (int kk = 0; kk < 6; kk++) s += double(sm.Am[e1 * 36 + r * 6 + kk]) * sm.cs[e1 * 6 + kk];
      atomicAdd(&bvec[6 * ip + r], -s);
    }
  }
  if (diag_item)
    for (int i = tid; i < d * 6; i += kBAThreads) {      // b_{j_e} += c_e
      const int e1 = i / 6;
      if (sm.jp[e1] >= 0) atomicAdd(&bvec[6 * sm.jp[e1] + (i - e1 * 6)], sm.cs[i]);
    }
}

// grid = Mcap * S CTAs; CTA (k, s) handles chunks s, s+S, ... of source frame kx[k]; launches are bucketed by
// degree: only frames with dlo <= nd[k] <= dcap work here (shared memory and NSEG grow with the bucket).
template <int NSEG>
__global__ void __launch_bounds__(kBAThreads, (NSEG <= 2 ? 2 : 1))
ba_schur_kernel(BAParams p, int S, int dlo, int dcap, int necap) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  if (p.hdr->status < 0) return;
  const int k = blockIdx.x / S, split = blockIdx.x - k * S;
  if (k >= p.hdr->M) return;
  const int d = p.nd[k];
  if (d < dlo || d > dcap) return;
  const int f = p.kx[k];
  if (f < p.own_lo || f >= p.own_hi) return;       // depth map owned by another rank
  if (frame_is_hd(d, p.deg_off[k + 1] - p.deg_off[k])) return;     // -> ba_schur_hd_kernel
  schur_item<NSEG, false>(p, smem_raw, k, split, S, dcap, necap, 0, 0);
}

// ---- high-degree frames (more than kMaxDeg pose edges or kMaxStereo stereo edges) -------------------------------------
// depth system of the frame over ALL its edges (:1396-1400): thread = pixel, edges in edge order; Q, w -> HBM
__global__ void __launch_bounds__(kBAThreads)
ba_hd_depth_kernel(BAParams p) {
  if (p.hdr->status < 0) return;
  const int k = blockIdx.y;
  if (k >= p.hdr->M) return;
  const int q_beg = p.deg_off[k], ne = p.deg_off[k + 1] - q_beg;
  if (!frame_is_hd(p.nd[k], ne)) return;
  const int f = p.kx[k];
  if (f < p.own_lo || f >= p.own_hi) return;
  const int pix = blockIdx.x * kBAThreads + threadIdx.x;
  if (pix >= p.HW) return;
  const int HW = p.HW;
  const float fx = p.intr[0], fy = p.intr[1], cx = p.intr[2], cy = p.intr[3];
  const float h = p.disps[size_t(f) * HW + pix];
  const float X0 = (float(pix % p.wd) - cx) / fx, X1 = (float(pix / p.wd) - cy) / fy;
  float Csum = 0.0f, wsum = 0.0f;
  for (int qq = 0; qq < ne; qq++) {
    const EdgeRec r = load_rec(p.erec + q_beg + qq);
    const float* tg = p.target + size_t(r.eid) * 2 * HW;
    const float* wg = p.weight + size_t(r.eid) * 2 * HW;
    PixEdge pe_;
    pix_edge(r, X0, X1, h, fx, fy, cx, cy, __ldg(tg + pix), __ldg(tg + HW + pix), __ldg(wg + pix), __ldg(wg + HW + pix), pe_);
    Csum += pe_.wu * pe_.Jzu * pe_.Jzu + pe_.wv * pe_.Jzv * pe_.Jzv;
    wsum += pe_.wu * pe_.ru * pe_.Jzu + pe_.wv * pe_.rv * pe_.Jzv;
  }
  const float ds = p.sens[size_t(f) * HW + pix];
  const float m = ds > 0.0f ? 1.0f : 0.0f;
  const float C = Csum + m * 0.05f + (1.0f - m) * p.eta[size_t(k) * HW + pix];
  p.Q[size_t(k) * HW + pix] = 1.0f / C;
  p.w[size_t(k) * HW + pix] = wsum - m * 0.05f * (h - ds);
}

// persistent CTAs walk the work items (frame, chunk pair, pixel split); hd_off[] = prefix sums of items per frame
template <int NSEG>
__global__ void __launch_bounds__(kBAThreads, 1)
ba_schur_hd_kernel(BAParams p, int S, int dcap, int necap) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  if (p.hdr->status < 0) return;
  const int M = p.hdr->M;
  const long long total = static_cast<long long>(p.hdr->hd_items) * S;
  for (long long wi = blockIdx.x; wi < total; wi += gridDim.x) {
    const int item = int(wi / S), split = int(wi - static_cast<long long>(item) * S);
    int lo = 0, hi = M;                                  // largest k with hd_off[k] <= item
    while (hi - lo > 1) { const int mid = (lo + hi) >> 1; if (p.hd_off[mid] <= item) lo = mid; else hi = mid; }
    const int k = lo, t = item - p.hd_off[k];
    int ca = int((sqrtf(8.0f * float(t) + 1.0f) - 1.0f) * 0.5f);
    while (ca * (ca + 1) / 2 > t) ca--;
    while ((ca + 1) * (ca + 2) / 2 <= t) ca++;
    const int cb = t - ca * (ca + 1) / 2;
    const int f = p.kx[k];
    if (f >= p.own_lo && f < p.own_hi) schur_item<NSEG, true>(p, smem_raw, k, split, S, dcap, necap, ca, cb);
    __syncthreads();                                     // the next item reuses the shared-memory carve-up
  }
}

// ------------------------------------------------------------------------------------------------
// depth back-substitution + disparity retraction (EvT6x1_kernel, dz = Q (w - sum E^T dx), disp_retr_kernel;
// droid_kernels.cu:1095-1115, :1408-1428).  E_e is recomputed from the edge records of THIS iteration
// (old poses, old disparities); y_e = [0<j_e-t0<P] dx_j - [0<i-t0<P] A_e^T dx_i comes from ba_post_solve.
__global__ void __launch_bounds__(kBAThreads)
ba_backsub_kernel(BAParams p, float* __restrict__ dz_out) {
  if (p.hdr->status < 0) return;
  const int k = blockIdx.y;
  if (k >= p.hdr->M) return;
  const int pix = blockIdx.x * kBAThreads + threadIdx.x;
  if (pix >= p.HW) return;
  const int f = p.kx[k];
  if (f < p.own_lo || f >= p.own_hi) return;
  const int q_beg = p.deg_off[k], ne = p.deg_off[k + 1] - q_beg;
  const float fx = p.intr[0], fy = p.intr[1], cx = p.intr[2], cy = p.intr[3];
  const float h = p.disps[size_t(f) * p.HW + pix];
  const float X0 = (float(pix % p.wd) - cx) / fx, X1 = (float(pix / p.wd) - cy) / fy;
  float acc = 0.0f;
  for (int qq = 0; qq < ne; qq++) {
    const EdgeRec r = load_rec(p.erec + q_beg + qq);
    if (r.slot < 0) continue;
    const float* wg = p.weight + size_t(r.eid) * 2 * p.HW;
    PixEdge pe_;
    pix_edge(r, X0, X1, h, fx, fy, cx, cy, 0.0f, 0.0f, __ldg(wg + pix), __ldg(wg + p.HW + pix), pe_);
    const float au = pe_.wu * pe_.Jzu, av = pe_.wv * pe_.Jzv;
    const float* y = p.yvec + size_t(q_beg + qq) * 6;
    float dw = 0.0f;
#pragma unroll
    for (int n = 0; n < 6; n++) dw = fmaf(au * pe_.Ju[n] + av * pe_.Jv[n], __ldg(y + n), dw);
    acc += dw;
  }
  const size_t o = size_t(k) * p.HW + pix;
  const float dz = p.Q[o] * (p.w[o] - acc);
  p.disps[size_t(f) * p.HW + pix] = h + dz;
  if (dz_out) dz_out[o] = dz;
}

static int launch_motion(const BAParams& p, int S, int dcap, cudaStream_t st) {
  const size_t smem = frame_smem_bytes(dcap);
  if (smem > 48 * 1024) DB200_CUDA(cudaFuncSetAttribute(ba_motion_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
  ba_motion_kernel<<<p.Mcap * S, kBAThreads, smem, st>>>(p, S, dcap);
  DB200_CHECK_LAUNCH();
  return DB200_OK;
}

template <int NSEG>
static int launch_schur(const BAParams& p, int S, int dlo, int dcap, cudaStream_t st) {
  const int necap = dcap + 4;
  const size_t smem = schur_smem_bytes(dcap, necap);
  auto kern = ba_schur_kernel<NSEG>;
  if (smem > 48 * 1024) DB200_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
  kern<<<p.Mcap * S, kBAThreads, smem, st>>>(p, S, dlo, dcap, necap);
  DB200_CHECK_LAUNCH();
  return DB200_OK;
}

static int launch_schur_hd(const BAParams& p, int S, cudaStream_t st) {
  dim3 grid(cdiv(p.HW, kBAThreads), p.Mcap);
  ba_hd_depth_kernel<<<grid, kBAThreads, 0, st>>>(p);
  DB200_CHECK_LAUNCH();
  const int dcap = kMaxDeg, necap = kMaxDeg + kMaxStereo;
  const size_t smem = schur_smem_bytes(dcap, necap);
  auto kern = ba_schur_hd_kernel<5>;
  DB200_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
  kern<<<kNumSMs, kBAThreads, smem, st>>>(p, S, dcap, necap);
  DB200_CHECK_LAUNCH();
  return DB200_OK;
}

// side streams + fork/join events of the accumulate stage, one set per device (a process may drive several GPUs)
namespace {
struct AuxStreams {
  cudaStream_t aux[4] = {};
  cudaEvent_t fork = nullptr, join[4] = {};
  bool ok = false;
};
AuxStreams* aux_streams() {
  static AuxStreams table[64];
  static std::mutex mu;
  int dev = -1;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return nullptr;
  std::lock_guard<std::mutex> lock(mu);
  AuxStreams& a = table[dev];
  if (!a.ok) {
    if (cudaEventCreateWithFlags(&a.fork, cudaEventDisableTiming) != cudaSuccess) return nullptr;
    for (int i = 0; i < 4; i++)
      if (cudaStreamCreateWithFlags(&a.aux[i], cudaStreamNonBlocking) != cudaSuccess ||
          cudaEventCreateWithFlags(&a.join[i], cudaEventDisableTiming) != cudaSuccess)
        return nullptr;
    a.ok = true;
  }
  return &a;
}
}  // namespace

int ba_accumulate_frames(const BAParams& p, cudaStream_t st) {
  DB200_CUDA(cudaMemsetAsync(p.Hb, 0, sizeof(double) * sys_elems(p.n), st));
  if (p.motion_only) {
    const int nsub = cdiv(p.HW, kBAThreads);
    int S = cdiv(4 * kNumSMs, p.Mcap);
    S = S < 1 ? 1 : (S > nsub ? nsub : S);
    return launch_motion(p, S, 32, st);
  }
  const int nchunks = cdiv(p.HW, kChunk);
  // CTAs per launch: ~6 per SM for backend graphs (hundreds of frames); a frontend window (8..25 frames) is fastest with
  // ~1.3 per SM — its CTAs have few chunks each and the per-CTA set-up dominates (measured: 25 frames best at 8 splits,
  // 8 frames at 24, 512 frames at 2).  The result does not depend on the split beyond fp32 round-off since the chunk
  // products are added round-to-nearest (schur_segment): tools/parity_vs_split.py, 6.7e-6 vs 5.8e-6 against the fp64 oracle
  // at 8 / 36 splits on the frontend problem.
  int ctas = 2 * p.Mcap;
  ctas = ctas < 192 ? 192 : (ctas > 6 * kNumSMs ? 6 * kNumSMs : ctas);
  int S = cdiv(ctas, p.Mcap);
  static const int s_env = getenv("DB200_BA_SPLIT") ? atoi(getenv("DB200_BA_SPLIT")) : 0;   // tuning aid
  if (s_env > 0) S = s_env;
  // ... but at most 12 chunks per CTA: the per-thread fp32 sums and the running G sums are chains over a CTA's chunks.  On
  // an ill-conditioned backend graph (160 keyframes / 1600 edges; the fp32 oracle's dx is 1.0e-3 from the fp64 oracle's)
  // splits 2, 3 (24, 16 chunks per CTA) give 1.9e-3 against the fp64 oracle, splits 4, 6, 8, 12, 24 give 1.0e-3, 2e-4,
  // 8e-4, 5e-4, 3e-4 — the scatter of any fp32 grouping, the reference's included (tools/parity_vs_split.py).  +5 % time at c3.
  if (s_env <= 0 && S < cdiv(nchunks, 12)) S = cdiv(nchunks, 12);
  S = S < 1 ? 1 : (S > nchunks ? nchunks : S);
  // degree buckets (shared memory and accumulator registers grow with the bucket's cap); a frame works in exactly one.
  // NSEG = ceil(schur_segs(dcap) / 8): 6, 12, 25, 36 segments of <= 4 tiles for dcap = 8, 14, 22, 28.
  const int dmax_host = p.N < kMaxDeg ? p.N : kMaxDeg;   // the host only knows d <= N
  // The buckets are independent (disjoint frames; the reduced system is updated with atomics), and each one alone
  // leaves SMs idle in its last wave: fork them onto side streams so that they overlap, join before returning.
  AuxStreams* ax = aux_streams();
  if (!ax) return DB200_ERR_CUDA;
  const bool b1 = dmax_host > 8, b2 = dmax_host > 14, b3 = dmax_host > 22;
  const bool hd = p.N > kMaxStereo;                      // a frame can only exceed the fused path's limits with > 4 edges
  if (!b1 && !hd) return launch_schur<1>(p, S, 0, 8, st);
  DB200_CUDA(cudaEventRecord(ax->fork, st));
  int rc = b1 ? launch_schur<2>(p, S, 9, 14, st)         // the bucket that carries most frames of a backend graph
              : launch_schur<1>(p, S, 0, 8, st);
  auto side = [&](int i, auto fn) -> int {
    DB200_CUDA(cudaStreamWaitEvent(ax->aux[i], ax->fork, 0));
    const int r = fn(ax->aux[i]);
    DB200_CUDA(cudaEventRecord(ax->join[i], ax->aux[i]));
    DB200_CUDA(cudaStreamWaitEvent(st, ax->join[i], 0));
    return r;
  };
  if (rc == DB200_OK && b1) rc = side(0, [&](cudaStream_t s2) { return launch_schur<1>(p, S, 0, 8, s2); });
  if (rc == DB200_OK && b2) rc = side(1, [&](cudaStream_t s2) { return launch_schur<4>(p, S, 15, 22, s2); });
  if (rc == DB200_OK && b3) rc = side(2, [&](cudaStream_t s2) { return launch_schur<5>(p, S, 23, kMaxDeg, s2); });
  // frames beyond kMaxDeg pose edges / kMaxStereo stereo edges (known on the device only: both launches return at once
  // when there are none): depth system over all edges of the frame, then the (frame, chunk pair) work items
  if (rc == DB200_OK && hd) rc = side(3, [&](cudaStream_t s2) { return launch_schur_hd(p, S, s2); });
  return rc;
}

int ba_backsub_frames(const BAParams& p, float* dz_out, cudaStream_t st) {
  dim3 grid(cdiv(p.HW, kBAThreads), p.Mcap);
  ba_backsub_kernel<<<grid, kBAThreads, 0, st>>>(p, dz_out);
  DB200_CHECK_LAUNCH();
  return DB200_OK;
}

}  // namespace db200
