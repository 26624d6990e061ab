// ba_frame.cu — fused per-source-frame accumulation of the dense BA and the depth back-substitution.
// Replaces projective_transform_kernel, accum_kernel, EEt6x6_kernel, Ev6x1_kernel, EvT6x1_kernel,
// disp_retr_kernel and the host-side block assembly of /root/reference/src/droid_kernels.cu
// (:176-424, :854-874, :1001-1115, :933-946, :1117-1190, :1222-1311).
//
// One CTA = (source frame k, pixel split s), 256 threads = 256 pixels per sub-tile.
//   phase 1 (thread = pixel): loop over the frame's edges; reprojection, residual, J_j, J_z;
//           C,w summed in registers; E_e[6] written to the shared-memory stash; the 27 per-edge sums
//           (Hjj lower triangle + vj) reduced with a transposing shuffle butterfly (31 shuffles for 32
//           values) into per-warp shared accumulators — no global atomics, no HBM round trip for E.
//   phase 2 (thread = (edge pair, k-group)): G_ef += (E_e Q) E_f^T over the sub-tile from the stash,
//           36 register accumulators per pair, persistent over the CTA's sub-tiles.
//   epilogue: fp64 6x6 transforms with A_e and scatter of this CTA's contribution into the reduced
//           system with fp64 red.global.add (the only atomics; they carry ~d^2*36 values per CTA).
// HBM traffic per edge-pixel: target 8 B + weight 8 B (+ disp 4 B per keyframe-pixel) — the algorithmic
// 20*HW*N + 16*HW*M of SURVEY.md §8d; the back-substitution re-reads weight (8 B) to recompute E.
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
  o.wu = bad ? 0.0f : float(0.001 * double(wur));   // `.001 * weight` is a double product in the source (:305)
  o.wv = bad ? 0.0f : float(0.001 * double(wvr));
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

__device__ __forceinline__ int pair_index(int e, int f, int d) {  // e <= f
  return e * d - (e * (e - 1)) / 2 + (f - e);
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
// shared-memory carve-up (all sizes derive from `dcap` = max pose edges per frame in this launch)
struct FrameSmem {
  float* stash;   // [dcap][6][kEStride]   (aliased by the epilogue item buffer)
  float* Qs;      // [256]
  float* Qws;     // [256]
  float* wacc;    // [8][dcap][27]
  double* Ws;     // [dcap][36]
  double* cs;     // [dcap][6]
  float* Am;      // [dcap][36]
  float* Hs;      // [dcap][27]
  int* jp;        // [dcap]
  int* slotq;     // [dcap]
  unsigned short* pr;  // [np_cap][2]
};

__host__ __device__ inline size_t stash_floats(int dcap, bool motion) {
  if (motion) return 64;
  size_t s = size_t(dcap) * 6 * kEStride;
  const size_t min_items = 42 * 4 * 8;   // tiny d: room for a few epilogue items
  return s < min_items ? min_items : s;
}

__host__ __device__ inline size_t frame_smem_bytes(int dcap, bool motion) {
  const int np = dcap * (dcap + 1) / 2;
  size_t b = 0;
  b += sizeof(double) * size_t(dcap) * 42;                       // Ws, cs (first: 8-byte aligned)
  b += sizeof(float) * (stash_floats(dcap, motion) + 512 + size_t(8) * dcap * 27 + size_t(dcap) * 63);
  b += sizeof(int) * size_t(dcap) * 2;
  b += sizeof(unsigned short) * 2 * size_t(np > 0 ? np : 1);
  return (b + 15) / 16 * 16;
}

__device__ __forceinline__ FrameSmem carve(unsigned char* base, int dcap, bool motion) {
  FrameSmem s;
  s.Ws = reinterpret_cast<double*>(base);
  s.cs = s.Ws + size_t(dcap) * 36;
  s.stash = reinterpret_cast<float*>(s.cs + size_t(dcap) * 6);
  s.Qs = s.stash + stash_floats(dcap, motion);
  s.Qws = s.Qs + 256;
  s.wacc = s.Qws + 256;
  s.Am = s.wacc + size_t(8) * dcap * 27;
  s.Hs = s.Am + size_t(dcap) * 36;
  s.jp = reinterpret_cast<int*>(s.Hs + size_t(dcap) * 27);
  s.slotq = s.jp + dcap;
  s.pr = reinterpret_cast<unsigned short*>(s.slotq + dcap);
  return s;
}

// ------------------------------------------------------------------------------------------------
// grid = Mcap * S CTAs; CTA (k, s) handles sub-tiles s, s+S, ... of source frame kx[k].
// Non-motion launches are bucketed by degree: only frames with dlo <= nd[k] <= dcap do work here.
template <int PPT, bool MOTION>
__global__ void __launch_bounds__(kBAThreads)
ba_frame_kernel(BAParams p, int S, int dlo, int dcap) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  if (p.hdr->status < 0) return;
  const int k = blockIdx.x / S, split = blockIdx.x - k * S;
  if (k >= p.hdr->M) return;
  const int dtot = p.nd[k];
  if (!MOTION && (dtot < dlo || dtot > dcap)) return;
  const FrameSmem sm = carve(smem_raw, dcap, MOTION);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int f = p.kx[k];
  if (f < p.own_lo || f >= p.own_hi) return;       // depth map owned by another rank
  const int q_beg = p.deg_off[k], ne = p.deg_off[k + 1] - q_beg;
  const int HW = p.HW, wd = p.wd;
  const float fx = p.intr[0], fy = p.intr[1], cx = p.intr[2], cy = p.intr[3];
  const int ip = f - p.t0;                          // pose row of the source frame (<0: fixed)
  const int nsub = cdiv(HW, kBAThreads);
  const float* dsp = p.disps + size_t(f) * HW;

  // MOTION handles its pose edges in chunks of `dcap` slots (no cross-edge coupling); the fused path
  // has exactly one chunk (dtot <= dcap).
  for (int c0 = 0; c0 < (MOTION ? max(dtot, 1) : 1); c0 += dcap) {
    const int d = MOTION ? min(dcap, dtot - c0) : dtot;   // pose edges in this chunk
    const int np = d * (d + 1) / 2;
    int kg = 1;
    if (!MOTION && np > 0) {
      kg = min(4, (kBAThreads * PPT) / np);
      while (kg > 1 && size_t(np) * kg * 42 > stash_floats(dcap, false)) kg--;
    }
    const int items = np * kg;
    __syncthreads();
    for (int i = tid; i < 8 * dcap * 27; i += kBAThreads) sm.wacc[i] = 0.0f;
    if (!MOTION) {
      for (int e = tid; e < d; e += kBAThreads) {      // pair table: rows of pair (e, f>=e) are consecutive
        int pidx = pair_index(e, e, d);
        for (int g = e; g < d; g++) { sm.pr[2 * pidx] = (unsigned short)e; sm.pr[2 * pidx + 1] = (unsigned short)g; pidx++; }
      }
    }
    for (int qq = tid; qq < ne; qq += kBAThreads) {    // slot -> sorted edge position
      const int sl = p.erec[q_beg + qq].slot - c0;
      if (sl >= 0 && sl < d) sm.slotq[sl] = q_beg + qq;
    }
    __syncthreads();

    float acc[PPT][36], gacc[PPT][6];
    int pe[PPT], pf[PPT], pg[PPT];
#pragma unroll
    for (int s = 0; s < PPT; s++) {
#pragma unroll
      for (int i = 0; i < 36; i++) acc[s][i] = 0.0f;
#pragma unroll
      for (int i = 0; i < 6; i++) gacc[s][i] = 0.0f;
      const int item = tid + kBAThreads * s;
      if (!MOTION && item < items) { const int pi = item / kg; pe[s] = sm.pr[2 * pi]; pf[s] = sm.pr[2 * pi + 1]; pg[s] = item - pi * kg; }
      else { pe[s] = -1; pf[s] = 0; pg[s] = 0; }
    }

    for (int sub = split; sub < nsub; sub += S) {
      const int pix = sub * kBAThreads + tid;
      const bool inb = pix < HW;
      const int pi_ = inb ? pix : 0;
      const float h = inb ? dsp[pi_] : 1.0f;
      const float X0 = (float(pi_ % wd) - cx) / fx, X1 = (float(pi_ / wd) - cy) / fy;
      float Csum = 0.0f, wsum = 0.0f;
      // ---------------- phase 1
      for (int qq = 0; qq < ne; qq++) {
        const EdgeRec r = load_rec(p.erec + q_beg + qq);
        const int sl = (r.slot < 0) ? -1 : r.slot - c0;
        if (MOTION && (sl < 0 || sl >= d)) continue;
        if (!MOTION && c0 != 0) continue;
        const float* tg = p.target + size_t(r.eid) * 2 * HW;
        const float* wg = p.weight + size_t(r.eid) * 2 * HW;
        const float tu = inb ? __ldg(tg + pi_) : 0.0f, tv = inb ? __ldg(tg + HW + pi_) : 0.0f;
        const float wur = inb ? __ldg(wg + pi_) : 0.0f, wvr = inb ? __ldg(wg + HW + pi_) : 0.0f;
        PixEdge pe_;
        pix_edge(r, X0, X1, h, fx, fy, cx, cy, tu, tv, wur, wvr, pe_);
        if (!MOTION) {
          Csum += pe_.wu * pe_.Jzu * pe_.Jzu + pe_.wv * pe_.Jzv * pe_.Jzv;   // Cii (:320,353)
          wsum += pe_.wu * pe_.ru * pe_.Jzu + pe_.wv * pe_.rv * pe_.Jzv;     // bz  (:321,354)
        }
        if (sl < 0) continue;                                                 // stereo edge: depth terms only
        const float au = pe_.wu * pe_.Jzu, av = pe_.wv * pe_.Jzv;
        if (!MOTION) {
          float* st = sm.stash + size_t(sl) * 6 * kEStride + tid;
#pragma unroll
          for (int n = 0; n < 6; n++) st[n * kEStride] = au * pe_.Ju[n] + av * pe_.Jv[n];   // Eij (:339,372)
        }
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
      if (!MOTION) {
        // depth system of this pixel (:1396-1400): C = sum Cii + m*alpha + (1-m)*eta; w = sum bz - m*alpha*(d - d_sens)
        float Q = 0.0f, wv_ = 0.0f;
        if (inb) {
          const float ds = p.sens[size_t(f) * HW + pix];
          const float m = ds > 0.0f ? 1.0f : 0.0f;
          const float C = Csum + m * 0.05f + (1.0f - m) * p.eta[size_t(k) * HW + pix];
          wv_ = wsum - m * 0.05f * (h - ds);
          Q = 1.0f / C;
          p.Q[size_t(k) * HW + pix] = Q;
          p.w[size_t(k) * HW + pix] = wv_;
        }
        sm.Qs[tid] = inb ? Q : 0.0f;
        sm.Qws[tid] = inb ? Q * wv_ : 0.0f;
        if (!inb) {   // keep the stash finite for padded pixels
          for (int sl = 0; sl < d; sl++)
            for (int n = 0; n < 6; n++) sm.stash[(size_t(sl) * 6 + n) * kEStride + tid] = 0.0f;
        }
        __syncthreads();
        // ---------------- phase 2: G_ef += (E_e Q) E_f^T, g_e += E_e (Q w)   (EEt6x6/Ev6x1, :1001-1093)
#pragma unroll
        for (int s = 0; s < PPT; s++) {
          if (pe[s] < 0) continue;
          const float* Ee = sm.stash + size_t(pe[s]) * 6 * kEStride;
          const float* Ef = sm.stash + size_t(pf[s]) * 6 * kEStride;
          const bool diag = pe[s] == pf[s];
          for (int px = pg[s]; px < kBAThreads; px += kg) {
            const float q = sm.Qs[px];
            float ee[6], ff[6];
#pragma unroll
            for (int n = 0; n < 6; n++) { ee[n] = Ee[n * kEStride + px]; ff[n] = Ef[n * kEStride + px]; }
            if (diag) {
              const float qw = sm.Qws[px];
#pragma unroll
              for (int n = 0; n < 6; n++) gacc[s][n] = fmaf(qw, ee[n], gacc[s][n]);
            }
#pragma unroll
            for (int n = 0; n < 6; n++) {
              const float eq = ee[n] * q;
#pragma unroll
              for (int m = 0; m < 6; m++) acc[s][n * 6 + m] = fmaf(eq, ff[m], acc[s][n * 6 + m]);
            }
          }
        }
        __syncthreads();
      }
    }
    __syncthreads();

    // ---------------- epilogue: reduce, transform with A_e in fp64, scatter into the reduced system
    float* itembuf = sm.stash;                       // [items][42], aliases the stash (phase 2 is done)
    if (!MOTION) {
#pragma unroll
      for (int s = 0; s < PPT; s++) {
        const int item = tid + kBAThreads * s;
        if (pe[s] >= 0) {
#pragma unroll
          for (int i = 0; i < 36; i++) itembuf[size_t(item) * 42 + i] = acc[s][i];
#pragma unroll
          for (int i = 0; i < 6; i++) itembuf[size_t(item) * 42 + 36 + i] = gacc[s][i];
        }
      }
    }
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
    if (!MOTION) {                                     // sum the k-groups of each pair into its first item
      for (int i = tid; i < np * 42; i += kBAThreads) {
        const int pi = i / 42, c = i - pi * 42;
        float s = 0.0f;
        for (int g = 0; g < kg; g++) s += itembuf[(size_t(pi) * kg + g) * 42 + c];
        itembuf[size_t(pi) * kg * 42 + c] = s;
      }
      __syncthreads();
    }
    // D_ef[r][c] = [e==f] Hjj_e[r][c] - G_ef[r][c];  c_e = vj_e - g_e
    auto Dval = [&](int e, int f2, int r, int c) -> double {
      double v = (e == f2) ? double(sm.Hs[e * 27 + hs_index(r, c)]) : 0.0;
      if (!MOTION) {
        v -= (e <= f2) ? double(itembuf[size_t(pair_index(e, f2, d)) * kg * 42 + r * 6 + c])
                       : double(itembuf[size_t(pair_index(f2, e, d)) * kg * 42 + c * 6 + r]);
      }
      return v;
    };
    for (int i = tid; i < d * 6; i += kBAThreads) {
      const int e = i / 6, r = i - e * 6;
      double v = double(sm.Hs[e * 27 + 21 + r]);
      if (!MOTION) v -= double(itembuf[size_t(pair_index(e, e, d)) * kg * 42 + 36 + r]);
      sm.cs[i] = v;
    }
    for (int i = tid; i < d * 36; i += kBAThreads) {   // W_f = sum_e A_e D_ef
      const int f2 = i / 36, rc = i - f2 * 36, r = rc / 6, c = rc - r * 6;
      double s = 0.0;
      for (int e = 0; e < d; e++)
        for (int kk = 0; kk < 6; kk++) s += double(sm.Am[e * 36 + r * 6 + kk]) * Dval(e, f2, kk, c);
      sm.Ws[i] = s;
    }
    __syncthreads();
    const int n = p.n;
    double* H = p.Hb;
    double* bvec = p.Hb + size_t(n) * n;
    for (int i = tid; i < d * d * 36; i += kBAThreads) {   // (j_e, j_f) += D_ef
      const int ef = i / 36, rc = i - ef * 36, e = ef / d, f2 = ef - e * d, r = rc / 6, c = rc - r * 6;
      const int je = sm.jp[e], jf = sm.jp[f2];
      if (je >= 0 && jf >= 0) atomicAdd(&H[size_t(6 * je + r) * n + 6 * jf + c], Dval(e, f2, r, c));
    }
    if (ip >= 0) {
      for (int i = tid; i < d * 36; i += kBAThreads) {     // (i, j_f) += -W_f ; (j_f, i) += -W_f^T
        const int f2 = i / 36, rc = i - f2 * 36, r = rc / 6, c = rc - r * 6;
        const int jf = sm.jp[f2];
        if (jf >= 0) {
          const double v = -sm.Ws[i];
          atomicAdd(&H[size_t(6 * ip + r) * n + 6 * jf + c], v);
          atomicAdd(&H[size_t(6 * jf + c) * n + 6 * ip + r], v);
        }
      }
      if (tid < 36) {                                      // (i, i) += sum_f W_f A_f^T
        const int r = tid / 6, c = tid - r * 6;
        double s = 0.0;
        for (int f2 = 0; f2 < d; f2++)
          for (int kk = 0; kk < 6; kk++) s += sm.Ws[f2 * 36 + r * 6 + kk] * double(sm.Am[f2 * 36 + c * 6 + kk]);
        atomicAdd(&H[size_t(6 * ip + r) * n + 6 * ip + c], s);
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

template <int PPT, bool MOTION>
static int launch_frames(const BAParams& p, int S, int dlo, int dcap, cudaStream_t st) {
  const size_t smem = frame_smem_bytes(dcap, MOTION);
  auto kern = ba_frame_kernel<PPT, MOTION>;
  if (smem > 48 * 1024) DB200_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
  kern<<<p.Mcap * S, kBAThreads, smem, st>>>(p, S, dlo, dcap);
  DB200_CHECK_LAUNCH();
  return DB200_OK;
}

int ba_accumulate_frames(const BAParams& p, cudaStream_t st) {
  DB200_CUDA(cudaMemsetAsync(p.Hb, 0, sizeof(double) * (size_t(p.n) * p.n + p.n), st));
  const int nsub = cdiv(p.HW, kBAThreads);
  int S = cdiv(4 * kNumSMs, p.Mcap);               // aim at >= ~4 CTAs per SM
  S = S < 1 ? 1 : (S > nsub ? nsub : S);
  if (p.motion_only) return launch_frames<1, true>(p, S, 0, 32, st);
  // degree buckets (shared memory per CTA grows with the bucket's cap); a frame does its work in exactly one
  const int dmax_host = p.N < kMaxDeg ? p.N : kMaxDeg;   // the host only knows d <= N
  int rc = launch_frames<1, false>(p, S, 0, 8, st);
  if (rc == DB200_OK && dmax_host > 8) rc = launch_frames<1, false>(p, S, 9, 14, st);
  if (rc == DB200_OK && dmax_host > 14) rc = launch_frames<1, false>(p, S, 15, 22, st);
  if (rc == DB200_OK && dmax_host > 22) rc = launch_frames<2, false>(p, S, 23, kMaxDeg, st);
  return rc;
}

int ba_backsub_frames(const BAParams& p, float* dz_out, cudaStream_t st) {
  dim3 grid(cdiv(p.HW, kBAThreads), p.Mcap);
  ba_backsub_kernel<<<grid, kBAThreads, 0, st>>>(p, dz_out);
  DB200_CHECK_LAUNCH();
  return DB200_OK;
}

}  // namespace db200
