// corr_lookup.cu — correlation-pyramid window lookup, forward and backward.
// Replaces /root/reference/src/correlation_kernels.cu:20-124 and the per-level loop + torch.cat of
// droid_slam/modules/corr.py:40-50.
//
// Layout/roofline: volume[n][y][x][h2][w2] — every source pixel (n,y,x) owns a contiguous h2*w2 slice
// and reads a (2r+2)^2 window of it; output corr[n][ch][y][x] is x-contiguous.  The reference maps one
// THREAD to one source pixel (neighbouring threads read slices h2*w2 elements apart, 4 RMWs per output
// cell).  Here one WARP owns one source pixel: lanes cover the window rows (coalesced per row), the
// window goes through shared memory, every output cell is produced once in registers in the reference's
// accumulation order (bit-identical fp16/fp32/fp64), and a CTA covers 64 consecutive x so that each
// output channel row is written as one full coalesced line.  HBM-bound: algorithmic bytes per lookup
// = L*(2r+2)^2*s + 8 + L*(2r+1)^2*s.
#include <cstdlib>
#include "common.cuh"

namespace db200 {

constexpr int kLookupTX = 64;      // source pixels (consecutive x) per CTA
constexpr int kLookupWarps = 8;    // 256 threads
constexpr int kMaxLevels = 4;

struct LevelPtrs {
  const void* p[kMaxLevels];
};

// One warp computes the (2r+1)^2 outputs of one source pixel for one level into `sout`
// (sout[ch * kLookupTX + px]).  `win` is this warp's (2r+2)^2 staging buffer.
template <typename T>
__device__ __forceinline__ void lookup_pixel(const T* __restrict__ slice, int h2, int w2, float x0, float y0, int r,
                                             T* __restrict__ win, T* __restrict__ sout, int px, int lane) {
  const int rd = 2 * r + 1, wn = rd + 1;
  const float fx = floorf(x0), fy = floorf(y0);
  const float dx = x0 - fx, dy = y0 - fy;
  const int xb = static_cast<int>(fx) - r, yb = static_cast<int>(fy) - r;
  // window: win[i*wn + j] = volume[y1 = yb + j][x1 = xb + i]  (i = x offset, j = y offset)
  for (int idx = lane; idx < wn * wn; idx += 32) {
    const int j = idx / wn, i = idx - j * wn;   // lanes run along x (contiguous in memory)
    const int x1 = xb + i, y1 = yb + j;
    T v = Elem<T>::zero();
    if (x1 >= 0 && x1 < w2 && y1 >= 0 && y1 < h2) v = slice[y1 * w2 + x1];
    win[i * wn + j] = v;
  }
  __syncwarp();
  const T w00 = Elem<T>::from_float((1.0f - dx) * (1.0f - dy));
  const T w01 = Elem<T>::from_float((1.0f - dx) * dy);
  const T w10 = Elem<T>::from_float(dx * (1.0f - dy));
  const T w11 = Elem<T>::from_float(dx * dy);
  for (int o = lane; o < rd * rd; o += 32) {
    const int i = o / rd, j = o - i * rd;
    // reference accumulation order for output (i,j): samples (i,j), (i,j+1), (i+1,j), (i+1,j+1);
    // out-of-range samples are skipped there, which equals adding a zero product here only when the
    // product is exactly +0 — keep the skip to stay bit-identical for -0/NaN inputs.
    T acc = Elem<T>::zero();
    const int x1 = xb + i, y1 = yb + j;
    const bool xin0 = (x1 >= 0 && x1 < w2), xin1 = (x1 + 1 >= 0 && x1 + 1 < w2);
    const bool yin0 = (y1 >= 0 && y1 < h2), yin1 = (y1 + 1 >= 0 && y1 + 1 < h2);
    if (xin0 && yin0) acc = Elem<T>::mad(win[i * wn + j], w00, acc);
    if (xin0 && yin1) acc = Elem<T>::mad(win[i * wn + j + 1], w01, acc);
    if (xin1 && yin0) acc = Elem<T>::mad(win[(i + 1) * wn + j], w10, acc);
    if (xin1 && yin1) acc = Elem<T>::mad(win[(i + 1) * wn + j + 1], w11, acc);
    sout[o * kLookupTX + px] = acc;
  }
  __syncwarp();
}

// grid: (ceil(w1/TX), h1, N); block 256.  out[n][L*rd*rd][h1][w1]
template <typename T>
__global__ void __launch_bounds__(kLookupWarps * 32)
corr_lookup_fwd_kernel(LevelPtrs levels, const float* __restrict__ coords, T* __restrict__ out, int N, int h1, int w1,
                       int h2, int w2, int num_levels, int r) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int rd = 2 * r + 1, wn = rd + 1, nch = rd * rd;
  T* sout = reinterpret_cast<T*>(smem_raw);                       // [num_levels*nch][TX]
  T* swin = sout + size_t(num_levels) * nch * kLookupTX;          // [warps][wn*wn]
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int x_base = blockIdx.x * kLookupTX, y = blockIdx.y, n = blockIdx.z;
  const int npx = min(kLookupTX, w1 - x_base);
  T* win = swin + warp * wn * wn;
  const size_t plane = size_t(h1) * w1;
  for (int px = warp; px < npx; px += kLookupWarps) {
    const int x = x_base + px;
    const float cx0 = coords[(size_t(n) * 2 + 0) * plane + size_t(y) * w1 + x];
    const float cy0 = coords[(size_t(n) * 2 + 1) * plane + size_t(y) * w1 + x];
    float scale = 1.0f;
    for (int l = 0; l < num_levels; l++) {
      const int h2l = h2 >> l, w2l = w2 >> l;
      const T* slice = static_cast<const T*>(levels.p[l]) + ((size_t(n) * h1 + y) * w1 + x) * size_t(h2l) * w2l;
      lookup_pixel<T>(slice, h2l, w2l, cx0 / scale, cy0 / scale, r, win, sout + size_t(l) * nch * kLookupTX, px, lane);
      scale *= 2.0f;
    }
  }
  __syncthreads();
  // coalesced store: each channel row = npx consecutive x
  const int total = num_levels * nch * kLookupTX;
  for (int idx = threadIdx.x; idx < total; idx += blockDim.x) {
    const int ch = idx / kLookupTX, px = idx - ch * kLookupTX;
    if (px < npx) out[(size_t(n) * num_levels * nch + ch) * plane + size_t(y) * w1 + x_base + px] = sout[idx];
  }
}

// ---- radius-3 fast path (the only radius DROID-SLAM uses: corr.py:23, droid_net.py) ---------------------------
// CTA = PX consecutive source pixels x all levels, 256 threads, four stages separated by barriers:
//   0  thread = (level, pixel): window origin (floor(x)-3, floor(y)-3) and bilinear fractions from the coords;
//   1  warp = (level, pixel) window: the 8x8 samples are fetched with lanes running along the window rows
//      (16-byte rows of the volume slice; out-of-range samples are zero-filled — adding a +0 product is
//      bit-identical to the reference's skipped `within_bounds` branch), four windows in flight per warp;
//   2  thread = (level, pixel): the window comes back from shared memory as eight 16-byte vectors and all
//      49 outputs are formed in registers in the reference's accumulation order;
//   3  the [levels*49][PX] output tile leaves as full coalesced lines.
// Per lookup this is ~130 warp instructions instead of ~1600 (profiles/r1_v1_lookup_ncu.txt), so the kernel
// is bound by the DRAM sectors the 16-byte window rows touch, not by issue slots.
constexpr int kWinStride = 72;     // elements per staged window (64 + 8 pad): conflict-free 16-byte reads at 32 lanes

template <typename T> struct WinVec;            // eight window samples (one x offset, y offsets 0..7) as one aligned vector
template <> struct WinVec<__half> { static constexpr int kLoads = 1; };
template <> struct WinVec<float> { static constexpr int kLoads = 2; };
template <> struct WinVec<double> { static constexpr int kLoads = 4; };

template <typename T>
__device__ __forceinline__ void load_win_col(const T* __restrict__ p, T (&v)[8]) {
  constexpr int n16 = 8 * sizeof(T) / 16;
  const uint4* q = reinterpret_cast<const uint4*>(p);
  uint4 tmp[n16];
#pragma unroll
  for (int i = 0; i < n16; i++) tmp[i] = q[i];
  const T* t = reinterpret_cast<const T*>(tmp);
#pragma unroll
  for (int i = 0; i < 8; i++) v[i] = t[i];
}

template <typename T, int PX>
__global__ void __launch_bounds__(256)
corr_lookup_r3_kernel(LevelPtrs levels, const float* __restrict__ coords, T* __restrict__ out, int N, int h1, int w1,
                      int h2, int w2, int num_levels) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  constexpr int kCh = 49;
  T* win = reinterpret_cast<T*>(smem_raw);                                   // [L*PX][kWinStride]
  T* sout = win + size_t(num_levels) * PX * kWinStride;                      // [L*49][PX]
  int* sxb = reinterpret_cast<int*>(sout + size_t(num_levels) * kCh * PX);   // [L*PX]
  int* syb = sxb + num_levels * PX;
  float* sdx = reinterpret_cast<float*>(syb + num_levels * PX);
  float* sdy = sdx + num_levels * PX;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int x_base = blockIdx.x * PX, y = blockIdx.y, n = blockIdx.z;
  const int npx = min(PX, w1 - x_base);
  const size_t plane = size_t(h1) * w1;
  const int nwin = num_levels * PX;

  // ---- stage 0
  for (int wi = tid; wi < nwin; wi += 256) {
    const int l = wi / PX, px = wi - l * PX;
    int xb = 0, yb = 0;
    float dx = 0.0f, dy = 0.0f;
    if (px < npx) {
      const size_t o = size_t(y) * w1 + x_base + px;
      const float scale = float(1 << l);
      const float x0 = coords[(size_t(n) * 2 + 0) * plane + o] / scale, y0 = coords[(size_t(n) * 2 + 1) * plane + o] / scale;
      const float fx = floorf(x0), fy = floorf(y0);
      dx = x0 - fx; dy = y0 - fy;
      xb = static_cast<int>(fx) - 3; yb = static_cast<int>(fy) - 3;
    }
    sxb[wi] = xb; syb[wi] = yb; sdx[wi] = dx; sdy[wi] = dy;
  }
  __syncthreads();

  // ---- stage 1: gather.  lane -> window element (x offset i = lane & 7, y offsets j = lane >> 3 and +4)
  {
    const int i = lane & 7, j0 = lane >> 3;
    for (int wb = warp * 4; wb < nwin; wb += 32) {           // 4 windows per warp and step
      T v[4][2];
#pragma unroll
      for (int u = 0; u < 4; u++) {
        const int wi = wb + u;
        const int l = wi / PX, px = wi - l * PX;
        v[u][0] = Elem<T>::zero(); v[u][1] = Elem<T>::zero();
        if (wi < nwin && px < npx) {
          const int h2l = h2 >> l, w2l = w2 >> l;
          const T* slice = static_cast<const T*>(levels.p[l]) + ((size_t(n) * h1 + y) * w1 + x_base + px) * size_t(h2l) * w2l;
          const int x1 = sxb[wi] + i, y1 = syb[wi] + j0;
          const bool xin = x1 >= 0 && x1 < w2l;
          if (xin && y1 >= 0 && y1 < h2l) v[u][0] = slice[size_t(y1) * w2l + x1];
          if (xin && y1 + 4 >= 0 && y1 + 4 < h2l) v[u][1] = slice[size_t(y1 + 4) * w2l + x1];
        }
      }
#pragma unroll
      for (int u = 0; u < 4; u++) {
        const int wi = wb + u;
        if (wi < nwin) {
          win[size_t(wi) * kWinStride + i * 8 + j0] = v[u][0];
          win[size_t(wi) * kWinStride + i * 8 + j0 + 4] = v[u][1];
        }
      }
    }
  }
  __syncthreads();

  // ---- stage 2: thread = window; outputs in registers, reference order (i,j), (i,j+1), (i+1,j), (i+1,j+1)
  for (int wi = tid; wi < nwin; wi += 256) {
    const int l = wi / PX, px = wi - l * PX;
    if (px >= npx) continue;
    const float dx = sdx[wi], dy = sdy[wi];
    const T w00 = Elem<T>::from_float((1.0f - dx) * (1.0f - dy));
    const T w01 = Elem<T>::from_float((1.0f - dx) * dy);
    const T w10 = Elem<T>::from_float(dx * (1.0f - dy));
    const T w11 = Elem<T>::from_float(dx * dy);
    const T* wp = win + size_t(wi) * kWinStride;
    T c0[8], c1[8];
    load_win_col<T>(wp, c0);
#pragma unroll
    for (int i = 0; i < 7; i++) {
      load_win_col<T>(wp + (i + 1) * 8, c1);
#pragma unroll
      for (int j = 0; j < 7; j++) {
        T acc = Elem<T>::zero();
        acc = Elem<T>::mad(c0[j], w00, acc);
        acc = Elem<T>::mad(c0[j + 1], w01, acc);
        acc = Elem<T>::mad(c1[j], w10, acc);
        acc = Elem<T>::mad(c1[j + 1], w11, acc);
        sout[size_t(l * kCh + i * 7 + j) * PX + px] = acc;
      }
#pragma unroll
      for (int j = 0; j < 8; j++) c0[j] = c1[j];
    }
  }
  __syncthreads();

  // ---- stage 3: coalesced store, out[n][L*49][h1][w1]
  const int nch = num_levels * kCh;
  T* obase = out + size_t(n) * nch * plane + size_t(y) * w1 + x_base;
  constexpr int kVec = 16 / sizeof(T);
  const bool vec_ok = (npx == PX) && (w1 % kVec == 0) && (reinterpret_cast<uintptr_t>(out) % 16 == 0);
  if (vec_ok) {
    constexpr int per_row = PX / kVec;
    for (int idx = tid; idx < nch * per_row; idx += 256) {
      const int ch = idx / per_row, v = idx - ch * per_row;
      *reinterpret_cast<uint4*>(obase + size_t(ch) * plane + v * kVec) = *reinterpret_cast<const uint4*>(sout + size_t(ch) * PX + v * kVec);
    }
  } else {
    for (int idx = tid; idx < nch * PX; idx += 256) {
      const int ch = idx / PX, px = idx - ch * PX;
      if (px < npx) obase[size_t(ch) * plane + px] = sout[idx];
    }
  }
}

// ---- pipelined radius-3 path (fp16 / fp32 volumes whose level widths are multiples of a 16-byte chunk) ----------
// Persistent CTAs (128 threads, tile = 32 consecutive source pixels x all levels = 128 windows) run a two-stage
// software pipeline so that HBM always has a full tile of window rows in flight per CTA while the previous tile is
// being interpolated and stored:
//   issue(k+1)  every 8-sample window row is fetched as the aligned 16-byte chunks that cover it with
//               cp.async.cg (L2 only); rows/chunks outside the slice are zero-filled by the copy itself (src-size 0),
//               which reproduces the reference's skipped `within_bounds` samples (+0 products);
//   coords(k+2) the coordinates of the tile after next are loaded into registers;
//   compute(k)  thread = window: two 16-byte shared-memory reads per row, a register funnel to the window's element
//               offset, and (fp16) packed __hmul2_rn/__hadd2_rn in the reference's accumulation order — two
//               x-offsets per instruction, each half rounded exactly like the scalar c10::Half sequence;
//   store(k)    the [levels*49][32] tile leaves as 16-byte stores; then the descriptors of tile k+2 are written.
// Window origin/descriptor arithmetic is the same float sequence as the reference (coords / 2^l, floorf).
template <typename T> struct PipeCfg;
template <> struct PipeCfg<__half> { static constexpr int CE = 8, CH = 2, LOG_CE = 3; };
template <> struct PipeCfg<float> { static constexpr int CE = 4, CH = 3, LOG_CE = 2; };
constexpr int kPipePX = 32, kPipeThreads = 128, kPipeWin = kMaxLevels * kPipePX;
template <typename T> __host__ __device__ constexpr int pipe_win_bytes() { return 8 * PipeCfg<T>::CH * 16 + 16; }

struct PipeDesc {                    // one stage of window descriptors
  long long base[kPipeWin];          // element offset (from the level base) of window row 0, chunk 0; may be out of range
  unsigned mask[kPipeWin];           // bits 0-7 valid rows, 8.. valid chunks, 16-19 element offset inside chunk 0
  float dx[kPipeWin], dy[kPipeWin];
};

template <typename T>
__host__ __device__ constexpr size_t pipe_smem_bytes(bool tiled = false) {
  return 2 * size_t(kPipeWin) * pipe_win_bytes<T>() + size_t(kMaxLevels) * 49 * kPipePX * sizeof(T) + 2 * sizeof(PipeDesc) +
         (tiled ? 2 * size_t(kPipeWin) * sizeof(int) : 0);
}


// .L2::64B: ask the L2 for the whole 64-byte DRAM burst around the chunk — the neighbouring chunk / window row that shares the
// burst is requested a few instructions later and then hits (2 % on the row-major lookup, 5 % on the tiled one)
__device__ __forceinline__ void cp_async16_zfill(void* smem_dst, const void* gsrc, bool valid) {
  const unsigned d = static_cast<unsigned>(__cvta_generic_to_shared(smem_dst));
  const int n = valid ? 16 : 0;
  asm volatile("cp.async.cg.shared.global.L2::64B [%0], [%1], 16, %2;\n" ::"r"(d), "l"(gsrc), "r"(n) : "memory");
}

// Row of a staged window -> for fp16: A[p] = samples (x=2p, 2p+1), S[p] = samples (x=2p+1, 2p+2) as half2 bit patterns.
__device__ __forceinline__ void pipe_row_half(const unsigned char* rowp, int q, unsigned sel_a, unsigned sel_s,
                                              unsigned (&A)[4], unsigned (&S)[4]) {
  const uint4 a = *reinterpret_cast<const uint4*>(rowp), b = *reinterpret_cast<const uint4*>(rowp + 16);
  const bool s2 = q & 2, s1 = q & 1;
  const unsigned t0 = s2 ? a.z : a.x, t1 = s2 ? a.w : a.y, t2 = s2 ? b.x : a.z, t3 = s2 ? b.y : a.w, t4 = s2 ? b.z : b.x,
                 t5 = s2 ? b.w : b.y;
  const unsigned u0 = s1 ? t1 : t0, u1 = s1 ? t2 : t1, u2 = s1 ? t3 : t2, u3 = s1 ? t4 : t3, u4 = s1 ? t5 : t4;
  A[0] = __byte_perm(u0, u1, sel_a); A[1] = __byte_perm(u1, u2, sel_a); A[2] = __byte_perm(u2, u3, sel_a); A[3] = __byte_perm(u3, u4, sel_a);
  S[0] = __byte_perm(u0, u1, sel_s); S[1] = __byte_perm(u1, u2, sel_s); S[2] = __byte_perm(u2, u3, sel_s); S[3] = __byte_perm(u3, u4, sel_s);
}

__device__ __forceinline__ __half2 as_h2(unsigned u) { return *reinterpret_cast<__half2*>(&u); }

template <typename T, bool TILED>
__global__ void __launch_bounds__(kPipeThreads)
corr_lookup_r3_pipe_kernel(LevelPtrs levels, const float* __restrict__ coords, T* __restrict__ out, int N, int h1, int w1,
                           int h2, int w2, int num_levels, int num_tiles, unsigned tiled_mask) {
  using Cfg = PipeCfg<T>;
  constexpr int CE = Cfg::CE, CH = Cfg::CH, WINB = pipe_win_bytes<T>(), PX = kPipePX, kCh = 49;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  unsigned char* win = smem_raw;                                                        // [2][kPipeWin][WINB]
  T* sout = reinterpret_cast<T*>(smem_raw + 2 * size_t(kPipeWin) * WINB);               // [L*49][PX]
  PipeDesc* desc = reinterpret_cast<PipeDesc*>(reinterpret_cast<unsigned char*>(sout) + size_t(kMaxLevels) * kCh * PX * sizeof(T));
  int* ykbuf = reinterpret_cast<int*>(desc + 2);                   // TILED: [2][kPipeWin] window origin (row | first chunk << 16)
  const int tid = threadIdx.x;
  const int nwin = num_levels * PX;
  const int tiles_x = (w1 + PX - 1) / PX;
  const size_t plane = size_t(h1) * w1;
  const int my_l = tid / PX, my_px = tid - my_l * PX;              // the window this thread describes and interpolates

  // coordinates of (tile, my_px) -> registers
  auto load_coords = [&](int tile, float& cx, float& cy) {
    cx = 0.0f; cy = 0.0f;
    if (tile < num_tiles && tid < nwin) {
      const int xt = tile % tiles_x, rest = tile / tiles_x, y = rest % h1, n = rest / h1;
      const int x = xt * PX + my_px;
      if (x < w1) {
        const size_t o = size_t(y) * w1 + x;
        cx = __ldg(coords + (size_t(n) * 2 + 0) * plane + o);
        cy = __ldg(coords + (size_t(n) * 2 + 1) * plane + o);
      }
    }
  };
  auto write_desc = [&](int tile, int s, float cx, float cy) {
    if (tile >= num_tiles || tid >= nwin) return;
    const int xt = tile % tiles_x, rest = tile / tiles_x, y = rest % h1, n = rest / h1;
    const int x = xt * PX + my_px;
    const int h2l = h2 >> my_l, w2l = w2 >> my_l;
    const float scale = float(1 << my_l);
    const float x0 = cx / scale, y0 = cy / scale;
    const float fx = floorf(x0), fy = floorf(y0);
    constexpr int kFar = 1 << 20;                                  // anything this far out is all zeros either way
    const int xb = max(-kFar, min(kFar, static_cast<int>(fx))) - 3, yb = max(-kFar, min(kFar, static_cast<int>(fy))) - 3;
    const int k0 = xb >> Cfg::LOG_CE, o = xb & (CE - 1);
    unsigned m = 0;
    if (x < w1) {
#pragma unroll
      for (int r = 0; r < 8; r++) m |= (static_cast<unsigned>(yb + r) < static_cast<unsigned>(h2l)) ? (1u << r) : 0u;
#pragma unroll
      for (int c = 0; c < CH; c++)
        m |= (static_cast<unsigned>(k0 + c) < static_cast<unsigned>(w2l >> Cfg::LOG_CE) && c * CE < o + 8) ? (1u << (8 + c)) : 0u;
      m |= static_cast<unsigned>(o) << 16;
      m |= 1u << 31;                                               // live pixel
    }
    const long long slice = ((static_cast<long long>(n) * h1 + y) * w1 + x) * (static_cast<long long>(h2l) * w2l);
    if (TILED && ((tiled_mask >> my_l) & 1u)) {                    // tile addresses are formed per row/chunk in issue()
      desc[s].base[tid] = slice;
      ykbuf[s * kPipeWin + tid] = (max(-8, min(32767, yb)) & 0xffff) | (max(-8, min(32767, k0)) << 16);
    } else {
      desc[s].base[tid] = slice + static_cast<long long>(yb) * w2l + static_cast<long long>(k0) * CE;
    }
    desc[s].mask[tid] = m;
    desc[s].dx[tid] = x0 - fx;
    desc[s].dy[tid] = y0 - fy;
  };
  auto issue = [&](int tile, int s) {
    if (tile < num_tiles) {
      unsigned char* wbase = win + size_t(s) * kPipeWin * WINB;
#pragma unroll
      for (int l = 0; l < kMaxLevels; l++) {
        if (l < num_levels) {
          const T* lp = static_cast<const T*>(levels.p[l]);
          const int w2l = w2 >> l;
#pragma unroll
          for (int it = 0; it < (PX * 8 * CH) / kPipeThreads; it++) {
            const int idx = it * kPipeThreads + tid;
            const int wl = idx / (8 * CH), rc = idx - wl * (8 * CH), row = rc / CH, c = rc - row * CH;
            const int w = l * PX + wl;
            const unsigned m = desc[s].mask[w];
            const bool ok = ((m >> row) & (m >> (8 + c)) & 1u) != 0;
            const T* src = lp;
            if (TILED && ((tiled_mask >> l) & 1u)) {
              const int yk = ykbuf[s * kPipeWin + w];
              const int yy = static_cast<short>(yk & 0xffff) + row, kc = (yk >> 16) + c;
              if (ok) src = lp + (desc[s].base[w] + ((((yy >> 2) * (w2l >> 3) + kc) << 5) + ((yy & 3) << 3)));
            } else if (ok) {
              src = lp + (desc[s].base[w] + static_cast<long long>(row) * w2l + c * CE);
            }
            cp_async16_zfill(wbase + size_t(w) * WINB + row * (CH * 16) + c * 16, src, ok);
          }
        }
      }
    }
    asm volatile("cp.async.commit_group;\n" ::: "memory");
  };

  int tile = blockIdx.x;
  const int stride = gridDim.x;
  {
    float cx, cy;
    load_coords(tile, cx, cy);
    write_desc(tile, 0, cx, cy);
    load_coords(tile + stride, cx, cy);
    write_desc(tile + stride, 1, cx, cy);
  }
  __syncthreads();
  issue(tile, 0);

  for (int k = 0; tile < num_tiles; k++, tile += stride) {
    const int s = k & 1;
    issue(tile + stride, s ^ 1);                                   // desc[s^1] was written before the last barrier
    float ncx, ncy;
    load_coords(tile + 2 * stride, ncx, ncy);
    asm volatile("cp.async.wait_group 1;\n" ::: "memory");
    __syncthreads();                                               // tile k's windows are in shared memory

    // ---- compute: thread = window
    if (tid < nwin && (desc[s].mask[tid] >> 31)) {
      const unsigned m = desc[s].mask[tid];
      const int o = (m >> 16) & 15;
      const float dx = desc[s].dx[tid], dy = desc[s].dy[tid];
      const unsigned char* wp = win + (size_t(s) * kPipeWin + tid) * WINB;
      T* so = sout + size_t(my_l) * kCh * PX + my_px;
      if constexpr (sizeof(T) == 2) {
        const __half2 w00 = __half2half2(__float2half_rn((1.0f - dx) * (1.0f - dy)));
        const __half2 w01 = __half2half2(__float2half_rn((1.0f - dx) * dy));
        const __half2 w10 = __half2half2(__float2half_rn(dx * (1.0f - dy)));
        const __half2 w11 = __half2half2(__float2half_rn(dx * dy));
        const __half2 zero = __half2half2(__float2half_rn(0.0f));
        const int q = o >> 1;
        const unsigned sel_a = (o & 1) ? 0x5432u : 0x3210u, sel_s = (o & 1) ? 0x7654u : 0x5432u;
        unsigned A0[4], S0[4], A1[4], S1[4];
        pipe_row_half(wp, q, sel_a, sel_s, A0, S0);
#pragma unroll
        for (int j = 0; j < 7; j++) {
          pipe_row_half(wp + (j + 1) * (CH * 16), q, sel_a, sel_s, A1, S1);
#pragma unroll
          for (int p = 0; p < 4; p++) {
            __half2 acc = __hadd2_rn(zero, __hmul2_rn(as_h2(A0[p]), w00));     // x offsets i = 2p (low), 2p+1 (high)
            acc = __hadd2_rn(acc, __hmul2_rn(as_h2(A1[p]), w01));
            acc = __hadd2_rn(acc, __hmul2_rn(as_h2(S0[p]), w10));
            acc = __hadd2_rn(acc, __hmul2_rn(as_h2(S1[p]), w11));
            so[size_t((2 * p) * 7 + j) * PX] = __low2half(acc);
            if (p < 3) so[size_t((2 * p + 1) * 7 + j) * PX] = __high2half(acc);
          }
#pragma unroll
          for (int p = 0; p < 4; p++) { A0[p] = A1[p]; S0[p] = S1[p]; }
        }
      } else {
        const T w00 = Elem<T>::from_float((1.0f - dx) * (1.0f - dy));
        const T w01 = Elem<T>::from_float((1.0f - dx) * dy);
        const T w10 = Elem<T>::from_float(dx * (1.0f - dy));
        const T w11 = Elem<T>::from_float(dx * dy);
        T r0[8], r1[8];
        const T* rp = reinterpret_cast<const T*>(wp) + o;
#pragma unroll
        for (int i = 0; i < 8; i++) r0[i] = rp[i];
#pragma unroll
        for (int j = 0; j < 7; j++) {
          const T* rn = rp + (j + 1) * (CH * 16 / int(sizeof(T)));
#pragma unroll
          for (int i = 0; i < 8; i++) r1[i] = rn[i];
#pragma unroll
          for (int i = 0; i < 7; i++) {
            T acc = Elem<T>::zero();
            acc = Elem<T>::mad(r0[i], w00, acc);
            acc = Elem<T>::mad(r1[i], w01, acc);
            acc = Elem<T>::mad(r0[i + 1], w10, acc);
            acc = Elem<T>::mad(r1[i + 1], w11, acc);
            so[size_t(i * 7 + j) * PX] = acc;
          }
#pragma unroll
          for (int i = 0; i < 8; i++) r0[i] = r1[i];
        }
      }
    }
    __syncthreads();                                               // sout complete; win[s], desc[s] free

    // ---- store tile k, then describe tile k+2 into desc[s]
    {
      const int xt = tile % tiles_x, rest = tile / tiles_x, y = rest % h1, n = rest / h1;
      const int x_base = xt * PX, npx = min(PX, w1 - x_base);
      const int nch = num_levels * kCh;
      T* obase = out + size_t(n) * nch * plane + size_t(y) * w1 + x_base;
      constexpr int kVec = 16 / sizeof(T), per_row = PX / kVec;
      const bool vec_ok = (npx == PX) && (w1 % kVec == 0) && (reinterpret_cast<uintptr_t>(out) % 16 == 0);
      if (vec_ok) {
        for (int idx = tid; idx < nch * per_row; idx += kPipeThreads) {
          const int ch = idx / per_row, v = idx - ch * per_row;
          *reinterpret_cast<uint4*>(obase + size_t(ch) * plane + v * kVec) = *reinterpret_cast<const uint4*>(sout + size_t(ch) * PX + v * kVec);
        }
      } else {
        for (int idx = tid; idx < nch * PX; idx += kPipeThreads) {
          const int ch = idx / PX, px = idx - ch * PX;
          if (px < npx) obase[size_t(ch) * plane + px] = sout[idx];
        }
      }
    }
    write_desc(tile + 2 * stride, s, ncx, ncy);
    __syncthreads();
  }
  asm volatile("cp.async.wait_group 0;\n" ::: "memory");
}

// ---- direct radius-3 fp16 path: thread = window, the window lives in registers -----------------------------------------------
// No shared memory, no barriers, no staging pass: every thread loads the 8 rows x 2 aligned 16-byte chunks that cover its window
// straight into registers (16 independent 128-bit loads in flight per thread; rows/chunks outside the slice are never read and
// stay zero = the reference's skipped `within_bounds` samples), interpolates with the same packed half2 sequence as the
// pipelined kernel (bit-identical), and writes its 49 values itself: the 32 lanes of a warp are 32 consecutive source pixels
// of ONE level, so every channel store of the warp is one contiguous 64-byte run.  With 80 registers per thread 6 CTAs are
// resident per SM (24 warps) — against 8 warps for the staged kernel, which holds two tiles of windows in shared memory — and a
// lookup costs 33 M warp instructions per 96 edges instead of 57 M (no descriptor round trip, no cp.async address pass, no
// shared-memory transposition).  That matters once the tiled slice layout (common.cuh) has cut the DRAM bytes: the staged
// kernel then runs at 42 % DRAM utilisation, bound by its instruction stream (0.122 ms; ncu, tools/tiled_ncu.py), this one at
// 0.093 ms and back on the DRAM pins (377 MB read + 116 MB written per launch at 5.3 TB/s).  On row-major levels the staged
// kernel stays the default (0.108 against 0.112 ms: there the bytes, not the instructions, are the bound).
// 16 bytes from global memory (read-only path, no L1 allocation) if row yy lies inside the level and the chunk is valid, else
// zeros.  The predicate is formed right at the load from plain integers: with 16 such loads in flight per thread the compiler
// otherwise keeps 16 predicates alive at once and spills them through general registers.
__device__ __forceinline__ uint4 ldg16_if(const void* p, unsigned yy, unsigned h2l, int chunk_ok) {
  uint4 v;
  asm("{\n\t.reg .pred p, q;\n\t"
      "setp.ne.b32 q, %7, 0;\n\t"
      "setp.lt.and.u32 p, %5, %6, q;\n\t"
      "mov.b32 %0, 0;\n\tmov.b32 %1, 0;\n\tmov.b32 %2, 0;\n\tmov.b32 %3, 0;\n\t"
      "@p ld.global.nc.L1::no_allocate.L2::64B.v4.u32 {%0,%1,%2,%3}, [%4];\n\t}"
      : "=&r"(v.x), "=&r"(v.y), "=&r"(v.z), "=&r"(v.w) : "l"(p), "r"(yy), "r"(h2l), "r"(chunk_ok));
  return v;
}

__device__ __forceinline__ void direct_row_half(const uint4& a, const uint4& b, int q, unsigned sel_a, unsigned sel_s,
                                                unsigned (&A)[4], unsigned (&S)[4]) {
  const bool s2 = q & 2, s1 = q & 1;
  const unsigned t0 = s2 ? a.z : a.x, t1 = s2 ? a.w : a.y, t2 = s2 ? b.x : a.z, t3 = s2 ? b.y : a.w, t4 = s2 ? b.z : b.x,
                 t5 = s2 ? b.w : b.y;
  const unsigned u0 = s1 ? t1 : t0, u1 = s1 ? t2 : t1, u2 = s1 ? t3 : t2, u3 = s1 ? t4 : t3, u4 = s1 ? t5 : t4;
  A[0] = __byte_perm(u0, u1, sel_a); A[1] = __byte_perm(u1, u2, sel_a); A[2] = __byte_perm(u2, u3, sel_a); A[3] = __byte_perm(u3, u4, sel_a);
  S[0] = __byte_perm(u0, u1, sel_s); S[1] = __byte_perm(u1, u2, sel_s); S[2] = __byte_perm(u2, u3, sel_s); S[3] = __byte_perm(u3, u4, sel_s);
}

constexpr int kDirectThreads = 128;
constexpr unsigned kDirectPrefetch = 12 * kNumSMs;   // tiles ahead whose coordinates are prefetched to L2 (two waves of CTAs)
#ifndef DIRECT_CTAS
#define DIRECT_CTAS 6                    // resident CTAs per SM the register budget is set for (80 registers)
#endif

// block = 32 * num_levels threads: warp = level, lane = source pixel; one CTA per source tile (32 pixels of one row), tiles in
// memory order.  Launched one-shot on purpose: CTAs that are resident together then work on neighbouring tiles, whose output
// rows and coordinates share DRAM pages — a persistent grid-stride walk of the same kernel (6 CTAs per SM) measured 0.105 ms
// against 0.093 ms one-shot for 96 edges (the loop stays for grids clipped to 2^31 - 1 CTAs)
template <bool TILED>
__global__ void __launch_bounds__(kDirectThreads, DIRECT_CTAS)
corr_lookup_r3_direct_kernel(LevelPtrs levels, const float* __restrict__ coords, __half* __restrict__ out, int num_tiles, int h1,
                             int w1, int h2, int w2, int num_levels, unsigned tiled_mask) {
  const int lane = threadIdx.x & 31, l = threadIdx.x >> 5;
  const unsigned tiles_x = unsigned(w1 + 31) >> 5;
  const size_t plane = size_t(h1) * w1;
 for (unsigned tile = blockIdx.x; tile < unsigned(num_tiles); tile += gridDim.x) {
  // the coordinates of a tile that starts a couple of CTA lifetimes from now -> L2: its first, dependent load then hits (1.5 %)
  if (threadIdx.x < 2) {
    const unsigned t2 = tile + kDirectPrefetch;
    if (t2 < unsigned(num_tiles)) {
      const unsigned r2 = t2 / tiles_x;
      const unsigned n2 = r2 / unsigned(h1), y2 = r2 - n2 * unsigned(h1), x2 = (t2 - r2 * tiles_x) * 32u;
      const float* pf = coords + (size_t(n2) * 2 + threadIdx.x) * plane + size_t(y2) * w1 + x2;
      asm volatile("prefetch.global.L2 [%0];" ::"l"(pf));
    }
  }
  const unsigned rest = tile / tiles_x;
  const int n = int(rest / unsigned(h1)), y = int(rest - unsigned(n) * unsigned(h1));
  const int x = int(tile - rest * tiles_x) * 32 + lane;
  if (x >= w1) continue;                                           // no warp-wide operation below
  const size_t o = size_t(y) * w1 + x;
  const float cx = __ldg(coords + (size_t(n) * 2 + 0) * plane + o), cy = __ldg(coords + (size_t(n) * 2 + 1) * plane + o);
  const int h2l = h2 >> l, w2l = w2 >> l;
  // coords / 2^l (corr.py:46): the quotient by a power of two is the exact product with its reciprocal (both are the correctly
  // rounded value of the same real number), so this is the reference's float sequence bit for bit
  const float inv = __int_as_float((127 - l) << 23);
  const float x0 = cx * inv, y0 = cy * inv;
  const float fx = floorf(x0), fy = floorf(y0);
  constexpr int kFar = 1 << 20;                                    // anything this far out is all zeros either way
  const int xb = max(-kFar, min(kFar, static_cast<int>(fx))) - 3, yb = max(-kFar, min(kFar, static_cast<int>(fy))) - 3;
  const int k0 = xb >> 3, off = xb & 7;
  const int nck = w2l >> 3;
  const int c0ok = static_cast<unsigned>(k0) < static_cast<unsigned>(nck);
  const int c1ok = static_cast<unsigned>(k0 + 1) < static_cast<unsigned>(nck) && off > 0;
  const void* lp = l == 0 ? levels.p[0] : l == 1 ? levels.p[1] : l == 2 ? levels.p[2] : levels.p[3];   // no local copy of the array
  const __half* sl = static_cast<const __half*>(lp) + ((static_cast<long long>(n) * h1 + y) * w1 + x) * (static_cast<long long>(h2l) * w2l);
  const bool tl = TILED && ((tiled_mask >> l) & 1u);
  // element offset of (row yy, chunk k) inside the slice: row-major yy * w2l + 8 k; tiled 8 yy + (yy >> 2) * (32 nck - 32) + 32 k
  const int rs = tl ? 8 : w2l, ts = tl ? 32 * nck - 32 : 0, cs = tl ? 32 : 8;
  const __half* b0 = sl + static_cast<long long>(k0) * cs;
  const __half* b1 = b0 + cs;
  uint4 ra[8], rb[8];
#pragma unroll
  for (int r = 0; r < 8; r++) {
    const int yy = yb + r;
    const int roff = yy * rs + (yy >> 2) * ts;                     // never dereferenced when yy is outside the level
    ra[r] = ldg16_if(b0 + roff, unsigned(yy), unsigned(h2l), c0ok);
    rb[r] = ldg16_if(b1 + roff, unsigned(yy), unsigned(h2l), c1ok);
  }
  const float dx = x0 - fx, dy = y0 - fy;
  const __half2 w00 = __half2half2(__float2half_rn((1.0f - dx) * (1.0f - dy)));
  const __half2 w01 = __half2half2(__float2half_rn((1.0f - dx) * dy));
  const __half2 w10 = __half2half2(__float2half_rn(dx * (1.0f - dy)));
  const __half2 w11 = __half2half2(__float2half_rn(dx * dy));
  const __half2 zero = __half2half2(__float2half_rn(0.0f));
  const int q = off >> 1;
  const unsigned sel_a = (off & 1) ? 0x5432u : 0x3210u, sel_s = (off & 1) ? 0x7654u : 0x5432u;
  __half* ob = out + (size_t(n) * num_levels + l) * 49 * plane + o;        // channel (l*49 + i*7 + j) of this pixel: ob[(i*7 + j) * plane]
  unsigned A0[4], S0[4], A1[4], S1[4];
  direct_row_half(ra[0], rb[0], q, sel_a, sel_s, A0, S0);
#pragma unroll
  for (int j = 0; j < 7; j++) {
    direct_row_half(ra[j + 1], rb[j + 1], q, sel_a, sel_s, A1, S1);
#pragma unroll
    for (int p = 0; p < 4; p++) {
      __half2 acc = __hadd2_rn(zero, __hmul2_rn(as_h2(A0[p]), w00));       // x offsets i = 2p (low), 2p+1 (high)
      acc = __hadd2_rn(acc, __hmul2_rn(as_h2(A1[p]), w01));
      acc = __hadd2_rn(acc, __hmul2_rn(as_h2(S0[p]), w10));
      acc = __hadd2_rn(acc, __hmul2_rn(as_h2(S1[p]), w11));
      ob[size_t((2 * p) * 7 + j) * plane] = __low2half(acc);
      if (p < 3) ob[size_t((2 * p + 1) * 7 + j) * plane] = __high2half(acc);
    }
#pragma unroll
    for (int p = 0; p < 4; p++) { A0[p] = A1[p]; S0[p] = S1[p]; }
  }
 }
}

template <bool TILED>
int launch_direct(const LevelPtrs& lv, const float* coords, void* out, int N, int h1, int w1, int h2, int w2, int L, unsigned tiled_mask,
                  cudaStream_t st) {
  const long long tiles = static_cast<long long>(N) * h1 * cdiv(w1, 32);
  if (tiles > 0x7fffffffLL) return DB200_ERR_ARG;
  static const int per_sm = getenv("DB200_LOOKUP_CTAS") ? atoi(getenv("DB200_LOOKUP_CTAS")) : 0;   // > 0: persistent walk (experiments)
  const long long cap = per_sm > 0 ? static_cast<long long>(per_sm) * kNumSMs : tiles;
  const int grid = int(tiles < cap ? tiles : cap);
  corr_lookup_r3_direct_kernel<TILED><<<grid, 32 * L, 0, st>>>(lv, coords, static_cast<__half*>(out), int(tiles), h1, w1, h2, w2, L,
                                                               tiled_mask);
  DB200_CHECK_LAUNCH();
  return DB200_OK;
}

// Backward: one warp per source pixel writes the WHOLE h2*w2 slice of volume_grad (zeros outside the
// window), so no separate memset pass is needed.  Gradient value per window sample follows
// correlation_kernels.cu:108-121 (g accumulates the 4 contributions in that order, then 0 + g).
template <typename T>
__global__ void __launch_bounds__(256)
corr_lookup_bwd_kernel(const float* __restrict__ coords, const T* __restrict__ corr_grad, T* __restrict__ volume_grad,
                       int N, int h1, int w1, int h2, int w2, int r) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int rd = 2 * r + 1, nch = rd * rd;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  T* sg = reinterpret_cast<T*>(smem_raw) + warp * nch;  // this pixel's rd*rd incoming grads
  const size_t plane = size_t(h1) * w1;
  const size_t pix = size_t(blockIdx.x) * (blockDim.x >> 5) + warp;
  if (pix >= size_t(N) * plane) return;
  const int n = int(pix / plane);
  const size_t yx = pix - size_t(n) * plane;
  const float x0 = coords[(size_t(n) * 2 + 0) * plane + yx];
  const float y0 = coords[(size_t(n) * 2 + 1) * plane + yx];
  const float fx = floorf(x0), fy = floorf(y0);
  const float dx = x0 - fx, dy = y0 - fy;
  const int xb = static_cast<int>(fx) - r, yb = static_cast<int>(fy) - r;
  for (int o = lane; o < nch; o += 32) sg[o] = corr_grad[(size_t(n) * nch + o) * plane + yx];
  __syncwarp();
  const T w11 = Elem<T>::from_float(dx * dy), w10 = Elem<T>::from_float(dx * (1.0f - dy));
  const T w01 = Elem<T>::from_float((1.0f - dx) * dy), w00 = Elem<T>::from_float((1.0f - dx) * (1.0f - dy));
  T* slice = volume_grad + pix * size_t(h2) * w2;
  auto grad_at = [&](int x1, int y1) -> T {
    const int i = x1 - xb, j = y1 - yb;
    T g = Elem<T>::zero();
    if (i >= 0 && i <= rd && j >= 0 && j <= rd) {
      if (i > 0 && j > 0) g = Elem<T>::mad(sg[(i - 1) * rd + (j - 1)], w11, g);
      if (i > 0 && j < rd) g = Elem<T>::mad(sg[(i - 1) * rd + j], w10, g);
      if (i < rd && j > 0) g = Elem<T>::mad(sg[i * rd + (j - 1)], w01, g);
      if (i < rd && j < rd) g = Elem<T>::mad(sg[i * rd + j], w00, g);
    }
    return g;
  };
  // 16 bytes per lane and step: almost every vector of the slice lies outside the window and is a plain zero store (the
  // element-wise loop spent ~25 instructions per 4-byte store: 865 M warp instructions for the 4 levels of 88 fp32 edges)
  constexpr int VEC = 16 / int(sizeof(T));
  if (w2 % VEC == 0 && reinterpret_cast<uintptr_t>(volume_grad) % 16 == 0) {
    const int vpr = w2 / VEC, nvec = h2 * vpr;                       // vectors per row / per slice (slices stay 16-byte aligned)
    const int sh = (vpr & (vpr - 1)) == 0 ? __ffs(vpr) - 1 : -1;
    for (int v = lane; v < nvec; v += 32) {
      const int y1 = sh >= 0 ? (v >> sh) : (v / vpr);
      const int x1 = (v - y1 * vpr) * VEC;
      union { uint4 q; T t[VEC]; } o;
      o.q = make_uint4(0u, 0u, 0u, 0u);                              // +0 in every format
      const int j = y1 - yb;
      if (j >= 0 && j <= rd && x1 + VEC > xb && x1 <= xb + rd) {
#pragma unroll
        for (int u = 0; u < VEC; u++) o.t[u] = grad_at(x1 + u, y1);
      }
      reinterpret_cast<uint4*>(slice)[v] = o.q;
    }
    return;
  }
  for (int e = lane; e < h2 * w2; e += 32) {
    const int y1 = e / w2, x1 = e - y1 * w2;
    slice[e] = grad_at(x1, y1);
  }
}

template <typename T>
int launch_fwd(const LevelPtrs& lv, const float* coords, void* out, int N, int h1, int w1, int h2, int w2, int L, int r,
               cudaStream_t st, unsigned tiled_mask = 0) {
  if (tiled_mask != 0) {
    // tiled levels exist for the pipelined fp16 radius-3 kernel only (CorrBlock's private pyramid); no other path reads them
    if constexpr (sizeof(T) == 2) {
      bool ok = (r == 3) && (tiled_mask >> L) == 0;
      for (int l = 0; l < L && ok; l++) {
        ok = ((w2 >> l) % 8 == 0) && ((h2 >> l) > 0) && (reinterpret_cast<uintptr_t>(lv.p[l]) % 16 == 0);
        if (ok && ((tiled_mask >> l) & 1u)) ok = level_is_tiled(h2 >> l, w2 >> l) && (h2 >> l) < 32768 && (w2 >> l) < 32768 * 8;
      }
      const long long tiles = static_cast<long long>(N) * h1 * cdiv(w1, kPipePX);
      if (!ok || tiles > 0x7fffffff) return DB200_ERR_ARG;
      static const bool staged = getenv("DB200_LOOKUP_STAGED") != nullptr;     // the cp.async-staged kernel (comparison runs)
      if (!staged) return launch_direct<true>(lv, coords, out, N, h1, w1, h2, w2, L, tiled_mask, st);
      const size_t smem = pipe_smem_bytes<T>(true);
      auto kern = corr_lookup_r3_pipe_kernel<T, true>;
      DB200_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
      const int grid = int(tiles < 2LL * kNumSMs ? tiles : 2LL * kNumSMs);
      kern<<<grid, kPipeThreads, smem, st>>>(lv, coords, static_cast<T*>(out), N, h1, w1, h2, w2, L, int(tiles), tiled_mask);
      DB200_CHECK_LAUNCH();
      return DB200_OK;
    }
    return DB200_ERR_ARG;
  }
  if constexpr (sizeof(T) <= 4) {
    // pipelined path: every level's width must be whole 16-byte chunks and the level bases 16-byte aligned
    constexpr int CE = 16 / int(sizeof(T));
    bool ok = (r == 3) && !getenv("DB200_LOOKUP_NO_PIPE");
    for (int l = 0; l < L && ok; l++)
      ok = ((w2 >> l) % CE == 0) && ((h2 >> l) > 0) && (reinterpret_cast<uintptr_t>(lv.p[l]) % 16 == 0);
    if (ok) {
      const long long tiles = static_cast<long long>(N) * h1 * cdiv(w1, kPipePX);
      if (tiles > 0x7fffffff) return DB200_ERR_ARG;
      if constexpr (sizeof(T) == 2) {
        static const bool direct = getenv("DB200_LOOKUP_DIRECT") != nullptr;   // register-resident windows on the row-major layout
        if (direct) return launch_direct<false>(lv, coords, out, N, h1, w1, h2, w2, L, 0u, st);
      }
      const size_t smem = pipe_smem_bytes<T>();
      auto kern = corr_lookup_r3_pipe_kernel<T, false>;
      DB200_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
      const int per_sm = smem <= 110 * 1024 ? 2 : 1;
      const int grid = int(tiles < static_cast<long long>(per_sm) * kNumSMs ? tiles : static_cast<long long>(per_sm) * kNumSMs);
      kern<<<grid, kPipeThreads, smem, st>>>(lv, coords, static_cast<T*>(out), N, h1, w1, h2, w2, L, int(tiles), 0u);
      DB200_CHECK_LAUNCH();
      return DB200_OK;
    }
  }
  if (r == 3) {
    constexpr int PX = sizeof(T) == 8 ? 32 : 64;
    const size_t smem = (size_t(L) * PX * kWinStride + size_t(L) * 49 * PX) * sizeof(T) + size_t(L) * PX * 16;
    auto kern = corr_lookup_r3_kernel<T, PX>;
    if (smem > 48 * 1024) DB200_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
    dim3 grid(cdiv(w1, PX), h1, N);
    kern<<<grid, 256, smem, st>>>(lv, coords, static_cast<T*>(out), N, h1, w1, h2, w2, L);
    DB200_CHECK_LAUNCH();
    return DB200_OK;
  }
  const int rd = 2 * r + 1, wn = rd + 1;
  const size_t smem = (size_t(L) * rd * rd * kLookupTX + size_t(kLookupWarps) * wn * wn) * sizeof(T);
  if (smem > 200 * 1024) return DB200_ERR_ARG;
  auto kern = corr_lookup_fwd_kernel<T>;
  if (smem > 48 * 1024) DB200_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
  dim3 grid(cdiv(w1, kLookupTX), h1, N);
  kern<<<grid, kLookupWarps * 32, smem, st>>>(lv, coords, static_cast<T*>(out), N, h1, w1, h2, w2, L, r);
  DB200_CHECK_LAUNCH();
  return DB200_OK;
}

template <typename T>
int launch_bwd(const float* coords, const void* g, void* vg, int N, int h1, int w1, int h2, int w2, int r, cudaStream_t st) {
  const int rd = 2 * r + 1;
  const size_t smem = size_t(8) * rd * rd * sizeof(T);
  const size_t npix = size_t(N) * h1 * w1;
  corr_lookup_bwd_kernel<T><<<unsigned((npix + 7) / 8), 256, smem, st>>>(coords, static_cast<const T*>(g),
                                                                        static_cast<T*>(vg), N, h1, w1, h2, w2, r);
  DB200_CHECK_LAUNCH();
  return DB200_OK;
}

}  // namespace db200

using namespace db200;

static int pyramid_lookup(const void* const* levels, const float* coords, void* out, int N, int h1, int w1, int h2, int w2,
                          int num_levels, int radius, int dtype, unsigned tiled_mask, db200_stream_t stream) {
  if (!levels || !coords || !out || N < 0 || h1 <= 0 || w1 <= 0 || h2 <= 0 || w2 <= 0 || radius < 0 || radius > 7 ||
      num_levels < 1 || num_levels > kMaxLevels)
    return DB200_ERR_ARG;
  if (N == 0) return DB200_OK;
  LevelPtrs lv{};
  for (int l = 0; l < num_levels; l++) {
    if (!levels[l]) return DB200_ERR_ARG;
    lv.p[l] = levels[l];
  }
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  switch (dtype) {
    case DB200_F32: return launch_fwd<float>(lv, coords, out, N, h1, w1, h2, w2, num_levels, radius, st, tiled_mask);
    case DB200_F16: return launch_fwd<__half>(lv, coords, out, N, h1, w1, h2, w2, num_levels, radius, st, tiled_mask);
    case DB200_F64: return launch_fwd<double>(lv, coords, out, N, h1, w1, h2, w2, num_levels, radius, st, tiled_mask);
  }
  return DB200_ERR_ARG;
}

extern "C" int db200_corr_pyramid_lookup(const void* const* levels, const float* coords, void* out, int N, int h1, int w1,
                                         int h2, int w2, int num_levels, int radius, int dtype, db200_stream_t stream) {
  return pyramid_lookup(levels, coords, out, N, h1, w1, h2, w2, num_levels, radius, dtype, 0u, stream);
}

extern "C" int db200_corr_pyramid_lookup_tiled(const void* const* levels, const float* coords, void* out, int N, int h1, int w1,
                                               int h2, int w2, int num_levels, int radius, int dtype, unsigned tiled_mask,
                                               db200_stream_t stream) {
  return pyramid_lookup(levels, coords, out, N, h1, w1, h2, w2, num_levels, radius, dtype, tiled_mask, stream);
}

extern "C" int db200_corr_index_forward(const void* volume, const float* coords, void* corr, int N, int h1, int w1,
                                        int h2, int w2, int radius, int dtype, db200_stream_t stream) {
  const void* lv[1] = {volume};
  return db200_corr_pyramid_lookup(lv, coords, corr, N, h1, w1, h2, w2, 1, radius, dtype, stream);
}

extern "C" int db200_corr_index_backward(const float* coords, const void* corr_grad, void* volume_grad, int N, int h1,
                                         int w1, int h2, int w2, int radius, int dtype, db200_stream_t stream) {
  if (!coords || !corr_grad || !volume_grad || N < 0 || h1 <= 0 || w1 <= 0 || h2 <= 0 || w2 <= 0 || radius < 0 || radius > 7)
    return DB200_ERR_ARG;
  if (N == 0) return DB200_OK;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  switch (dtype) {
    case DB200_F32: return launch_bwd<float>(coords, corr_grad, volume_grad, N, h1, w1, h2, w2, radius, st);
    case DB200_F16: return launch_bwd<__half>(coords, corr_grad, volume_grad, N, h1, w1, h2, w2, radius, st);
    case DB200_F64: return launch_bwd<double>(coords, corr_grad, volume_grad, N, h1, w1, h2, w2, radius, st);
  }
  return DB200_ERR_ARG;
}
