// altcorr_tc.cu — low-memory correlation (AltCorrBlock) on the 5th-generation tensor cores.
// Replaces /root/reference/src/altcorr_kernel.cu:27-149 (altcorr_forward) and the per-level index_select + .float()
// copies of AltCorrBlock.corr_fn (/root/reference/droid_slam/modules/corr.py:109-131) for C = 128, radius = 3.
//
// The reference computes, per source pixel, 64 dot products of 128 channels against an 8x8 window of fmap2 (one warp per
// 4x8 pixel tile, a gather + __syncthreads per tap and channel slab).  Neighbouring pixels reproject to neighbouring
// windows, so for a TILE of 8x16 = 128 source pixels the windows cover a small region of fmap2; the dot products of the
// whole tile against a 16x8 = 128-position CELL of that region are one 128x128x128 GEMM:
//
//     S[p, n] = sum_c f1[tile pixel p, c] * f2[cell position n, c]            (tcgen05.mma, fp32 accumulator in TMEM)
//
// and every pixel then picks its window taps out of its row of S.  Which cells a tile touches is data dependent (the
// bounding corner of the tile's windows + a 64-bit mask of 8x8 cells relative to it); arbitrary coordinates just touch
// more cells, so there is no "fast path only for smooth flow" — results are the reference's for any input.
//
// altcorr_plan_kernel (one warp per tile) writes the per-level cell origin + mask of every tile first; then one persistent
// CTA per SM, 19 warps:
//   warp 0      producer: reads the tile's plan entry (the next one is prefetched), publishes it to the other roles, then
//               feeds the A tile (128 pixels x 128 channels) and a ring of B cells, each as ONE 5-D TMA box
//               [64 ch x 16 x 8 x 2 channel blocks x 1 frame] of the channel-blocked fp16 pyramid, 128B swizzle (the TMA
//               zero-fills positions outside the image = the reference's `within_bounds` skip);
//   warps 1, 18 two single-thread MMA issuers (even / odd cells): 8 (x3 when operands carry lo parts) tcgen05.mma
//               M=128,N=128,K=16 per cell into one of FOUR 128-column TMEM accumulators; tcgen05.commit hands stages back /
//               accumulators on;
//   warps 2..17 epilogue (the bottleneck: issue bound, the tensor pipe is 30 % busy): four warps per TMEM lane quarter, each
//               reads ONE 32-column chunk (2 cell rows) with tcgen05.ld — skipped when no lane has a window row there — and
//               scatters the taps that belong to its lane's pixel into the tap buffer D[64 taps][128 pixels] (conflict-free,
//               one base address, predicates from a column bit mask); per level the four warps of a pixel meet at a
//               128-thread named barrier, combine bilinearly (out-of-image taps are selected to zero by per-thread
//               row/column masks: D is never cleared) and store 49 coalesced channel rows.
// All 4 pyramid levels of a tile are done by the same CTA (the A tile is loaded once per tile).
//
// Precision: operands are fp16.  The fused AltCorrBlock path feeds the fp16 feature pyramid directly (the reference casts
// the same fp16 values to fp32: products are exact either way).  fp32 callers (droid_backends.altcorr_forward) go through
// a split kernel x = hi + lo (both fp16, |x - hi - lo| <= 2^-22 |x|); the GEMM then adds hi*lo + lo*hi to hi*hi
// (3 products, skipped at run time when every lo is zero, i.e. fp16-exact inputs).  Values outside the fp16 range or
// non-finite set a flag that routes the whole call to the SIMT kernel below, which is also the path for levels whose
// cells do not fit the 8x8 mask (images wider than 128 / taller than 64 positions with scattered coordinates).
#include <cstdlib>
#include "tc_common.cuh"

namespace db200 {

constexpr int kATPy = 8, kATPx = 16;                   // pixel tile
constexpr int kATCy = 8, kATCx = 16;                   // fmap2 cell
constexpr int kATMaxL = 4;
constexpr int kATEpiWarps = 16;                      // 4 per TMEM lane quarter: one 32-column chunk (2 cell rows) each
constexpr int kATMmaWarpB = 2 + kATEpiWarps;            // warp 18: the second MMA issuer (odd cells); warp 1 takes the even ones
constexpr int kATThreads = 64 + 32 * kATEpiWarps + 32;  // 608
constexpr uint32_t kATBlk = 128 * 64 * 2;              // one [128 rows][64 halfs] swizzled block: 16 KB
constexpr uint32_t kATARegion = 4 * kATBlk;            // 64 KB: one hi+lo A tile, or two hi-only A tiles
constexpr uint32_t kATBRegion = 8 * kATBlk;            // 128 KB: two hi+lo cells, or four hi-only cells
constexpr uint32_t kATDBytes = 64 * 128 * 4;           // 32 KB tap buffer
constexpr size_t kATSmem = kATARegion + kATBRegion + kATDBytes + 1024;
constexpr uint32_t kATIdesc = (1u << 4) | (uint32_t(128 >> 3) << 17) | (uint32_t(128 >> 4) << 24);   // f16 x f16 -> f32, M = N = 128

struct AltTaskInfo {                                   // 80 bytes: the plan entry of one pixel tile
  int bx0[kATMaxL], by0[kATMaxL];                      // bounding corner of the tile's windows in the target map, per level
  unsigned long long mask[kATMaxL];                    // 8x8 mask of 16x8 cells relative to it
  int skip[kATMaxL];                                   // cells beyond the mask: this (tile, level) is on the SIMT list
};
static_assert(sizeof(AltTaskInfo) == 80, "plan entry size");

struct AltTcParams {
  CUtensorMap a_hi, a_lo;
  CUtensorMap b_hi[kATMaxL], b_lo[kATMaxL];
  const float* coords;
  long long c_sb, c_ss, c_sp;                          // coords strides (floats): edge, sample, pixel; (x,y) contiguous
  float* out;
  long long o_sb, o_ss, o_sc, o_sp;                    // output strides (floats): edge, sample, channel, pixel
  const long long* idx1;                               // frame of fmap1 per edge (null: edge index)
  const long long* idx2;                               // frame of fmap2 per edge (null: edge index)
  const unsigned* flags;                               // bit 0: lo parts present and non-zero; bit 1: whole call on the SIMT path
  int* status;                                         // [0] error word, [1] number of (task, level) entries for the SIMT kernel,
                                                       // [2] cells (tile x cell GEMMs) issued — instrumentation
  int* list;                                           // entries: task * kATMaxL + level
  int list_cap;
  AltTaskInfo* plan;                                   // [tasks], written by altcorr_plan_kernel
  int B, S, H1, W1, L;
  int H2[kATMaxL], W2[kATMaxL];
  float scale[kATMaxL];                                // level coordinates = coords * scale (2^-l: exact, = coords / 2**l)
  int tiles_x, tiles_y;
  int has_lo;
  int dbg;                                             // timing experiments only (env DB200_ALT_DBG): 1 no tap scatter, 2 no output
                                                       // stores, 4 no TMEM reads, 8 no per-level barriers; results are then wrong
};


// one box = both 64-channel blocks of a 16 x 8 position tile of frame f: [2][8][16][64] halfs = 32 KB, the two K-blocks of the operand
__device__ __forceinline__ void tma_load_tile(uint32_t dst, const CUtensorMap* map, int x, int y, int f, uint32_t bar) {
  asm volatile("cp.async.bulk.tensor.5d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4, %5, %6}], [%7];"
               ::"r"(dst), "l"(map), "r"(0), "r"(x), "r"(y), "r"(0), "r"(f), "r"(bar) : "memory");
}

// window origin of a level coordinate: floor, as int, clamped so that +-7 never overflows (far-away windows are out of
// every image anyway); NaN -> 0 like cvt.rzi
__device__ __forceinline__ int alt_floor_clamped(float v, float* frac) {
  const float f = floorf(v);
  *frac = v - f;
  const float c = fminf(fmaxf(f, -1048576.0f), 1048576.0f);
  return (f == f) ? static_cast<int>(c) : 0;
}

// the four epilogue warps that share a TMEM lane quarter q (one per 32-column chunk of the accumulator) are the only threads
// that touch the tap columns of pixels 32q..32q+31: they synchronise among themselves (named barrier 1 + q, 128 threads)
__device__ __forceinline__ void alt_pair_bar(int q) { asm volatile("bar.sync %0, 128;" ::"r"(1 + q) : "memory"); }

// ---- plan: which target cells every pixel tile needs, per level (one warp per tile; thousands of warps in flight) --------
// In the first version the persistent kernel's producer warp did this itself, tile after tile: ~2000 dependent instructions of
// ONE warp per tile (~6 us) in front of that tile's loads — with the whole epilogue switched off the kernel still took 82 % of
// its time.  Computed up front it costs a few microseconds for the whole call.
__global__ void __launch_bounds__(256) altcorr_plan_kernel(const __grid_constant__ AltTcParams P) {
  const unsigned flags = P.flags ? *P.flags : 0u;
  if (flags & 2u) return;
  const int lane = threadIdx.x & 31;
  const int tiles = P.tiles_x * P.tiles_y;
  const long long total = static_cast<long long>(P.B) * P.S * tiles;
  const long long t = static_cast<long long>(blockIdx.x) * 8 + (threadIdx.x >> 5);
  if (t >= total) return;
  const int tile = int(t % tiles);
  const long long bs = t / tiles;
  const int s = int(bs % P.S), b = int(bs / P.S);
  const int ty0 = (tile / P.tiles_x) * kATPy, tx0 = (tile % P.tiles_x) * kATPx;
  AltTaskInfo* out = P.plan + t;
  {
      float cx[4], cy[4];
      bool pv[4];
#pragma unroll
      for (int j = 0; j < 4; j++) {
        const int p = lane + 32 * j, y = ty0 + (p >> 4), x = tx0 + (p & 15);
        pv[j] = y < P.H1 && x < P.W1;
        cx[j] = cy[j] = 0.0f;
        if (pv[j]) {
          const float2 c = *reinterpret_cast<const float2*>(P.coords + b * P.c_sb + s * P.c_ss + (static_cast<long long>(y) * P.W1 + x) * P.c_sp);
          cx[j] = c.x; cy[j] = c.y;
        }
      }
      unsigned long long masks[kATMaxL];
      int bx0s[kATMaxL], by0s[kATMaxL];
#pragma unroll
      for (int l = 0; l < kATMaxL; l++) {
        masks[l] = 0ull; bx0s[l] = by0s[l] = 0;
        if (l >= P.L) continue;
        const int W2 = P.W2[l], H2 = P.H2[l];
        int lx[4], hx[4], ly[4], hy[4];
        int minx = 0x7fffffff, miny = 0x7fffffff;
#pragma unroll
        for (int j = 0; j < 4; j++) {
          float fr;
          const int wx0 = alt_floor_clamped(cx[j] * P.scale[l], &fr) - 3, wy0 = alt_floor_clamped(cy[j] * P.scale[l], &fr) - 3;
          lx[j] = max(wx0, 0); hx[j] = min(wx0 + 7, W2 - 1);
          ly[j] = max(wy0, 0); hy[j] = min(wy0 + 7, H2 - 1);
          if (!pv[j] || lx[j] > hx[j] || ly[j] > hy[j]) { lx[j] = 1; hx[j] = 0; }     // empty
          else { minx = min(minx, lx[j]); miny = min(miny, ly[j]); }
        }
        minx = __reduce_min_sync(0xffffffffu, minx);
        miny = __reduce_min_sync(0xffffffffu, miny);
        unsigned mlo = 0u, mhi = 0u;
        bool ovf = false;
        if (minx != 0x7fffffff) {
#pragma unroll
          for (int j = 0; j < 4; j++) {
            if (lx[j] > hx[j]) continue;
            const int c0 = (lx[j] - minx) >> 4, c1 = (hx[j] - minx) >> 4, r0 = (ly[j] - miny) >> 3, r1 = (hy[j] - miny) >> 3;
            if (c1 >= 8 || r1 >= 8) { ovf = true; continue; }
            for (int r = r0; r <= r1; r++)
              for (int c = c0; c <= c1; c++) {
                const int bit = r * 8 + c;
                if (bit < 32) mlo |= 1u << bit; else mhi |= 1u << (bit - 32);
              }
          }
        }
        mlo = __reduce_or_sync(0xffffffffu, mlo);
        mhi = __reduce_or_sync(0xffffffffu, mhi);
        ovf = __any_sync(0xffffffffu, ovf);
        int skip = 0;
        if (ovf) {                                     // cells beyond the 8x8 mask: this (task, level) goes to the SIMT kernel
          mlo = mhi = 0u; skip = 1;
          if (lane == 0) {
            const int pos = atomicAdd(P.status + 1, 1);
            if (pos < P.list_cap) P.list[pos] = int(t * kATMaxL + l);
            else atomicExch(P.status, 2);
          }
        }
        masks[l] = (static_cast<unsigned long long>(mhi) << 32) | mlo;
        bx0s[l] = minx == 0x7fffffff ? 0 : minx;
        by0s[l] = miny == 0x7fffffff ? 0 : miny;
        if (lane == 0) {
          (*out).bx0[l] = bx0s[l]; (*out).by0[l] = by0s[l];
          (*out).mask[l] = masks[l]; (*out).skip[l] = skip;
        }
      }
    if (lane == 0) {
      int ncell = 0;
      for (int l = 0; l < P.L; l++) ncell += __popcll(masks[l]);
      if (ncell) atomicAdd(P.status + 2, ncell);
    }
  }
}

__global__ void __launch_bounds__(kATThreads, 1) altcorr_tc_fwd_kernel(const __grid_constant__ AltTcParams P) {
  extern __shared__ unsigned char smem_dyn[];
  __shared__ __align__(8) uint64_t bars[24];           // a_full[2] a_empty[2] b_full[4] b_empty[4] t_full[4] t_empty[4] i_full[2] i_empty[2]
  __shared__ uint32_t tmem_base_s;
  __shared__ AltTaskInfo tinfo[2];

  const unsigned flags = P.flags ? *P.flags : 0u;
  if (flags & 2u) return;                              // non-finite / out-of-range operands: the SIMT kernel does the call
  const bool lo = P.has_lo && (flags & 1u);
  const int NA = lo ? 1 : 2, NB = lo ? 2 : 4;            // powers of two: ring positions by mask / shift (no division per cell)
  const int na_shift = lo ? 0 : 1, nb_shift = lo ? 1 : 2;
  const uint32_t tile_bytes = lo ? 4 * kATBlk : 2 * kATBlk;

  const uint32_t s0 = (smem_u32(smem_dyn) + 1023u) & ~1023u;
  const uint32_t sA = s0, sB = s0 + kATARegion;
  float* D = reinterpret_cast<float*>(smem_dyn + (s0 - smem_u32(smem_dyn)) + kATARegion + kATBRegion);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  auto a_full = [&](int i) { return smem_u32(&bars[i]); };
  auto a_empty = [&](int i) { return smem_u32(&bars[2 + i]); };
  auto b_full = [&](int i) { return smem_u32(&bars[4 + i]); };
  auto b_empty = [&](int i) { return smem_u32(&bars[8 + i]); };
  auto t_full = [&](int i) { return smem_u32(&bars[12 + i]); };
  auto t_empty = [&](int i) { return smem_u32(&bars[16 + i]); };
  auto i_full = [&](int i) { return smem_u32(&bars[20 + i]); };
  auto i_empty = [&](int i) { return smem_u32(&bars[22 + i]); };

  if (tid == 0) {
    for (int i = 0; i < 16; i++) mbar_init(smem_u32(&bars[i]), 1);
    mbar_init(a_empty(0), 2); mbar_init(a_empty(1), 2);                    // both MMA issuers release an A tile
    for (int i = 0; i < 4; i++) mbar_init(t_empty(i), kATEpiWarps);
    mbar_init(i_full(0), 1); mbar_init(i_full(1), 1);
    mbar_init(i_empty(0), 2 + kATEpiWarps); mbar_init(i_empty(1), 2 + kATEpiWarps);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_s)), "n"(512) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem_base = tmem_base_s;

  const int tiles = P.tiles_x * P.tiles_y;
  const long long total = static_cast<long long>(P.B) * P.S * tiles;
  // debug (P.dbg & 16): cycles each role spends in each wait, summed over the CTAs into status words 8.. (long long[12])
  long long tw[6] = {0, 0, 0, 0, 0, 0};
  const bool timing = (P.dbg & 16) != 0;
  auto twait = [&](int slot_, uint32_t bar_, uint32_t par_) -> bool {
    if (!timing) return mbar_wait(bar_, par_);
    const long long c0 = clock64();
    const bool ok_ = mbar_wait(bar_, par_);
    tw[slot_] += clock64() - c0;
    return ok_;
  };
  const long long t_start = clock64();

  if (warp == 0) {
    // ===================================================================== producer
    // the cell origins / masks of every tile were computed by altcorr_plan_kernel: the producer reads its 80-byte plan
    // entry (the next tile's entry is prefetched), hands it to the other roles through shared memory and issues the loads
    uint32_t it = 0;                                   // cells loaded so far (ring position); only its low bits matter
    int k = 0;
    bool dead = false;
    int pre = 0;
    if (lane < 20 && static_cast<long long>(blockIdx.x) < total) pre = reinterpret_cast<const int*>(P.plan + blockIdx.x)[lane];
    for (long long t = blockIdx.x; t < total && !dead; t += gridDim.x, k++) {
      const int slot = k & 1;
      const int tile = int(t % tiles);
      const long long bs = t / tiles;
      const int b = int(bs / P.S);
      const int ty0 = (tile / P.tiles_x) * kATPy, tx0 = (tile % P.tiles_x) * kATPx;
      if (!twait(0, i_empty(slot), ((k >> 1) & 1) ^ 1)) { atomicExch(P.status, 1); break; }
      if (lane < 20) reinterpret_cast<int*>(&tinfo[slot])[lane] = pre;
      if (lane < 20 && t + gridDim.x < total) pre = reinterpret_cast<const int*>(P.plan + t + gridDim.x)[lane];
      __syncwarp();
      if (lane == 0) {
        mbar_arrive(i_full(slot));
        const int f1 = P.idx1 ? int(P.idx1[b]) : b, f2 = P.idx2 ? int(P.idx2[b]) : b;
        const int as = k & (NA - 1);
        if (!twait(1, a_empty(as), ((k >> na_shift) & 1) ^ 1)) { atomicExch(P.status, 1); dead = true; }
        if (!dead) {
          const uint32_t dst = sA + uint32_t(as) * tile_bytes;
          mbar_expect_tx(a_full(as), tile_bytes);
          tma_load_tile(dst, &P.a_hi, tx0, ty0, f1, a_full(as));
          if (lo) tma_load_tile(dst + 2 * kATBlk, &P.a_lo, tx0, ty0, f1, a_full(as));
          for (int l = 0; l < P.L && !dead; l++) {
            unsigned long long m = tinfo[slot].mask[l];
            const int bx0 = tinfo[slot].bx0[l], by0 = tinfo[slot].by0[l];
            while (m) {
              const int bit = __ffsll(static_cast<long long>(m)) - 1;
              m &= m - 1;
              const int X0 = bx0 + kATCx * (bit & 7), Y0 = by0 + kATCy * (bit >> 3);
              const int st = int(it & uint32_t(NB - 1));
              if (!twait(2, b_empty(st), ((it >> nb_shift) & 1u) ^ 1u)) { atomicExch(P.status, 1); dead = true; break; }
              const uint32_t d2 = sB + uint32_t(st) * tile_bytes;
              mbar_expect_tx(b_full(st), tile_bytes);
              tma_load_tile(d2, &P.b_hi[l], X0, Y0, f2, b_full(st));
              if (lo) tma_load_tile(d2 + 2 * kATBlk, &P.b_lo[l], X0, Y0, f2, b_full(st));
              it++;
            }
          }
        }
      }
      dead = __shfl_sync(0xffffffffu, int(dead), 0) != 0;
    }
    if (timing && lane == 0) {
      long long* dt = reinterpret_cast<long long*>(P.status + 8);
      atomicAdd(reinterpret_cast<unsigned long long*>(dt + 0), static_cast<unsigned long long>(tw[0]));
      atomicAdd(reinterpret_cast<unsigned long long*>(dt + 1), static_cast<unsigned long long>(tw[1]));
      atomicAdd(reinterpret_cast<unsigned long long*>(dt + 2), static_cast<unsigned long long>(tw[2]));
      atomicAdd(reinterpret_cast<unsigned long long*>(dt + 3), static_cast<unsigned long long>(clock64() - t_start));
    }
  } else if (warp == 1 || warp == kATMmaWarpB) {
    // ===================================================================== MMA issuers
    // Two single-thread issuers: the fixed cost of a cell in the issuing thread (8 tcgen05.mma, two commits, two or three
    // barrier waits: ~0.65 us) exceeded the epilogue's once that was trimmed; cells alternate between the two threads by the
    // parity of the running cell counter (ring stage and accumulator follow from the same counter, so each thread has its own
    // stages {par, par + 2} and accumulators {par, par + 2}); both wait for the A tile and both release it (count 2).
    if (lane == 0) {
      const uint32_t par = (warp == 1) ? 0u : 1u;
      uint32_t it = 0;
      int k = 0;
      bool dead = false;
      for (long long t = blockIdx.x; t < total && !dead; t += gridDim.x, k++) {
        const int slot = k & 1;
        if (!twait(0, i_full(slot), (k >> 1) & 1)) { atomicExch(P.status, 1); break; }
        int cells = 0;
        for (int l = 0; l < P.L; l++) cells += __popcll(tinfo[slot].mask[l]);
        mbar_arrive(i_empty(slot));
        const int as = k & (NA - 1);
        if (!twait(1, a_full(as), (k >> na_shift) & 1)) { atomicExch(P.status, 1); break; }
        const uint32_t aT = sA + uint32_t(as) * tile_bytes;
        for (int c = 0; c < cells; c++, it++) {
          if ((P.dbg & 64) ? (par != 0u) : ((it & 1u) != par)) continue;     // dbg 64: issuer A takes every cell
          const int acc = int(it & 3u), st = int(it & uint32_t(NB - 1));
          if (!twait(2, t_empty(acc), (uint32_t(it >> 2) & 1u) ^ 1u)) { atomicExch(P.status, 1); dead = true; break; }
          if (!twait(3, b_full(st), (it >> nb_shift) & 1u)) { atomicExch(P.status, 1); dead = true; break; }
          asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
          const uint32_t bT = sB + uint32_t(st) * tile_bytes;
          const uint32_t d = tmem_base + uint32_t(acc) * 128u;
          const int nprod = lo ? 3 : 1;
          uint32_t first = 0u;
          for (int pr = 0; pr < nprod; pr++) {           // hi.hi, hi.lo, lo.hi
            const uint32_t ao = (pr == 2) ? 2 * kATBlk : 0u, bo = (pr == 1) ? 2 * kATBlk : 0u;
#pragma unroll
            for (int kb = 0; kb < 2; kb++)
#pragma unroll
              for (int kk = 0; kk < 4; kk++) {
                if (!(P.dbg & 32)) umma_f16(d, umma_desc_sw128(aT + ao + kb * kATBlk + kk * 32), umma_desc_sw128(bT + bo + kb * kATBlk + kk * 32), kATIdesc, first);
                first = 1u;
              }
          }
          umma_commit(b_empty(st));
          umma_commit(t_full(acc));
        }
        if (!dead) umma_commit(a_empty(as));
      }
      if (timing && par == 0u) {
        long long* dt = reinterpret_cast<long long*>(P.status + 8);
        for (int i = 0; i < 4; i++) atomicAdd(reinterpret_cast<unsigned long long*>(dt + 4 + i), static_cast<unsigned long long>(tw[i]));
        atomicAdd(reinterpret_cast<unsigned long long*>(dt + 8), static_cast<unsigned long long>(clock64() - t_start));
      }
    }
  } else {
    // ===================================================================== epilogue
    const int ew = warp - 2;
    const int q = warp & 3;                            // TMEM lane quarter this warp may read
    const int h = ew >> 2;                             // which 32 accumulator columns (= cell rows 2h, 2h+1); output rows 2h, 2h+1
    const int p = 32 * q + lane;                       // pixel of the tile = TMEM lane = A row
    uint32_t it = 0;
    int k = 0;
    bool dead = false;
    for (long long t = blockIdx.x; t < total && !dead; t += gridDim.x, k++) {
      const int slot = k & 1;
      const int tile = int(t % tiles);
      const long long bs = t / tiles;
      const int s = int(bs % P.S), b = int(bs / P.S);
      const int y = (tile / P.tiles_x) * kATPy + (p >> 4), x = (tile % P.tiles_x) * kATPx + (p & 15);
      const bool valid = y < P.H1 && x < P.W1;
      float pcx = 0.0f, pcy = 0.0f;
      if (valid) {
        const float2 c = *reinterpret_cast<const float2*>(P.coords + b * P.c_sb + s * P.c_ss + (static_cast<long long>(y) * P.W1 + x) * P.c_sp);
        pcx = c.x; pcy = c.y;
      }
      if (!twait(0, i_full(slot), (k >> 1) & 1)) { if (lane == 0) atomicExch(P.status, 1); break; }
      float* outp = P.out + b * P.o_sb + s * P.o_ss + (static_cast<long long>(y) * P.W1 + x) * P.o_sp;
      for (int l = 0; l < P.L && !dead; l++) {
        if (!(P.dbg & 8)) alt_pair_bar(q);             // the partner warps' output stage of the previous level has read D
        float dx, dy;
        const int wx0 = alt_floor_clamped(pcx * P.scale[l], &dx) - 3, wy0 = alt_floor_clamped(pcy * P.scale[l], &dy) - 3;
        // taps inside the target image: exactly these are written by the visited cells (the masks cover every in-image tap
        // of every valid pixel); the others are selected to zero in the combination below — D is never cleared
        // bit i: 0 <= w0 + i < size, i = 0..7 — a contiguous run of bits [max(0, -w0), min(8, size - w0))
        auto inrange = [](int w0, int size) -> unsigned {
          const int lo_ = max(0, -w0), hi_ = min(8, size - w0);
          return hi_ > lo_ ? ((1u << hi_) - (1u << lo_)) : 0u;
        };
        const unsigned rowok = inrange(wy0, P.H2[l]), colok = inrange(wx0, P.W2[l]);
        unsigned long long m = tinfo[slot].mask[l];
        const int bx0 = tinfo[slot].bx0[l], by0 = tinfo[slot].by0[l];
        while (m) {
          const int bit = __ffsll(static_cast<long long>(m)) - 1;
          m &= m - 1;
          const int relx = wx0 - (bx0 + kATCx * (bit & 7)), rely = wy0 - (by0 + kATCy * (bit >> 3));
          const int acc = int(it & 3);
          const bool ok = twait(1, t_full(acc), uint32_t(it >> 2) & 1u);
          asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
          if (!ok) { if (lane == 0) atomicExch(P.status, 1); dead = true; break; }
          {
            // cell rows 2h, 2h + 1: skip the TMEM read when no lane of this warp has a window row there
            const int ty0 = 2 * h - rely;
            const bool need = valid && (static_cast<unsigned>(ty0) < 8u || static_cast<unsigned>(ty0 + 1) < 8u) &&
                              relx > -8 && relx < 16;
            if (__any_sync(0xffffffffu, need) && !(P.dbg & 4)) {
            uint32_t v[32];
            tmem_ld32(tmem_base + (uint32_t(32 * q) << 16) + uint32_t(acc * 128 + 32 * h), v);
            // D[(ty * 8 + tx) * 128 + p] with ty = ty0 + rr, tx = c - relx: ONE base address, the column c and the row rr are
            // immediate offsets (c * 128 and rr * 1024 floats); never dereferenced where the predicate is false
            if (!(P.dbg & 1)) {
              float* dbase = D + ((ty0 * 8 - relx) * 128 + p);
              // bit c of cm: column c of the cell row is one of the window's 8 columns (relx <= c <= relx + 7)
              const unsigned cm = (relx > -8 && relx < 16) ? (((0xFFu << (relx + 8)) >> 8) & 0xFFFFu) : 0u;
              if (valid && static_cast<unsigned>(ty0) < 8u) {
#pragma unroll
                for (int c = 0; c < 16; c++)
                  if (cm & (1u << c)) dbase[c * 128] = __uint_as_float(v[c]);
              }
              if (valid && static_cast<unsigned>(ty0 + 1) < 8u) {
#pragma unroll
                for (int c = 0; c < 16; c++)
                  if (cm & (1u << c)) dbase[c * 128 + 1024] = __uint_as_float(v[16 + c]);
              }
            }
            }
          }
          asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
          __syncwarp();
          if (lane == 0) mbar_arrive(t_empty(acc));
          it++;
        }
        if (dead) break;
        if (!(P.dbg & 8)) alt_pair_bar(q);             // all four chunks' taps of this level are in D
        if (valid && !tinfo[slot].skip[l] && !(P.dbg & 2)) {
          // bilinear combination (altcorr_kernel.cu:107-140): channel oy + 7 ox =
          //   (1-dy)(1-dx) S[oy][ox] + (1-dy) dx S[oy][ox+1] + dy (1-dx) S[oy+1][ox] + dy dx S[oy+1][ox+1]
          const float w00 = (1.0f - dy) * (1.0f - dx), w01 = (1.0f - dy) * dx, w10 = dy * (1.0f - dx), w11 = dy * dx;
          // the 7 output rows over the 4 warps of the pixel: 2, 2, 2, 1 — rows oy0 (and oy0 + 1 for h < 3), straight-line code
          const int oy0 = 2 * h;
          auto tap_row = [&](int ty, float (&r)[8]) {        // taps of window row ty, out-of-image ones as zeros
            const unsigned mk = ((rowok >> ty) & 1u) ? colok : 0u;
            const float* drow = D + (ty * 8) * 128 + p;
#pragma unroll
            for (int c = 0; c < 8; c++) r[c] = 0.0f;
#pragma unroll
            for (int c = 0; c < 8; c++) if (mk & (1u << c)) r[c] = drow[c * 128];
          };
          const long long ostep = 7 * P.o_sc;
          auto out_row = [&](float* po, const float (&up)[8], const float (&dn)[8]) {   // channels oy + 7 ox, ox = 0..6
#pragma unroll
            for (int ox = 0; ox < 7; ox++, po += ostep) {
              float r = w00 * up[ox];
              r = fmaf(w01, up[ox + 1], r);
              r = fmaf(w10, dn[ox], r);
              r = fmaf(w11, dn[ox + 1], r);
              *po = r;
            }
          };
          float* ol = outp + static_cast<long long>(l) * 49 * P.o_sc + static_cast<long long>(oy0) * P.o_sc;
          float ra[8], rb[8];
          tap_row(oy0, ra);
          tap_row(oy0 + 1, rb);
          out_row(ol, ra, rb);
          if (h < 3) {
            tap_row(oy0 + 2, ra);
            out_row(ol + P.o_sc, rb, ra);
          }
        }
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(i_empty(slot));
    }
    if (timing && ew == 0 && lane == 0) {
      long long* dt = reinterpret_cast<long long*>(P.status + 8);
      atomicAdd(reinterpret_cast<unsigned long long*>(dt + 9), static_cast<unsigned long long>(tw[0]));
      atomicAdd(reinterpret_cast<unsigned long long*>(dt + 10), static_cast<unsigned long long>(tw[1]));
      atomicAdd(reinterpret_cast<unsigned long long*>(dt + 11), static_cast<unsigned long long>(clock64() - t_start));
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 1) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(512) : "memory");
}

// ---- SIMT path: (task, level) entries the tensor-core kernel handed over, or the whole call when flags bit 1 is set ----
// One warp per source pixel: lane l holds channels 4l..4l+3 of the pixel's feature vector; a tap is one coalesced read of
// the fmap2 position, an xor-butterfly finishes the dot product.  T = float reads the caller's fp32 operands unchanged.
struct AltSimtParams {
  const void* f1; const void* f2[kATMaxL];
  const float* coords; long long c_sb, c_ss, c_sp;
  float* out; long long o_sb, o_ss, o_sc, o_sp;
  const long long* idx1; const long long* idx2;
  const unsigned* flags; const int* status; const int* list; int list_cap;
  int B, S, H1, W1, L, C;
  int H2[kATMaxL], W2[kATMaxL];
  float scale[kATMaxL];
  int tiles_x, tiles_y;
};

template <typename T> __device__ __forceinline__ float4 alt_ld4(const T* p);
template <> __device__ __forceinline__ float4 alt_ld4<float>(const float* p) { return *reinterpret_cast<const float4*>(p); }
// address of channels 4 * lane .. 4 * lane + 3 of position `pos` (= (frame * H + y) * W + x): channels-last fp32 / blocked fp16
template <typename T> __device__ __forceinline__ const T* alt_chan_ptr(const T* base, long long frame, long long hw, long long pix, int lane);
template <> __device__ __forceinline__ const float* alt_chan_ptr<float>(const float* base, long long frame, long long hw, long long pix, int lane) {
  return base + (frame * hw + pix) * 128 + lane * 4;
}
template <> __device__ __forceinline__ const __half* alt_chan_ptr<__half>(const __half* base, long long frame, long long hw, long long pix, int lane) {
  return base + ((frame * 2 + (lane >> 4)) * hw + pix) * 64 + (lane & 15) * 4;
}
template <> __device__ __forceinline__ float4 alt_ld4<__half>(const __half* p) {
  const uint2 u = *reinterpret_cast<const uint2*>(p);
  const float2 a = __half22float2(*reinterpret_cast<const __half2*>(&u.x)), b = __half22float2(*reinterpret_cast<const __half2*>(&u.y));
  return make_float4(a.x, a.y, b.x, b.y);
}

template <typename T>
__global__ void __launch_bounds__(256) altcorr_simt_list_kernel(const __grid_constant__ AltSimtParams P) {
  __shared__ float sdot[8][64];
  const unsigned flags = P.flags ? *P.flags : 0u;
  const int tiles = P.tiles_x * P.tiles_y;
  const long long total = static_cast<long long>(P.B) * P.S * tiles;
  const bool all = (flags & 2u) != 0u;
  const long long entries = all ? total * P.L : min(P.status[1], P.list_cap);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float* dots = sdot[warp];
  for (long long e = blockIdx.x; e < entries; e += gridDim.x) {
    long long t; int l;
    if (all) { t = e / P.L; l = int(e % P.L); } else { const int v = P.list[e]; t = v / kATMaxL; l = v % kATMaxL; }
    const int tile = int(t % tiles);
    const long long bs = t / tiles;
    const int s = int(bs % P.S), b = int(bs / P.S);
    const int ty0 = (tile / P.tiles_x) * kATPy, tx0 = (tile % P.tiles_x) * kATPx;
    const int H2 = P.H2[l], W2 = P.W2[l];
    const long long fr1 = P.idx1 ? P.idx1[b] : b, fr2 = P.idx2 ? P.idx2[b] : b;
    const T* f2b = static_cast<const T*>(P.f2[l]);
    const long long hw2 = static_cast<long long>(H2) * W2, hw1 = static_cast<long long>(P.H1) * P.W1;
    for (int pp = warp; pp < kATPy * kATPx; pp += 8) {
      const int y = ty0 + (pp >> 4), x = tx0 + (pp & 15);
      if (y >= P.H1 || x >= P.W1) continue;
      const float2 c = *reinterpret_cast<const float2*>(P.coords + b * P.c_sb + s * P.c_ss + (static_cast<long long>(y) * P.W1 + x) * P.c_sp);
      float dx, dy;
      const int wx0 = alt_floor_clamped(c.x * P.scale[l], &dx) - 3, wy0 = alt_floor_clamped(c.y * P.scale[l], &dy) - 3;
      const float4 a = alt_ld4<T>(alt_chan_ptr<T>(static_cast<const T*>(P.f1), fr1, hw1, static_cast<long long>(y) * P.W1 + x, lane));
      for (int tp = 0; tp < 64; tp++) {
        const int h2 = wy0 + (tp >> 3), w2 = wx0 + (tp & 7);
        float acc = 0.0f;
        if (h2 >= 0 && h2 < H2 && w2 >= 0 && w2 < W2) {
          const float4 qv = alt_ld4<T>(alt_chan_ptr<T>(f2b, fr2, hw2, static_cast<long long>(h2) * W2 + w2, lane));
          acc = fmaf(a.x, qv.x, acc); acc = fmaf(a.y, qv.y, acc); acc = fmaf(a.z, qv.z, acc); acc = fmaf(a.w, qv.w, acc);
          acc = warp_sum(acc);
        }
        if (lane == 0) dots[tp] = acc;
      }
      __syncwarp();
      const float w00 = (1.0f - dy) * (1.0f - dx), w01 = (1.0f - dy) * dx, w10 = dy * (1.0f - dx), w11 = dy * dx;
      float* ol = P.out + b * P.o_sb + s * P.o_ss + (static_cast<long long>(y) * P.W1 + x) * P.o_sp + static_cast<long long>(l) * 49 * P.o_sc;
      for (int o = lane; o < 49; o += 32) {
        const int ox = o / 7, oy = o - ox * 7;
        float r = w00 * dots[oy * 8 + ox];
        r = fmaf(w01, dots[oy * 8 + ox + 1], r);
        r = fmaf(w10, dots[(oy + 1) * 8 + ox], r);
        r = fmaf(w11, dots[(oy + 1) * 8 + ox + 1], r);
        ol[static_cast<long long>(o) * P.o_sc] = r;
      }
      __syncwarp();
    }
  }
}

// ---- fp32 -> fp16 hi + lo split (the lo parts make the tensor-core product fp32-faithful) -----------------------------
// x: channels-last [F][HW][128] fp32  ->  hi, lo: channel-blocked [F][2][HW][64] fp16 (hw = 0: same layout as x)
__global__ void __launch_bounds__(256) altcorr_split_kernel(const float* __restrict__ x, __half* __restrict__ hi, __half* __restrict__ lo,
                                                            size_t n4, unsigned* __restrict__ flags, size_t hw) {
  unsigned f = 0u;
  for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n4; i += size_t(gridDim.x) * blockDim.x) {
    const float4 v = reinterpret_cast<const float4*>(x)[i];
    const float a[4] = {v.x, v.y, v.z, v.w};
    __half hh[4], ll[4];
#pragma unroll
    for (int j = 0; j < 4; j++) {
      hh[j] = __float2half_rn(a[j]);
      const float back = __half2float(hh[j]);
      if (!(fabsf(back) <= 65504.0f)) f |= 2u;          // overflow to inf, or NaN
      const float r = a[j] - back;
      ll[j] = __float2half_rn(r);
      if (r != 0.0f) f |= 1u;
    }
    uint2 ph, pl;
    __half2 h01 = __halves2half2(hh[0], hh[1]), h23 = __halves2half2(hh[2], hh[3]);
    __half2 l01 = __halves2half2(ll[0], ll[1]), l23 = __halves2half2(ll[2], ll[3]);
    ph.x = *reinterpret_cast<uint32_t*>(&h01); ph.y = *reinterpret_cast<uint32_t*>(&h23);
    pl.x = *reinterpret_cast<uint32_t*>(&l01); pl.y = *reinterpret_cast<uint32_t*>(&l23);
    size_t o = i;
    if (hw) {                                           // i = (pos * 128 + c) / 4 -> ((f * 2 + c / 64) * hw + pix) * 64 + c % 64
      const size_t pos = i >> 5, c4 = i & 31, f = pos / hw, pix = pos - f * hw;
      o = (((f * 2 + (c4 >> 4)) * hw + pix) << 4) + (c4 & 15);
    }
    reinterpret_cast<uint2*>(hi)[o] = ph;
    reinterpret_cast<uint2*>(lo)[o] = pl;
  }
  f = __reduce_or_sync(0xffffffffu, f);
  if (f && (threadIdx.x & 31) == 0) atomicOr(flags, f);
}

// ======================================================================================================================
// Backward (altcorr_kernel.cu:152-286) on the tensor cores, same tile x cell decomposition:
//   G[p, n]     = d loss / d S[p, n]: the 4-corner combination of corr_grad of pixel p at the window tap that falls on
//                 cell position n, 0 elsewhere (fp32 -> fp16 hi + lo, written by the threads in the UMMA operand layout);
//   d f1[tile] += G  . f2[cell]          M = pixels,    K = positions   (A = G K-major,  B = f2 cell  MN-major)
//   d f2[cell]  = G^T . f1[tile]         M = positions, K = pixels      (A = G MN-major, B = f1 tile  MN-major)
// The SAME shared-memory image of G serves both products (K-major for the first, MN-major for the second), and the
// feature tiles are used as TMA delivers them ([rows][128 channels]) through MN-major descriptors.  d f1 accumulates in
// TMEM over all cells and coordinate samples of the tile and is stored once; d f2 of a cell is drained from TMEM with one
// vector red.global.add per 4 channels (rows nobody touched are skipped) instead of one atomicAdd per tap and channel.
struct AltTcBwdParams {
  CUtensorMap a_hi, a_lo, b_hi, b_lo;
  const float* coords; long long c_sb, c_ss, c_sp;
  const float* grad; long long g_sb, g_ss, g_sc, g_sp;
  float* g1; float* g2;
  const unsigned* flags;
  int* status;                                         // [0] error, [1] tiles whose cells left the 8x8 mask (whole call redone on SIMT)
  int B, S, H1, W1, H2, W2;
  int tiles_x, tiles_y;
};

__device__ __forceinline__ uint64_t umma_desc_sw128_mn(uint32_t saddr) {     // MN-major: LBO = next 64-element MN block (16 KB), SBO = next 8 K-rows (1 KB)
  return uint64_t((saddr >> 4) & 0x3FFFu) | (uint64_t(kATBlk >> 4) << 16) | (uint64_t(1024 >> 4) << 32) | (uint64_t(1) << 46) | (uint64_t(2) << 61);
}
__device__ __forceinline__ void red_add_v4(float* addr, float a, float b, float c, float d) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

constexpr size_t kATBwdSmem = 3 * size_t(4 * kATBlk) + kATDBytes + 1024;       // f1, f2, G (hi+lo each) + tap table

__global__ void __launch_bounds__(256, 1) altcorr_tc_bwd_kernel(const __grid_constant__ AltTcBwdParams P) {
  extern __shared__ unsigned char smem_dyn[];
  __shared__ __align__(8) uint64_t bars[3];            // f1 tile landed, f2 cell landed, MMAs of the cell done
  __shared__ uint32_t tmem_base_s;
  __shared__ int s_min[2], s_ovf, s_ncell;
  __shared__ unsigned s_mask[2];
  __shared__ short s_cx[64], s_cy[64];

  const unsigned flags = P.flags ? *P.flags : 0u;
  if (flags & 2u) return;
  const bool lo = (flags & 1u) != 0u;
  const uint32_t s0 = (smem_u32(smem_dyn) + 1023u) & ~1023u;
  unsigned char* g0 = smem_dyn + (s0 - smem_u32(smem_dyn));
  const uint32_t sF1 = s0, sF2 = s0 + 4 * kATBlk, sG = s0 + 8 * kATBlk;
  unsigned char* Gp = g0 + 8 * kATBlk;
  float* T = reinterpret_cast<float*>(g0 + 12 * kATBlk);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const uint32_t bar_f1 = smem_u32(&bars[0]), bar_f2 = smem_u32(&bars[1]), bar_mma = smem_u32(&bars[2]);
  if (tid == 0) {
    mbar_init(bar_f1, 1); mbar_init(bar_f2, 1); mbar_init(bar_mma, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_s)), "n"(256) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem_base = tmem_base_s;

  const int p = tid & 127, hh = tid >> 7;              // pixel of the tile, and which half of the rows (taps / cell rows) this thread builds
  const int q = warp & 3, h = warp >> 2;               // TMEM lane quarter / column half for the drains
  const int tiles = P.tiles_x * P.tiles_y;
  const long long total = static_cast<long long>(P.B) * tiles;
  const uint32_t tile_bytes = lo ? 4 * kATBlk : 2 * kATBlk;
  constexpr uint32_t kIdesc1 = kATIdesc | (1u << 16);                 // A K-major, B MN-major
  constexpr uint32_t kIdesc2 = kATIdesc | (1u << 15) | (1u << 16);    // both MN-major
  uint32_t n_cells_done = 0;                           // parity of bar_f2 / bar_mma
  int k = 0;
  bool dead = false;
  for (long long t = blockIdx.x; t < total && !dead; t += gridDim.x, k++) {
    const int tile = int(t % tiles), b = int(t / tiles);
    const int ty0 = (tile / P.tiles_x) * kATPy, tx0 = (tile % P.tiles_x) * kATPx;
    const int y = ty0 + (p >> 4), x = tx0 + (p & 15);
    const bool valid = y < P.H1 && x < P.W1;
    if (tid == 0) {
      mbar_expect_tx(bar_f1, tile_bytes);
      tma_load_tile(sF1, &P.a_hi, tx0, ty0, b, bar_f1);
      if (lo) tma_load_tile(sF1 + 2 * kATBlk, &P.a_lo, tx0, ty0, b, bar_f1);
    }
    int task_cells = 0;
    for (int s = 0; s < P.S && !dead; s++) {
      if (tid == 0) { s_min[0] = s_min[1] = 0x7fffffff; s_mask[0] = s_mask[1] = 0u; s_ovf = 0; }
      __syncthreads();
      float dx = 0.f, dy = 0.f;
      int wx0 = 0, wy0 = 0, lx = 1, hx = 0, ly = 1, hy = 0;
      if (valid) {
        const float2 c = *reinterpret_cast<const float2*>(P.coords + b * P.c_sb + s * P.c_ss + (static_cast<long long>(y) * P.W1 + x) * P.c_sp);
        wx0 = alt_floor_clamped(c.x, &dx) - 3; wy0 = alt_floor_clamped(c.y, &dy) - 3;
        lx = max(wx0, 0); hx = min(wx0 + 7, P.W2 - 1); ly = max(wy0, 0); hy = min(wy0 + 7, P.H2 - 1);
        if (lx > hx || ly > hy) { lx = 1; hx = 0; }
        else if (hh == 0) { atomicMin(&s_min[0], lx); atomicMin(&s_min[1], ly); }
      }
      // tap table T[ty][tx][p] = d loss / d S[p][tap]: rows 4hh..4hh+3 by this thread
      {
        const float w00 = (1.0f - dy) * (1.0f - dx), w01 = (1.0f - dy) * dx, w10 = dy * (1.0f - dx), w11 = dy * dx;
        const float* gp = P.grad + b * P.g_sb + s * P.g_ss + (static_cast<long long>(y) * P.W1 + x) * P.g_sp;
        float up[7], cur[7];
        const int r0 = 4 * hh;
#pragma unroll
        for (int ox = 0; ox < 7; ox++) up[ox] = (valid && r0 > 0) ? gp[static_cast<long long>((r0 - 1) + 7 * ox) * P.g_sc] : 0.0f;
#pragma unroll
        for (int ty = r0; ty < r0 + 4; ty++) {
#pragma unroll
          for (int ox = 0; ox < 7; ox++) cur[ox] = (valid && ty < 7) ? gp[static_cast<long long>(ty + 7 * ox) * P.g_sc] : 0.0f;
#pragma unroll
          for (int tx = 0; tx < 8; tx++) {
            float g = 0.0f;
            if (tx > 0) { g = fmaf(w11, up[tx - 1], g); g = fmaf(w01, cur[tx - 1], g); }
            if (tx < 7) { g = fmaf(w10, up[tx], g); g = fmaf(w00, cur[tx], g); }
            T[(ty * 8 + tx) * 128 + p] = g;
          }
#pragma unroll
          for (int ox = 0; ox < 7; ox++) up[ox] = cur[ox];
        }
      }
      __syncthreads();
      const int bx0 = s_min[0], by0 = s_min[1];
      if (hh == 0 && lx <= hx) {
        const int c0 = (lx - bx0) >> 4, c1 = (hx - bx0) >> 4, r0 = (ly - by0) >> 3, r1 = (hy - by0) >> 3;
        if (c1 >= 8 || r1 >= 8) s_ovf = 1;
        else
          for (int r = r0; r <= r1; r++)
            for (int c = c0; c <= c1; c++) { const int bit = r * 8 + c; atomicOr(&s_mask[bit >> 5], 1u << (bit & 31)); }
      }
      __syncthreads();
      const bool ovf = s_ovf != 0;                    // cells beyond the 8x8 mask: the SIMT kernel redoes the whole call
      if (tid == 0 && !ovf) {
        int n = 0;
        unsigned long long m = (static_cast<unsigned long long>(s_mask[1]) << 32) | s_mask[0];
        while (m) { const int bit = __ffsll(static_cast<long long>(m)) - 1; m &= m - 1; s_cx[n] = short(bit & 7); s_cy[n] = short(bit >> 3); n++; }
        s_ncell = n;
        if (n > 0) {                                   // first cell of this sample
          mbar_expect_tx(bar_f2, tile_bytes);
          const int X0 = bx0 + kATCx * s_cx[0], Y0 = by0 + kATCy * s_cy[0];
          tma_load_tile(sF2, &P.b_hi, X0, Y0, b, bar_f2);
          if (lo) tma_load_tile(sF2 + 2 * kATBlk, &P.b_lo, X0, Y0, b, bar_f2);
        }
      }
      __syncthreads();
      if (ovf) { if (tid == 0) atomicAdd(P.status + 1, 1); continue; }
      const int ncell = s_ncell;
      for (int c = 0; c < ncell && !dead; c++, n_cells_done++, task_cells++) {
        const int X0 = bx0 + kATCx * s_cx[c], Y0 = by0 + kATCy * s_cy[c];
        // ---- G rows of pixel p for cell rows 4hh..4hh+3 (= operand block hh), hi and lo parts
        {
          const int relx = wx0 - X0, rely = wy0 - Y0;
          unsigned char* ghi = Gp + hh * kATBlk + p * 128;
          unsigned char* glo = ghi + 2 * kATBlk;
#pragma unroll
          for (int yy = 0; yy < 4; yy++) {
            const int ty = 4 * hh + yy - rely;
            const bool rowok = valid && static_cast<unsigned>(ty) < 8u;
            uint32_t wh[8], wl[8];
#pragma unroll
            for (int xp = 0; xp < 8; xp++) {
              float v[2];
#pragma unroll
              for (int e = 0; e < 2; e++) {
                const int tx = 2 * xp + e - relx;
                v[e] = (rowok && static_cast<unsigned>(tx) < 8u) ? T[(ty * 8 + tx) * 128 + p] : 0.0f;
              }
              const __half2 hv = __floats2half2_rn(v[0], v[1]);
              const float2 back = __half22float2(hv);
              const __half2 lv = __floats2half2_rn(v[0] - back.x, v[1] - back.y);
              wh[xp] = *reinterpret_cast<const uint32_t*>(&hv);
              wl[xp] = *reinterpret_cast<const uint32_t*>(&lv);
            }
#pragma unroll
            for (int ch = 0; ch < 2; ch++) {
              const int chunk = ((yy * 2 + ch) ^ (p & 7)) << 4;
              *reinterpret_cast<uint4*>(ghi + chunk) = make_uint4(wh[4 * ch], wh[4 * ch + 1], wh[4 * ch + 2], wh[4 * ch + 3]);
              *reinterpret_cast<uint4*>(glo + chunk) = make_uint4(wl[4 * ch], wl[4 * ch + 1], wl[4 * ch + 2], wl[4 * ch + 3]);
            }
          }
        }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");      // generic-proxy writes of G -> visible to the tensor core
        __syncthreads();
        if (tid == 0) {
          bool ok = true;
          if (task_cells == 0) ok = mbar_wait(bar_f1, k & 1);
          ok = ok && mbar_wait(bar_f2, n_cells_done & 1u);
          if (!ok) atomicExch(P.status, 1);
          asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
          const int nprod = lo ? 3 : 2;
          // d f1 (TMEM columns 0..127) += G . f2cell
          uint32_t acc1 = task_cells == 0 ? 0u : 1u;
          for (int pr = 0; pr < nprod; pr++) {          // Gh.Fh, Gl.Fh, Gh.Fl
            const uint32_t go = (pr == 1) ? 2 * kATBlk : 0u, fo = (pr == 2) ? 2 * kATBlk : 0u;
#pragma unroll
            for (int kk = 0; kk < 8; kk++) {
              umma_f16(tmem_base, umma_desc_sw128(sG + go + (kk >> 2) * kATBlk + (kk & 3) * 32), umma_desc_sw128_mn(sF2 + fo + kk * 2048), kIdesc1, acc1);
              acc1 = 1u;
            }
          }
          // d f2cell (TMEM columns 128..255) = G^T . f1tile
          uint32_t acc2 = 0u;
          for (int pr = 0; pr < nprod; pr++) {
            const uint32_t go = (pr == 1) ? 2 * kATBlk : 0u, fo = (pr == 2) ? 2 * kATBlk : 0u;
#pragma unroll
            for (int kk = 0; kk < 8; kk++) {
              umma_f16(tmem_base + 128u, umma_desc_sw128_mn(sG + go + kk * 2048), umma_desc_sw128_mn(sF1 + fo + kk * 2048), kIdesc2, acc2);
              acc2 = 1u;
            }
          }
          umma_commit(bar_mma);
        }
        __syncwarp();
        const bool ok = mbar_wait(bar_mma, n_cells_done & 1u);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        if (!ok) { if (lane == 0) atomicExch(P.status, 1); dead = true; }
        if (tid == 0 && c + 1 < ncell && ok) {          // the f2 stage is free again: fetch the next cell under the drain
          mbar_expect_tx(bar_f2, tile_bytes);
          const int X1 = bx0 + kATCx * s_cx[c + 1], Y1 = by0 + kATCy * s_cy[c + 1];
          tma_load_tile(sF2, &P.b_hi, X1, Y1, b, bar_f2);
          if (lo) tma_load_tile(sF2 + 2 * kATBlk, &P.b_lo, X1, Y1, b, bar_f2);
        }
        __syncwarp();
        // ---- drain d f2cell: TMEM lane = cell position
        {
          const int n = 32 * q + lane, Y = Y0 + (n >> 4), X = X0 + (n & 15);
          const bool inb = Y < P.H2 && X < P.W2;
          float* dst = P.g2 + ((static_cast<long long>(b) * P.H2 + Y) * P.W2 + X) * 128 + 64 * h;
#pragma unroll
          for (int j = 0; j < 2; j++) {
            uint32_t v[32];
            tmem_ld32(tmem_base + (uint32_t(32 * q) << 16) + uint32_t(128 + 64 * h + 32 * j), v);
            uint32_t any = 0u;
#pragma unroll
            for (int i = 0; i < 32; i++) any |= v[i];
            if (ok && inb && (any & 0x7fffffffu))
#pragma unroll
              for (int i = 0; i < 8; i++)
                red_add_v4(dst + 32 * j + 4 * i, __uint_as_float(v[4 * i]), __uint_as_float(v[4 * i + 1]), __uint_as_float(v[4 * i + 2]), __uint_as_float(v[4 * i + 3]));
          }
        }
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
        __syncthreads();                               // G, T-dependent state and the d f2 accumulator may be reused
      }
    }
    // ---- d f1 of the tile: TMEM lane = pixel
    {
      const int pp = 32 * q + lane, yy = ty0 + (pp >> 4), xx = tx0 + (pp & 15);
      float* dst = P.g1 + ((static_cast<long long>(b) * P.H1 + yy) * P.W1 + xx) * 128 + 64 * h;
      const bool pv = yy < P.H1 && xx < P.W1;
      if (task_cells == 0) {                           // nothing accumulated (all windows outside the image): zeros
        if (tid == 0 && !mbar_wait(bar_f1, k & 1)) atomicExch(P.status, 1);     // keep the barrier phase in step
        if (pv)
#pragma unroll
          for (int i = 0; i < 16; i++) reinterpret_cast<float4*>(dst)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
      } else {
#pragma unroll
        for (int j = 0; j < 2; j++) {
          uint32_t v[32];
          tmem_ld32(tmem_base + (uint32_t(32 * q) << 16) + uint32_t(64 * h + 32 * j), v);
          if (pv)
#pragma unroll
            for (int i = 0; i < 8; i++)
              reinterpret_cast<float4*>(dst + 32 * j)[i] = make_float4(__uint_as_float(v[4 * i]), __uint_as_float(v[4 * i + 1]), __uint_as_float(v[4 * i + 2]), __uint_as_float(v[4 * i + 3]));
        }
      }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();                                   // f1 stage and the d f1 accumulator are free for the next tile
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 1) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(256) : "memory");
}

// ---- host ---------------------------------------------------------------------------------------------------------------
// Operand layout of the tensor-core kernels: CHANNEL-BLOCKED fp16 feature maps [F][2][H][W][64] (channel c of position
// (y, x) of frame f at ((f * 2 + c / 64) * H * W + y * W + x) * 64 + c % 64).  The 128-byte swizzle limits a TMA box to 64
// channels; in a channels-last [F][H][W][128] array such a box is 128 separate 128-byte segments (positions are 256 bytes
// apart) and the TMA unit of the SM, not the tensor pipe or the epilogue, bounded the kernel (1.5 us per cell: 82 % of the
// time remained with the whole epilogue switched off).  Blocked, a box [64 ch x 16 x 8 x 2 blocks x 1 frame] is 16 runs of 2 KB and ONE TMA instruction per operand tile.
// It lands in shared memory as [8*16 rows][64 halfs], the K-major UMMA operand block.
static bool alt_make_map(CUtensorMap* m, const void* base, long long F, int H, int W) {
  EncodeTiledFn fn = encode_fn();
  if (!fn) return false;
  const cuuint64_t gdim[5] = {64, cuuint64_t(W), cuuint64_t(H), 2, cuuint64_t(F)};
  const cuuint64_t gstride[4] = {128, cuuint64_t(W) * 128, cuuint64_t(H) * W * 128, cuuint64_t(H) * W * 256};
  const cuuint32_t box[5] = {64, kATCx, kATCy, 2, 1};
  const cuuint32_t estr[5] = {1, 1, 1, 1, 1};
  return fn(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 5, const_cast<void*>(base), gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
            CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

bool altcorr_tc_supported(int C, int radius) {
  static const bool off = getenv("DB200_ALTCORR_NO_TCGEN05") != nullptr;
  return !off && C == 128 && radius == 3 && encode_fn() != nullptr;
}

struct AltLaunch {
  const void* f1_hi; const void* f1_lo;               // [F1][H1][W1][128] fp16 (lo may be null)
  const void* f2_hi[kATMaxL]; const void* f2_lo[kATMaxL];
  const void* f1_raw; const void* f2_raw[kATMaxL];    // operands as the caller passed them (SIMT path)
  int raw_is_half;
  long long F1, F2;
  int H1, W1, L;
  int H2[kATMaxL], W2[kATMaxL];
  float scale[kATMaxL];
  const long long* idx1; const long long* idx2;
  const float* coords; long long c_sb, c_ss, c_sp;
  float* out; long long o_sb, o_ss, o_sc, o_sp;
  int B, S;
  unsigned* flags;                                    // device word (may be null when there are no lo parts)
  int* status;                                        // device: 2 ints + list
  int list_cap;
};

// status area of a call: [256 B status words][SIMT hand-over list: tasks * kATMaxL ints][plan: tasks * 80 B]
static inline long long alt_tasks(int B, int S, int H1, int W1) { return static_cast<long long>(B) * S * cdiv(H1, kATPy) * cdiv(W1, kATPx); }
static inline size_t alt_list_bytes(long long tasks) { return (size_t(tasks) * kATMaxL * sizeof(int) + 15) / 16 * 16; }
size_t altcorr_tc_status_bytes(int B, int S, int H1, int W1) {
  const long long tasks = alt_tasks(B, S, H1, W1);
  return 256 + alt_list_bytes(tasks) + size_t(tasks) * sizeof(AltTaskInfo);
}

int altcorr_tc_launch(const AltLaunch& a, cudaStream_t st) {
  AltTcParams P;
  AltSimtParams Q;
  if (!alt_make_map(&P.a_hi, a.f1_hi, a.F1, a.H1, a.W1)) return DB200_ERR_CUDA;
  if (a.f1_lo) { if (!alt_make_map(&P.a_lo, a.f1_lo, a.F1, a.H1, a.W1)) return DB200_ERR_CUDA; } else P.a_lo = P.a_hi;
  for (int l = 0; l < kATMaxL; l++) {
    const int ll = l < a.L ? l : 0;
    if (!alt_make_map(&P.b_hi[l], a.f2_hi[ll], a.F2, a.H2[ll], a.W2[ll])) return DB200_ERR_CUDA;
    if (a.f2_lo[ll]) { if (!alt_make_map(&P.b_lo[l], a.f2_lo[ll], a.F2, a.H2[ll], a.W2[ll])) return DB200_ERR_CUDA; } else P.b_lo[l] = P.b_hi[l];
    P.H2[l] = Q.H2[l] = a.H2[ll]; P.W2[l] = Q.W2[l] = a.W2[ll]; P.scale[l] = Q.scale[l] = a.scale[ll];
    Q.f2[l] = a.f2_raw[ll];
  }
  P.coords = Q.coords = a.coords; P.c_sb = Q.c_sb = a.c_sb; P.c_ss = Q.c_ss = a.c_ss; P.c_sp = Q.c_sp = a.c_sp;
  P.out = Q.out = a.out; P.o_sb = Q.o_sb = a.o_sb; P.o_ss = Q.o_ss = a.o_ss; P.o_sc = Q.o_sc = a.o_sc; P.o_sp = Q.o_sp = a.o_sp;
  P.idx1 = Q.idx1 = a.idx1; P.idx2 = Q.idx2 = a.idx2;
  P.flags = Q.flags = a.flags;
  P.status = a.status; Q.status = a.status;
  const long long tasks = alt_tasks(a.B, a.S, a.H1, a.W1);
  P.list = a.status + 64; Q.list = a.status + 64; P.list_cap = Q.list_cap = int(tasks * kATMaxL);
  P.plan = reinterpret_cast<AltTaskInfo*>(reinterpret_cast<char*>(a.status) + 256 + alt_list_bytes(tasks));
  P.B = Q.B = a.B; P.S = Q.S = a.S; P.H1 = Q.H1 = a.H1; P.W1 = Q.W1 = a.W1; P.L = Q.L = a.L; Q.C = 128;
  P.tiles_x = Q.tiles_x = cdiv(a.W1, kATPx); P.tiles_y = Q.tiles_y = cdiv(a.H1, kATPy);
  P.has_lo = a.f1_lo != nullptr;
  static const int dbg_env = getenv("DB200_ALT_DBG") ? atoi(getenv("DB200_ALT_DBG")) : 0;
  P.dbg = dbg_env;
  Q.f1 = a.f1_raw;
  DB200_CUDA(cudaMemsetAsync(a.status, 0, 32 * sizeof(int), st));
  static bool attr = false;
  if (!attr) {
    DB200_CUDA(cudaFuncSetAttribute(altcorr_tc_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, int(kATSmem)));
    attr = true;
  }
  altcorr_plan_kernel<<<unsigned((tasks + 7) / 8), 256, 0, st>>>(P);
  DB200_CHECK_LAUNCH();
  const int grid = int(tasks < kNumSMs ? tasks : kNumSMs);
  altcorr_tc_fwd_kernel<<<grid, kATThreads, kATSmem, st>>>(P);
  DB200_CHECK_LAUNCH();
  const int g2 = int(tasks * a.L < 8 * kNumSMs ? tasks * a.L : 8 * kNumSMs);
  if (a.raw_is_half) altcorr_simt_list_kernel<__half><<<g2, 256, 0, st>>>(Q);
  else altcorr_simt_list_kernel<float><<<g2, 256, 0, st>>>(Q);
  DB200_CHECK_LAUNCH();
  return DB200_OK;
}

int altcorr_split(const float* x, void* hi, void* lo, size_t n, unsigned* flags, cudaStream_t st, size_t hw = 0) {
  if (n % 4 != 0 || (hw && n % (hw * 128) != 0)) return DB200_ERR_ARG;
  const size_t n4 = n / 4;
  const int blocks = int((n4 + 255) / 256 < size_t(8 * kNumSMs) ? (n4 + 255) / 256 : size_t(8 * kNumSMs));
  altcorr_split_kernel<<<blocks > 0 ? blocks : 1, 256, 0, st>>>(x, static_cast<__half*>(hi), static_cast<__half*>(lo), n4, flags, hw);
  DB200_CHECK_LAUNCH();
  return DB200_OK;
}


// SIMT backward of altcorr.cu, run only when the device words ask for it (flags bit 1, or status[1] != 0)
int altcorr_bwd_simt_conditional(const float* fmap1, const float* fmap2, const float* coords, const float* corr_grad,
                                 float* fmap1_grad, float* fmap2_grad, int B, int S, int H1, int W1, int H2, int W2, int C,
                                 int radius, const unsigned* flags, const int* status, cudaStream_t st);

int altcorr_tc_bwd_launch(const void* h1, const void* l1, const void* h2, const void* l2, const float* coords, const float* grad,
                          float* g1, float* g2, int B, int S, int H1, int W1, int H2, int W2, unsigned* flags, int* status,
                          cudaStream_t st) {
  AltTcBwdParams P;
  if (!alt_make_map(&P.a_hi, h1, B, H1, W1) || !alt_make_map(&P.a_lo, l1, B, H1, W1) || !alt_make_map(&P.b_hi, h2, B, H2, W2) ||
      !alt_make_map(&P.b_lo, l2, B, H2, W2))
    return DB200_ERR_CUDA;
  P.coords = coords; P.c_sb = static_cast<long long>(S) * H1 * W1 * 2; P.c_ss = static_cast<long long>(H1) * W1 * 2; P.c_sp = 2;
  P.grad = grad; P.g_sb = static_cast<long long>(S) * 49 * H1 * W1; P.g_ss = static_cast<long long>(49) * H1 * W1;
  P.g_sc = static_cast<long long>(H1) * W1; P.g_sp = 1;
  P.g1 = g1; P.g2 = g2; P.flags = flags; P.status = status;
  P.B = B; P.S = S; P.H1 = H1; P.W1 = W1; P.H2 = H2; P.W2 = W2;
  P.tiles_x = cdiv(W1, kATPx); P.tiles_y = cdiv(H1, kATPy);
  static bool attr = false;
  if (!attr) {
    DB200_CUDA(cudaFuncSetAttribute(altcorr_tc_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, int(kATBwdSmem)));
    attr = true;
  }
  const long long tasks = static_cast<long long>(B) * P.tiles_x * P.tiles_y;
  altcorr_tc_bwd_kernel<<<int(tasks < kNumSMs ? tasks : kNumSMs), 256, kATBwdSmem, st>>>(P);
  DB200_CHECK_LAUNCH();
  return DB200_OK;
}

}  // namespace db200

using namespace db200;

static inline size_t al256(size_t x) { return (x + 255) / 256 * 256; }

// workspace of db200_altcorr_forward_ws: [flags 256 B][status + list][f1 hi][f1 lo][f2 hi][f2 lo]
extern "C" size_t db200_altcorr_workspace_bytes(int B, int S, int H1, int W1, int H2, int W2, int C) {
  if (B <= 0 || S <= 0 || H1 <= 0 || W1 <= 0 || H2 <= 0 || W2 <= 0 || C <= 0) return 0;
  return 256 + al256(altcorr_tc_status_bytes(B, S, H1, W1)) + 2 * al256(size_t(B) * H1 * W1 * C * 2) + 2 * al256(size_t(B) * H2 * W2 * C * 2);
}

extern "C" int db200_altcorr_forward_ws(const float* fmap1, const float* fmap2, const float* coords, float* corr, int B, int S,
                                        int H1, int W1, int H2, int W2, int C, int radius, void* workspace, size_t workspace_bytes,
                                        db200_stream_t stream) {
  if (!altcorr_tc_supported(C, radius) || !workspace)
    return db200_altcorr_forward(fmap1, fmap2, coords, corr, B, S, H1, W1, H2, W2, C, radius, stream);
  if (!fmap1 || !fmap2 || !coords || !corr || B < 0 || S < 0 || H1 <= 0 || W1 <= 0 || H2 <= 0 || W2 <= 0) return DB200_ERR_ARG;
  if (B == 0 || S == 0) return DB200_OK;
  if (workspace_bytes < db200_altcorr_workspace_bytes(B, S, H1, W1, H2, W2, C)) return DB200_ERR_WORKSPACE;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  char* ws = static_cast<char*>(workspace);
  unsigned* flags = reinterpret_cast<unsigned*>(ws); ws += 256;
  int* status = reinterpret_cast<int*>(ws); ws += al256(altcorr_tc_status_bytes(B, S, H1, W1));
  const size_t n1 = size_t(B) * H1 * W1 * C, n2 = size_t(B) * H2 * W2 * C;
  void* h1 = ws; ws += al256(n1 * 2);
  void* l1 = ws; ws += al256(n1 * 2);
  void* h2 = ws; ws += al256(n2 * 2);
  void* l2 = ws;
  DB200_CUDA(cudaMemsetAsync(flags, 0, sizeof(unsigned), st));
  int rc = altcorr_split(fmap1, h1, l1, n1, flags, st, size_t(H1) * W1);
  if (rc != DB200_OK) return rc;
  rc = altcorr_split(fmap2, h2, l2, n2, flags, st, size_t(H2) * W2);
  if (rc != DB200_OK) return rc;
  AltLaunch a{};
  a.f1_hi = h1; a.f1_lo = l1; a.f2_hi[0] = h2; a.f2_lo[0] = l2;
  a.f1_raw = fmap1; a.f2_raw[0] = fmap2; a.raw_is_half = 0;
  a.F1 = B; a.F2 = B; a.H1 = H1; a.W1 = W1; a.L = 1; a.H2[0] = H2; a.W2[0] = W2; a.scale[0] = 1.0f;
  a.idx1 = nullptr; a.idx2 = nullptr;
  a.coords = coords; a.c_sb = static_cast<long long>(S) * H1 * W1 * 2; a.c_ss = static_cast<long long>(H1) * W1 * 2; a.c_sp = 2;
  a.out = corr; a.o_sb = static_cast<long long>(S) * 49 * H1 * W1; a.o_ss = static_cast<long long>(49) * H1 * W1;
  a.o_sc = static_cast<long long>(H1) * W1; a.o_sp = 1;
  a.B = B; a.S = S; a.flags = flags; a.status = status;
  a.list_cap = 0;                                     // set by altcorr_tc_launch
  return altcorr_tc_launch(a, st);
}

extern "C" int db200_altcorr_backward_ws(const float* fmap1, const float* fmap2, const float* coords, const float* corr_grad,
                                         float* fmap1_grad, float* fmap2_grad, float* coords_grad, int B, int S, int H1, int W1,
                                         int H2, int W2, int C, int radius, void* workspace, size_t workspace_bytes,
                                         db200_stream_t stream) {
  if (!altcorr_tc_supported(C, radius) || !workspace)
    return db200_altcorr_backward(fmap1, fmap2, coords, corr_grad, fmap1_grad, fmap2_grad, coords_grad, B, S, H1, W1, H2, W2, C,
                                  radius, stream);
  if (!fmap1 || !fmap2 || !coords || !corr_grad || !fmap1_grad || !fmap2_grad || B < 0 || S < 0 || H1 <= 0 || W1 <= 0 || H2 <= 0 ||
      W2 <= 0)
    return DB200_ERR_ARG;
  if (B == 0) return DB200_OK;
  if (workspace_bytes < db200_altcorr_workspace_bytes(B, S > 0 ? S : 1, H1, W1, H2, W2, C)) return DB200_ERR_WORKSPACE;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  DB200_CUDA(cudaMemsetAsync(fmap2_grad, 0, sizeof(float) * size_t(B) * H2 * W2 * C, st));
  if (coords_grad && S > 0) DB200_CUDA(cudaMemsetAsync(coords_grad, 0, sizeof(float) * size_t(B) * S * H1 * W1 * 2, st));
  if (S == 0) { DB200_CUDA(cudaMemsetAsync(fmap1_grad, 0, sizeof(float) * size_t(B) * H1 * W1 * C, st)); return DB200_OK; }
  char* ws = static_cast<char*>(workspace);
  unsigned* flags = reinterpret_cast<unsigned*>(ws); ws += 256;
  int* status = reinterpret_cast<int*>(ws); ws += al256(altcorr_tc_status_bytes(B, S, H1, W1));
  const size_t n1 = size_t(B) * H1 * W1 * C, n2 = size_t(B) * H2 * W2 * C;
  void* h1 = ws; ws += al256(n1 * 2);
  void* l1 = ws; ws += al256(n1 * 2);
  void* h2 = ws; ws += al256(n2 * 2);
  void* l2 = ws;
  DB200_CUDA(cudaMemsetAsync(flags, 0, sizeof(unsigned), st));
  DB200_CUDA(cudaMemsetAsync(status, 0, 2 * sizeof(int), st));
  int rc = altcorr_split(fmap1, h1, l1, n1, flags, st, size_t(H1) * W1);
  if (rc != DB200_OK) return rc;
  rc = altcorr_split(fmap2, h2, l2, n2, flags, st, size_t(H2) * W2);
  if (rc != DB200_OK) return rc;
  rc = altcorr_tc_bwd_launch(h1, l1, h2, l2, coords, corr_grad, fmap1_grad, fmap2_grad, B, S, H1, W1, H2, W2, flags, status, st);
  if (rc != DB200_OK) return rc;
  return altcorr_bwd_simt_conditional(fmap1, fmap2, coords, corr_grad, fmap1_grad, fmap2_grad, B, S, H1, W1, H2, W2, C, radius, flags,
                                      status, st);
}

extern "C" size_t db200_altcorr_pyramid_workspace_bytes(int B, int S, int H1, int W1) {
  if (B <= 0 || S <= 0 || H1 <= 0 || W1 <= 0) return 0;
  return al256(altcorr_tc_status_bytes(B, S, H1, W1));
}

extern "C" int db200_altcorr_pyramid_forward(const void* fmap1_hi, const void* fmap1_lo, const void* const* fmap2_hi,
                                             const void* const* fmap2_lo, const unsigned* flags, const float* fmap1_f32,
                                             const float* const* fmap2_f32, long long frames, int H, int W, int C,
                                             int num_levels, const int64_t* idx1, const int64_t* idx2, const float* coords,
                                             long long c_sb, long long c_ss, long long c_sp, float* out, long long o_sb,
                                             long long o_ss, long long o_sc, long long o_sp, int B, int S, int radius,
                                             void* workspace, size_t workspace_bytes, db200_stream_t stream) {
  if (!fmap1_hi || !fmap2_hi || !coords || !out || frames <= 0 || H <= 0 || W <= 0 || num_levels < 1 || num_levels > kATMaxL ||
      B < 0 || S < 0)
    return DB200_ERR_ARG;
  if (!altcorr_tc_supported(C, radius)) return DB200_ERR_ARG;
  if (B == 0 || S == 0) return DB200_OK;
  if (!workspace || workspace_bytes < db200_altcorr_pyramid_workspace_bytes(B, S, H, W)) return DB200_ERR_WORKSPACE;
  const bool has_lo = fmap1_lo != nullptr;
  if (has_lo != (fmap2_lo != nullptr) || (has_lo && (!flags || !fmap1_f32 || !fmap2_f32))) return DB200_ERR_ARG;
  AltLaunch a{};
  a.f1_hi = fmap1_hi; a.f1_lo = fmap1_lo;
  a.f1_raw = has_lo ? static_cast<const void*>(fmap1_f32) : fmap1_hi; a.raw_is_half = has_lo ? 0 : 1;
  for (int l = 0; l < num_levels; l++) {
    if (!fmap2_hi[l] || (H >> l) <= 0 || (W >> l) <= 0 || (has_lo && (!fmap2_lo[l] || !fmap2_f32[l]))) return DB200_ERR_ARG;
    a.f2_hi[l] = fmap2_hi[l]; a.f2_lo[l] = has_lo ? fmap2_lo[l] : nullptr;
    a.f2_raw[l] = has_lo ? static_cast<const void*>(fmap2_f32[l]) : fmap2_hi[l];
    a.H2[l] = H >> l; a.W2[l] = W >> l; a.scale[l] = 1.0f / float(1 << l);
  }
  a.F1 = frames; a.F2 = frames; a.H1 = H; a.W1 = W; a.L = num_levels;
  a.idx1 = reinterpret_cast<const long long*>(idx1); a.idx2 = reinterpret_cast<const long long*>(idx2);
  a.coords = coords; a.c_sb = c_sb; a.c_ss = c_ss; a.c_sp = c_sp;
  a.out = out; a.o_sb = o_sb; a.o_ss = o_ss; a.o_sc = o_sc; a.o_sp = o_sp;
  a.B = B; a.S = S; a.flags = has_lo ? const_cast<unsigned*>(flags) : nullptr; a.status = static_cast<int*>(workspace);
  a.list_cap = 0;
  return altcorr_tc_launch(a, static_cast<cudaStream_t>(stream));
}

extern "C" int db200_split_hilo_f16(const float* x, void* hi, void* lo, size_t n, size_t hw, unsigned* flags, db200_stream_t stream) {
  if (!x || !hi || !lo || !flags) return DB200_ERR_ARG;
  if (n == 0) return DB200_OK;
  return altcorr_split(x, hi, lo, n, flags, static_cast<cudaStream_t>(stream), hw);
}

extern "C" int db200_debug_altcorr_waits(const void* status_words, long long* out12, db200_stream_t stream) {
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  DB200_CUDA(cudaMemcpyAsync(out12, static_cast<const char*>(status_words) + 32, 12 * sizeof(long long), cudaMemcpyDeviceToHost, st));
  DB200_CUDA(cudaStreamSynchronize(st));
  return DB200_OK;
}

extern "C" long long db200_altcorr_cells(const void* status_words, db200_stream_t stream) {
  int h[4] = {0, 0, 0, 0};
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (cudaMemcpyAsync(h, status_words, sizeof(h), cudaMemcpyDeviceToHost, st) != cudaSuccess || cudaStreamSynchronize(st) != cudaSuccess)
    return -1;
  return h[2];
}

extern "C" int db200_altcorr_status(const void* status_words, db200_stream_t stream) {
  int h[2] = {0, 0};
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  DB200_CUDA(cudaMemcpyAsync(h, status_words, sizeof(h), cudaMemcpyDeviceToHost, st));
  DB200_CUDA(cudaStreamSynchronize(st));
  return h[0] != 0 ? DB200_ERR_CUDA : h[1];
}

// ======================================================================================================================
// db200_altcorr_forward_static — altcorr_forward for callers that cannot allocate (CUDA graph capture, tight serving
// loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
// Same kernels as db200_altcorr_forward_ws; the scratch (fp16 hi/lo split of both feature maps, flag and status words,
// hand-over list) lives in ONE file-scope device array of kAltStaticBytes that every GPU that loads the library holds until
// exit.  All calls share it: two calls that overlap on the GPU (different streams, or graph branches) race on it.
// Nothing in it has to survive between calls; the words that must start at zero (flags, status) are cleared on `stream`
// at the start of every call.
constexpr size_t kAltStaticBytes = size_t(8) << 20;                   // 8 MiB: e.g. B*S = 2 edges at 48x64 -> 48x64, C = 128
static __device__ __align__(256) unsigned char g_altcorr_static_scratch[kAltStaticBytes];

namespace {
// In host code `g_altcorr_static_scratch` is a host-side stand-in, not a device address: resolve it per device, once,
// with that device current.
constexpr int kMaxDevices = 64;
void* g_alt_static_ptr[kMaxDevices] = {};
int alt_static_scratch(void** out) {
  int dev = -1;
  DB200_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kMaxDevices) return DB200_ERR_ARG;
  if (!g_alt_static_ptr[dev]) {
    void* p = nullptr;
    DB200_CUDA(cudaGetSymbolAddress(&p, g_altcorr_static_scratch));
    if (!p) return DB200_ERR_CUDA;
    g_alt_static_ptr[dev] = p;                                       // idempotent: racing host threads store the same value
  }
  *out = g_alt_static_ptr[dev];
  return DB200_OK;
}
}  // namespace

extern "C" size_t db200_altcorr_static_max_bytes(void) { return kAltStaticBytes; }

extern "C" int db200_altcorr_forward_static(const float* fmap1, const float* fmap2, const float* coords, float* corr, int B,
                                            int S, int H1, int W1, int H2, int W2, int C, int radius, db200_stream_t stream) {
  if (!fmap1 || !fmap2 || !coords || !corr || B < 0 || S < 0 || H1 <= 0 || W1 <= 0 || H2 <= 0 || W2 <= 0 || C <= 0) return DB200_ERR_ARG;
  if (!altcorr_tc_supported(C, radius)) return DB200_ERR_ARG;         // the scratch-free SIMT entry point serves other shapes
  if (B == 0 || S == 0) return DB200_OK;
  const size_t need = db200_altcorr_workspace_bytes(B, S, H1, W1, H2, W2, C);
  if (need == 0 || need > kAltStaticBytes) return DB200_ERR_STATIC_LIMIT;
  void* scratch = nullptr;
  const int rc = alt_static_scratch(&scratch);
  if (rc != DB200_OK) return rc;
  return db200_altcorr_forward_ws(fmap1, fmap2, coords, corr, B, S, H1, W1, H2, W2, C, radius, scratch, kAltStaticBytes, stream);
}
