// corr_volume_tc.cu — all-pairs 128-d correlation on the 5th-generation tensor cores (fp16 inference path).
// Replaces CorrBlock.corr's torch.matmul + the three avg_pool2d passes of
// /root/reference/droid_slam/modules/corr.py:30-38,63-71 for f16 feature maps.
//
//   corr_l[n, p1, p2] = (1/16) sum_c f1[n,c,p1] * pool_l(f2)[n,c,p2]          (pooling commutes with the contraction)
//
// prep kernel : feature maps [N][C][HW] -> channels-last K-major copies f1T[N*HW][C], f2T_l[N*HW_l][C] (2x2 mean per level)
// GEMM kernel : one CTA = one 128x128 output tile of one edge, 128 threads.
//   thread 0      : two TMA loads per operand (cp.async.bulk.tensor.2d, 128B swizzle, box 64 x 128: the whole K = 128
//                   of the tile is resident) completing on one mbarrier;
//   thread 32     : 8 x tcgen05.mma.cta_group::1.kind::f16 (M=128, N=128, K=16), fp32 accumulator in TMEM (128 columns),
//                   tcgen05.commit -> mbarrier;
//   all 4 warps   : tcgen05.ld 32x32b (warp w owns TMEM lanes 32w..32w+31 = output rows), x 1/16, cvt to f16, 64-byte
//                   row segments straight to HBM.
// 3 CTAs are resident per SM (64 KB smem, 128 of 512 TMEM columns each), so the TMA / MMA / store phases of different tiles
// overlap without an in-kernel pipeline.  The op is output-write bound (25 MB written per edge, 2.4 GFLOP).
// Every mbarrier wait is bounded: a broken descriptor raises DB200_ERR_CUDA through the status word instead of hanging.
#include <cstdlib>
#include "tc_common.cuh"

namespace db200 {

static inline size_t align_up256(size_t x) { return (x + 255) / 256 * 256; }

constexpr int kTM = 128, kTN = 128, kTK = 64;            // tile, and K per TMA box (64 halfs = one 128-byte swizzle row)

// ---- prep: [N][C][hs*ws] -> [N][(hs>>l)*(ws>>l)][C] with 2^l x 2^l mean (l = 0: plain transpose), fp32 mean rounded once per level
// (level l+1 is pooled from the ROUNDED level l, like F.avg_pool2d applied to an f16 tensor)
// 64 pixels x 64 channels per CTA: 128-byte reads along the pixels (half2 per lane), 128-byte writes along the channels.
// wt > 0: destination rows in the tiled order of a [HW / wt][wt] map (common.cuh tiled_index) — the GEMM then writes its output
// columns, i.e. the [h2][w2] slice of every source pixel, in that order with no other change.
// (the 32x32-element version moved 64-byte segments and took 47 us per 48-edge map: 1.6 TB/s)
__global__ void __launch_bounds__(256) transpose_cl_kernel(const __half* __restrict__ src, __half* __restrict__ dst, int C, int HW, int wt = 0) {
  __shared__ __half tileT[64][66];                                     // [pixel][channel], pitch 33 words
  const int n = blockIdx.z, c0 = blockIdx.y * 64, p0 = blockIdx.x * 64;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const bool pair_ok = (HW % 2 == 0);                                  // half2 reads need an even row pitch
  for (int r = warp; r < 64; r += 8) {
    const int c = c0 + r, p = p0 + 2 * lane;
    __half a = __float2half(0.0f), b = a;
    if (c < C) {
      const __half* sp = src + (size_t(n) * C + c) * HW + p;
      if (pair_ok && p + 1 < HW) { const __half2 v = *reinterpret_cast<const __half2*>(sp); a = __low2half(v); b = __high2half(v); }
      else { if (p < HW) a = sp[0]; if (p + 1 < HW) b = sp[1]; }
    }
    tileT[2 * lane][r] = a;
    tileT[2 * lane + 1][r] = b;
  }
  __syncthreads();
  const bool cpair_ok = (C % 2 == 0);
  for (int r = warp; r < 64; r += 8) {
    const int p = p0 + r, c = c0 + 2 * lane;
    if (p >= HW) continue;
    const size_t drow = size_t(n) * HW + (wt > 0 ? tiled_index(p / wt, p % wt, wt) : p);
    const __half2 v = *reinterpret_cast<const __half2*>(&tileT[r][2 * lane]);
    __half* dp = dst + drow * C + c;
    if (cpair_ok && c + 1 < C) *reinterpret_cast<__half2*>(dp) = v;
    else { if (c < C) dp[0] = __low2half(v); if (c + 1 < C) dp[1] = __high2half(v); }
  }
}

// channels-last pooling: dst[n][y][x][c] = mean of the 2x2 block of src (floor mode)
// (src_tiled / dst_tiled: that map's rows are in tiled order)
__global__ void pool_cl_kernel(const __half* __restrict__ src, __half* __restrict__ dst, int N, int hs, int ws, int C,
                               bool src_tiled = false, bool dst_tiled = false) {
  const int hd = hs >> 1, wd = ws >> 1;
  const size_t total = size_t(N) * hd * wd * (C / 2);
  for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += size_t(gridDim.x) * blockDim.x) {
    const int c2 = int(i % (C / 2));
    size_t r = i / (C / 2);
    const int x = int(r % wd); r /= wd;
    const int y = int(r % hd);
    const size_t n = r / hd;
    float2 a, b, c, d;
    if (src_tiled) {
      const __half2* sn = reinterpret_cast<const __half2*>(src + n * hs * ws * size_t(C)) + c2;
      a = __half22float2(sn[size_t(tiled_index(2 * y, 2 * x, ws)) * (C / 2)]);
      b = __half22float2(sn[size_t(tiled_index(2 * y, 2 * x + 1, ws)) * (C / 2)]);
      c = __half22float2(sn[size_t(tiled_index(2 * y + 1, 2 * x, ws)) * (C / 2)]);
      d = __half22float2(sn[size_t(tiled_index(2 * y + 1, 2 * x + 1, ws)) * (C / 2)]);
    } else {
      const __half2* s = reinterpret_cast<const __half2*>(src + ((n * hs + 2 * y) * ws + 2 * x) * size_t(C)) + c2;
      a = __half22float2(s[0]); b = __half22float2(s[C / 2]); c = __half22float2(s[size_t(ws) * (C / 2)]);
      d = __half22float2(s[size_t(ws) * (C / 2) + C / 2]);
    }
    const __half2 v = __floats2half2_rn((a.x + b.x + c.x + d.x) * 0.25f, (a.y + b.y + c.y + d.y) * 0.25f);
    const size_t o = dst_tiled ? ((n * hd * wd + tiled_index(y, x, wd)) * size_t(C / 2) + c2) : i;
    reinterpret_cast<__half2*>(dst)[o] = v;
  }
}

// instruction descriptor: D = F32 (bits 4-5 = 1), A = B = F16 (0), both K-major, N >> 3 at bit 17, M >> 4 at bit 24
constexpr uint32_t kIdescF16 = (1u << 4) | (uint32_t(kTN >> 3) << 17) | (uint32_t(kTM >> 4) << 24);

// grid (ceil(N2/128), ceil(M/128), edges); 128 threads; dynamic smem 64 KB + 1 KB alignment slack
__global__ void __launch_bounds__(128)
corr_gemm_tc_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, __half* __restrict__ out,
                    int M, int N2, int* __restrict__ err) {
  extern __shared__ unsigned char smem_dyn[];
  __shared__ __align__(8) uint64_t bars[2];
  __shared__ uint32_t tmem_base_s;
  const uint32_t s0 = (smem_u32(smem_dyn) + 1023u) & ~1023u;          // 1024-byte aligned: required by the 128B swizzle
  const uint32_t sA = s0, sB = s0 + 2 * kTM * kTK * 2;                 // A: 2 k-blocks of [128][64] halfs, then B likewise
  const uint32_t full_bar = smem_u32(&bars[0]), mma_bar = smem_u32(&bars[1]);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int n0 = blockIdx.x * kTN, m0 = blockIdx.y * kTM, e = blockIdx.z;

  if (tid == 0) {
    mbar_init(full_bar, 1);
    mbar_init(mma_bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {                                                     // TMEM: 128 columns x 128 lanes of fp32
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_s)), "n"(kTN) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem_base = tmem_base_s;

  if (tid == 0) {
    mbar_expect_tx(full_bar, 2u * (kTM + kTN) * kTK * 2u);
    tma_load_2d(sA, &tmA, 0, e * M + m0, full_bar);
    tma_load_2d(sA + kTM * kTK * 2, &tmA, kTK, e * M + m0, full_bar);
    tma_load_2d(sB, &tmB, 0, e * N2 + n0, full_bar);
    tma_load_2d(sB + kTN * kTK * 2, &tmB, kTK, e * N2 + n0, full_bar);
  }
  if (tid == 32) {
    if (mbar_wait(full_bar, 0)) {
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
#pragma unroll
      for (int kb = 0; kb < 2; kb++)
#pragma unroll
        for (int k = 0; k < kTK / 16; k++) {
          const uint64_t ad = umma_desc_sw128(sA + kb * kTM * kTK * 2 + k * 32);
          const uint64_t bd = umma_desc_sw128(sB + kb * kTN * kTK * 2 + k * 32);
          umma_f16(tmem_base, ad, bd, kIdescF16, (kb | k) != 0 ? 1u : 0u);
        }
    } else {
      atomicExch(err, 1);
    }
    umma_commit(mma_bar);                                             // arrives when the MMAs above have completed
  }
  __syncwarp();
  const bool ok = mbar_wait(mma_bar, 0);
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  if (!ok && lane == 0) atomicExch(err, 1);

  // ---- epilogue: TMEM lane = output row, 32 columns per load
  const int row = m0 + 32 * warp + lane;
  __half* orow = out + (size_t(e) * M + (row < M ? row : 0)) * N2 + n0;
  const bool vec = (N2 % 8 == 0);
#pragma unroll 1
  for (int cb = 0; cb < kTN; cb += 32) {
    uint32_t v[32];
    tmem_ld32(tmem_base + (uint32_t(32 * warp) << 16) + uint32_t(cb), v);
    if (ok && row < M) {
      if (vec && n0 + cb + 32 <= N2) {
        uint4 pk[4];
        uint32_t* pw = reinterpret_cast<uint32_t*>(pk);
#pragma unroll
        for (int i = 0; i < 16; i++) {
          const __half2 h = __floats2half2_rn(__uint_as_float(v[2 * i]) * 0.0625f, __uint_as_float(v[2 * i + 1]) * 0.0625f);
          pw[i] = *reinterpret_cast<const uint32_t*>(&h);
        }
        uint4* dstp = reinterpret_cast<uint4*>(orow + cb);
#pragma unroll
        for (int i = 0; i < 4; i++) dstp[i] = pk[i];
      } else {
#pragma unroll
        for (int i = 0; i < 32; i++)
          if (n0 + cb + i < N2) orow[cb + i] = __float2half_rn(__uint_as_float(v[i]) * 0.0625f);
      }
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 1) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(kTN) : "memory");
}

// ---- persistent, warp-specialised pipeline ---------------------------------------------------------------------
// One CTA per SM walks the tiles (n-tile fastest, so consecutive tiles share the A operand in L2):
//   warp 0 (one lane) : TMA producer — two shared-memory stages of [A 128x128 | B 128x128] halfs (64 KB each);
//   warp 1 (one lane) : issues the 16 tcgen05.mma of a tile into one of TWO 128-column TMEM accumulators, commits the
//                       stage back to the producer and the accumulator to the epilogue;
//   warps 2..5        : epilogue — tcgen05.ld of their TMEM lane quarter, x 1/16, f16, row segments to HBM, then hand
//                       the accumulator back.  The stores of tile k overlap the MMAs of tile k+1 and the loads of k+2.
// mbarrier phases: stage s = k & 1 is in its (k >> 1)-th use; "empty"-type barriers are waited with the inverted
// parity so that the first use passes.  Every wait is bounded (status word instead of a hang).
constexpr int kPipeThreadsTC = 192;
// two operand stages (64 KB each) + two output staging tiles (32 KB each) + alignment slack
constexpr size_t kPipeSmemTC = 2 * (2 * (kTM + kTN) * kTK * 2) + 2 * (kTM * kTN * 2) + 1024;
// FUSED: staging of the pooled levels for their TMA stores, two buffers each: level 1 [128 rows][2 tiles][16 halfs], level 2
// [128 rows][8 halfs]
constexpr uint32_t kFusedL1Stage = kTM * 64, kFusedL2Stage = kTM * 16;
constexpr size_t kPipeSmemFused = kPipeSmemTC + 2 * (kFusedL1Stage + kFusedL2Stage);

// MN-major operand tile, 128-byte swizzle: blocks of 64 MN-elements x K rows of 128 bytes; LBO = next 64-element MN block
// (16 KB: the second TMA box), SBO = next 8 K-rows (1 KB).  Used for fmap1 as it lies in memory ([C][HW], pixels contiguous).
__device__ __forceinline__ uint64_t umma_desc_sw128_mn(uint32_t saddr) {
  return uint64_t((saddr >> 4) & 0x3FFFu) | (uint64_t(16384 >> 4) << 16) | (uint64_t(1024 >> 4) << 32) | (uint64_t(1) << 46) | (uint64_t(2) << 61);
}

template <int THREADS>
__device__ __forceinline__ void epi_bar() { asm volatile("bar.sync 1, %0;" ::"n"(THREADS) : "memory"); }   // the epilogue warps

// TMA_STORE: the f16 tile is staged in shared memory in the 128-byte-swizzled box layout (bank-conflict free for
// thread = row) and leaves as two asynchronous bulk tensor stores (full lines, clipped per edge by the 3-D tensor map);
// otherwise every thread writes its row segments directly (any N2).
//
// FUSED (CorrBlock's private pyramid, 64-wide maps, level 0 and 1 tiled): the pooled levels come out of the SAME launch.  The
// columns of the product are in tiled order, so the 32 accumulator columns of one tcgen05.ld are one 4x8-pixel tile of the
// (h2, w2) map: its 2x2 / 4x4 means are level-1 / level-2 values, complete in the thread's registers; level 3 (8x8) needs
// the tile row below, which is two n-tiles later in the SAME CTA — a CTA takes the n-tiles in groups of four (= two tile
// rows), carrying eight partial sums.  The pooled values are means of the fp32 accumulators, rounded to fp16 once (the
// unfused path rounds pooled fp16 operands first; the reference rounds at every level: all within fp16 rounding of the
// exact mean).  No pooled operand copies, no further GEMM launches, fmap1 is read once.
struct FusedOut {
  __half* l1; __half* l2; __half* l3;     // level 1 (tiled slices), level 2 and 3 (row-major slices); NULL = level not wanted
};

template <bool TMA_STORE, bool FUSED>
__global__ void __launch_bounds__(kPipeThreadsTC, 1)
corr_gemm_tc_pipe_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                         const __grid_constant__ CUtensorMap tmO, __half* __restrict__ out, int M, int N2, int E,
                         int* __restrict__ err, FusedOut fo, const __grid_constant__ CUtensorMap tmL1,
                         const __grid_constant__ CUtensorMap tmL2) {
  extern __shared__ unsigned char smem_dyn[];
  __shared__ __align__(8) uint64_t bars[8];          // full[2], empty[2], tfull[2], tempty[2]
  __shared__ uint32_t tmem_base_s;
  constexpr uint32_t kStageBytes = 2u * (kTM + kTN) * kTK * 2u;        // 64 KB
  const uint32_t s0 = (smem_u32(smem_dyn) + 1023u) & ~1023u;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int MT = (M + kTM - 1) / kTM, NT = (N2 + kTN - 1) / kTN;
  const long long tiles = static_cast<long long>(E) * MT * NT;
  // tile sequence of this CTA: plain = blockIdx.x, + gridDim.x, ...; FUSED = units of 4 consecutive n-tiles (NT % 4 == 0, so a
  // unit stays inside one (edge, m-tile)): unit u -> tiles 4u .. 4u+3
  auto tile_at = [&](long long k) -> long long {
    if (FUSED) return 4 * (blockIdx.x + (k >> 2) * static_cast<long long>(gridDim.x)) + (k & 3);
    return blockIdx.x + k * static_cast<long long>(gridDim.x);
  };
  auto bar = [&](int i) { return smem_u32(&bars[i]); };

  if (tid == 0) {
    for (int i = 0; i < 6; i++) mbar_init(bar(i), 1);
    mbar_init(bar(6), 4); mbar_init(bar(7), 4);                        // four epilogue warps release an accumulator
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_s)), "n"(2 * kTN) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem_base = tmem_base_s;

  if (warp == 0) {
    if (lane == 0) {
      int k = 0;
      for (long long t = tile_at(0); t < tiles; t = tile_at(++k)) {
        const int s = k & 1, use = k >> 1;
        const int nt = int(t % NT), mt = int((t / NT) % MT), e = int(t / (static_cast<long long>(NT) * MT));
        if (!mbar_wait(bar(2 + s), (use & 1) ^ 1)) { atomicExch(err, 1); break; }
        const uint32_t sA = s0 + s * kStageBytes, sB = sA + 2 * kTM * kTK * 2;
        mbar_expect_tx(bar(s), kStageBytes);
        if constexpr (FUSED) {                                         // fmap1 untransposed: [64 pixels][128 channels] per box
          tma_load_2d(sA, &tmA, mt * kTM, e * 2 * kTK, bar(s));
          tma_load_2d(sA + kTM * kTK * 2, &tmA, mt * kTM + 64, e * 2 * kTK, bar(s));
        } else {
          tma_load_2d(sA, &tmA, 0, e * M + mt * kTM, bar(s));
          tma_load_2d(sA + kTM * kTK * 2, &tmA, kTK, e * M + mt * kTM, bar(s));
        }
        tma_load_2d(sB, &tmB, 0, e * N2 + nt * kTN, bar(s));
        tma_load_2d(sB + kTN * kTK * 2, &tmB, kTK, e * N2 + nt * kTN, bar(s));
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      int k = 0;
      for (long long t = tile_at(0); t < tiles; t = tile_at(++k)) {
        const int s = k & 1, use = k >> 1;
        if (!mbar_wait(bar(6 + s), (use & 1) ^ 1)) { atomicExch(err, 1); break; }     // accumulator s drained
        if (!mbar_wait(bar(s), use & 1)) { atomicExch(err, 1); break; }               // operands landed
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        const uint32_t sA = s0 + s * kStageBytes, sB = sA + 2 * kTM * kTK * 2;
#pragma unroll
        for (int kb = 0; kb < 2; kb++)
#pragma unroll
          for (int kk = 0; kk < kTK / 16; kk++) {
            const uint64_t ad = FUSED ? umma_desc_sw128_mn(sA + (kb * (kTK / 16) + kk) * 2048) : umma_desc_sw128(sA + kb * kTM * kTK * 2 + kk * 32);
            const uint64_t bd = umma_desc_sw128(sB + kb * kTN * kTK * 2 + kk * 32);
            umma_f16(tmem_base + uint32_t(s * kTN), ad, bd, FUSED ? (kIdescF16 | (1u << 15)) : kIdescF16, (kb | kk) != 0 ? 1u : 0u);
          }
        umma_commit(bar(2 + s));                                       // stage free once these MMAs have read it
        umma_commit(bar(4 + s));                                       // accumulator ready for the epilogue
      }
    }
  } else {
    const int quarter = warp & 3;                                      // TMEM lane quarter this warp may read
    const bool vec = (N2 % 8 == 0);
    constexpr int kEpiThreads = 128;
    float p3[8], f3[4];                                                // FUSED: level-3 partial sums of the upper tile row, finished values
    uint32_t w1[2][4], w2[4];                                          // FUSED: packed level-1 / level-2 values of the running tile
    int k = 0;
    for (long long t = tile_at(0); t < tiles; t = tile_at(++k)) {
      const int s = k & 1, use = k >> 1;
      const int nt = int(t % NT), mt = int((t / NT) % MT), e = int(t / (static_cast<long long>(NT) * MT));
      const bool ok = mbar_wait(bar(4 + s), use & 1);
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      if (!ok) { if (lane == 0) atomicExch(err, 1); break; }
      const int n0 = nt * kTN, row = mt * kTM + 32 * quarter + lane;
      if constexpr (TMA_STORE) {
        const int trow = 32 * quarter + lane;                          // row inside the tile
        const uint32_t stg = s0 + 2 * kStageBytes + uint32_t(s) * (kTM * kTN * 2);
        const uint32_t l1stg = s0 + 2 * kStageBytes + 2 * (kTM * kTN * 2) + uint32_t(s) * kFusedL1Stage;   // FUSED only
        const uint32_t l2stg = s0 + 2 * kStageBytes + 2 * (kTM * kTN * 2) + 2 * kFusedL1Stage + uint32_t(s) * kFusedL2Stage;
        if (tid == 64) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");   // the store that last read this buffer
        epi_bar<kEpiThreads>();
#pragma unroll(FUSED ? 4 : 1)
        for (int cb = 0; cb < kTN; cb += 32) {
          uint32_t v[32];
          tmem_ld32(tmem_base + (uint32_t(32 * quarter) << 16) + uint32_t(s * kTN + cb), v);
          if constexpr (FUSED) {
            // v = one 4x8 tile (ty, tx) of the (h2, w2) map, v[8 * yi + xi]; ty = nt >> 1, tx = 4 * (nt & 1) + cb / 32.
            // Levels 1 and 2 are staged in shared memory (thread = volume row) and leave with the tile's TMA stores: direct
            // stores would be 32 different slices per warp instruction, and the L1/TEX path of ~10 such stores per tile and
            // thread set the tile period (68 % L1/TEX, 308 us per 48 edges instead of the 202 us of level 0 alone).
            const int ty = nt >> 1, q = cb >> 5;                       // q: static after unrolling
            float s1[2][4];
#pragma unroll
            for (int r = 0; r < 2; r++)
#pragma unroll
              for (int c = 0; c < 4; c++)
                s1[r][c] = (__uint_as_float(v[16 * r + 2 * c]) + __uint_as_float(v[16 * r + 2 * c + 1])) +
                           (__uint_as_float(v[16 * r + 8 + 2 * c]) + __uint_as_float(v[16 * r + 8 + 2 * c + 1]));
            const float s2a = (s1[0][0] + s1[0][1]) + (s1[1][0] + s1[1][1]), s2b = (s1[0][2] + s1[0][3]) + (s1[1][2] + s1[1][3]);
            const float s3 = s2a + s2b;
            auto pk = [](float a, float b) { const __half2 h = __floats2half2_rn(a, b); return *reinterpret_cast<const uint32_t*>(&h); };
#pragma unroll
            for (int r = 0; r < 2; r++) {                              // level 1: rows 2ty + r, columns 4tx .. 4tx + 3
              w1[r][2 * (q & 1)] = pk(s1[r][0] * 0.015625f, s1[r][1] * 0.015625f);
              w1[r][2 * (q & 1) + 1] = pk(s1[r][2] * 0.015625f, s1[r][3] * 0.015625f);
            }
            w2[q] = pk(s2a * 0.00390625f, s2b * 0.00390625f);          // level 2: row ty, columns 2tx, 2tx + 1
            if (q & 1) {                                               // chunks (q-1, q) = the row pair of one 4x8 level-1 tile: 32 bytes
              const uint32_t a = l1stg + uint32_t(trow) * 64 + uint32_t(q >> 1) * 32;
              asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(a), "r"(w1[0][0]), "r"(w1[0][1]), "r"(w1[0][2]), "r"(w1[0][3]) : "memory");
              asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(a + 16), "r"(w1[1][0]), "r"(w1[1][1]), "r"(w1[1][2]), "r"(w1[1][3]) : "memory");
            }
            if (q == 3)                                                // 8 level-2 columns of row ty: 16 bytes
              asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(l2stg + uint32_t(trow) * 16), "r"(w2[0]), "r"(w2[1]), "r"(w2[2]), "r"(w2[3]) : "memory");
            if ((ty & 1) == 0) {                                       // level 3 = 8x8 means: upper tile row kept, lower one completes it
              if (nt & 1) p3[4 + q] = s3; else p3[q] = s3;             // static register indices
            } else {
              f3[q] = (((nt & 1) ? p3[4 + q] : p3[q]) + s3) * 0.0009765625f;
              if (q == 3 && fo.l3 && row < M)                          // 4 level-3 columns of row ty >> 1 (1/64 of the bytes: direct store)
                *reinterpret_cast<uint2*>(fo.l3 + (size_t(e) * M + row) * size_t(N2 >> 6) + (ty >> 1) * 8 + 4 * (nt & 1)) = make_uint2(pk(f3[0], f3[1]), pk(f3[2], f3[3]));
            }
          }
          const uint32_t box = stg + uint32_t(cb >> 6) * (kTM * 128) + uint32_t(trow) * 128;
          const int ch0 = (cb & 63) >> 3;
#pragma unroll
          for (int c4 = 0; c4 < 4; c4++) {
            uint32_t w[4];
#pragma unroll
            for (int i = 0; i < 4; i++) {
              const __half2 h = __floats2half2_rn(__uint_as_float(v[8 * c4 + 2 * i]) * 0.0625f, __uint_as_float(v[8 * c4 + 2 * i + 1]) * 0.0625f);
              w[i] = *reinterpret_cast<const uint32_t*>(&h);
            }
            const uint32_t a = box + uint32_t(((ch0 + c4) ^ (trow & 7)) << 4);
            asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(a), "r"(w[0]), "r"(w[1]), "r"(w[2]), "r"(w[3]) : "memory");
          }
        }
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
        __syncwarp();
        if (lane == 0) mbar_arrive(bar(6 + s));                        // accumulator drained: the next MMAs may start
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic-proxy writes -> visible to the TMA
        epi_bar<kEpiThreads>();
        if (tid == 64) {
          tma_store_3d(&tmO, stg, n0, mt * kTM, e);
          if (n0 + 64 < N2) tma_store_3d(&tmO, stg + kTM * 128, n0 + 64, mt * kTM, e);
          if constexpr (FUSED) {                                       // same bulk group: the staging buffers are reused together
            const int ty = nt >> 1;
            // level 1 as [16 halfs = 2 tile rows][row pair][tile column][tile row][volume row]; level 2 as [16][12][volume row]
            if (fo.l1) tma_store_5d(&tmL1, l1stg, 0, ty & 1, 2 * (nt & 1), ty >> 1, e * M + mt * kTM);
            if (fo.l2) tma_store_3d(&tmL2, l2stg, 8 * (nt & 1), ty, e * M + mt * kTM);
          }
          asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        }
        continue;
      }
      __half* orow = out + (size_t(e) * M + (row < M ? row : 0)) * N2 + n0;
#pragma unroll 1
      for (int cb = 0; cb < kTN; cb += 32) {
        uint32_t v[32];
        tmem_ld32(tmem_base + (uint32_t(32 * quarter) << 16) + uint32_t(s * kTN + cb), v);
        if (row < M) {
          if (vec && n0 + cb + 32 <= N2) {
            uint4 pk[4];
            uint32_t* pw = reinterpret_cast<uint32_t*>(pk);
#pragma unroll
            for (int i = 0; i < 16; i++) {
              const __half2 h = __floats2half2_rn(__uint_as_float(v[2 * i]) * 0.0625f, __uint_as_float(v[2 * i + 1]) * 0.0625f);
              pw[i] = *reinterpret_cast<const uint32_t*>(&h);
            }
            uint4* dstp = reinterpret_cast<uint4*>(orow + cb);
#pragma unroll
            for (int i = 0; i < 4; i++) dstp[i] = pk[i];
          } else {
#pragma unroll
            for (int i = 0; i < 32; i++)
              if (n0 + cb + i < N2) orow[cb + i] = __float2half_rn(__uint_as_float(v[i]) * 0.0625f);
          }
        }
      }
      asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
      __syncwarp();
      if (lane == 0) mbar_arrive(bar(6 + s));
    }
    if (TMA_STORE && tid == 64) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 1) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(2 * kTN) : "memory");
}

// ---- fp32 volumes (training path) on the tensor cores: 3xTF32 ----------------------------------------------------
// a = a_hi + a_lo with a_hi = tf32(a), a_lo = tf32(a - a_hi):  a.b ~= a_hi.b_hi + a_hi.b_lo + a_lo.b_hi (the dropped
// lo.lo term is ~2^-22 relative), accumulated in fp32 in TMEM — fp32-faithful (|err| ~ 1e-6 of the fp32 SGEMM the
// reference calls), at tensor-core rate.  Operands: channels-last fp32 copies split by the prep kernels; the GEMM is
// a K = 3 x 128 chain streamed in 12 blocks of 32 (one 128-byte swizzle row) through a 4-stage TMA ring.
__global__ void __launch_bounds__(256) transpose_cl_f32_kernel(const float* __restrict__ src, float* __restrict__ dst, int C, int HW) {
  __shared__ float tile[32][33];
  const int n = blockIdx.z, c0 = blockIdx.y * 32, p0 = blockIdx.x * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  for (int r = ty; r < 32; r += 8) {
    const int c = c0 + r, p = p0 + tx;
    tile[r][tx] = (c < C && p < HW) ? src[(size_t(n) * C + c) * HW + p] : 0.0f;
  }
  __syncthreads();
  for (int r = ty; r < 32; r += 8) {
    const int p = p0 + r, c = c0 + tx;
    if (p < HW && c < C) dst[(size_t(n) * HW + p) * C + c] = tile[tx][r];
  }
}
__global__ void pool_cl_f32_kernel(const float* __restrict__ src, float* __restrict__ dst, int N, int hs, int ws, int C) {
  const int hd = hs >> 1, wd = ws >> 1;
  const size_t total = size_t(N) * hd * wd * C;
  for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += size_t(gridDim.x) * blockDim.x) {
    const int c = int(i % C);
    size_t r = i / C;
    const int x = int(r % wd); r /= wd;
    const int y = int(r % hd);
    const size_t n = r / hd;
    const float* s = src + ((n * hs + 2 * y) * ws + 2 * x) * size_t(C) + c;
    dst[i] = (s[0] + s[C] + s[size_t(ws) * C] + s[size_t(ws) * C + C]) * 0.25f;
  }
}
__global__ void split_tf32_kernel(const float* __restrict__ x, float* __restrict__ hi, float* __restrict__ lo, size_t n) {
  for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += size_t(gridDim.x) * blockDim.x) {
    const float v = x[i];
    uint32_t h, l;
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(h) : "f"(v));
    const float r = v - __uint_as_float(h);
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(l) : "f"(r));
    hi[i] = __uint_as_float(h);
    lo[i] = __uint_as_float(l);
  }
}

constexpr int kTKf = 32;                                 // fp32 elements per 128-byte swizzle row = K per TMA box
constexpr int kStagesF = 4;
constexpr uint32_t kStageBytesF = (kTM + kTN) * kTKf * 4;           // 32 KB: one K-block of A and of B
constexpr size_t kSmemTF32 = kStagesF * kStageBytesF + size_t(kTM) * kTN * 4 + 1024;   // + one fp32 output staging tile
// instruction descriptor: D = F32, A = B = TF32 (format 2), both K-major
constexpr uint32_t kIdescTF32 = (1u << 4) | (2u << 7) | (2u << 10) | (uint32_t(kTN >> 3) << 17) | (uint32_t(kTM >> 4) << 24);

__device__ __forceinline__ void umma_tf32(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
               ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}

__global__ void __launch_bounds__(kPipeThreadsTC, 1)
corr_gemm_tc_tf32_kernel(const __grid_constant__ CUtensorMap tmAhi, const __grid_constant__ CUtensorMap tmAlo,
                         const __grid_constant__ CUtensorMap tmBhi, const __grid_constant__ CUtensorMap tmBlo,
                         const __grid_constant__ CUtensorMap tmO, int M, int N2, int E, int* __restrict__ err) {
  extern __shared__ unsigned char smem_dyn[];
  __shared__ __align__(8) uint64_t bars[2 * kStagesF + 4];   // full[4], empty[4], tfull[2], tempty[2]
  __shared__ uint32_t tmem_base_s;
  const uint32_t s0 = (smem_u32(smem_dyn) + 1023u) & ~1023u;
  const uint32_t stg = s0 + kStagesF * kStageBytesF;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int MT = (M + kTM - 1) / kTM, NT = (N2 + kTN - 1) / kTN;
  const long long tiles = static_cast<long long>(E) * MT * NT;
  constexpr int kKB = 12;                                   // 3 products x (128 / 32) K-blocks
  auto full = [&](int i) { return smem_u32(&bars[i]); };
  auto empty = [&](int i) { return smem_u32(&bars[kStagesF + i]); };
  auto tfull = [&](int i) { return smem_u32(&bars[2 * kStagesF + i]); };
  auto tempty = [&](int i) { return smem_u32(&bars[2 * kStagesF + 2 + i]); };

  if (tid == 0) {
    for (int i = 0; i < 2 * kStagesF + 2; i++) mbar_init(smem_u32(&bars[i]), 1);
    mbar_init(tempty(0), 4); mbar_init(tempty(1), 4);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_s)), "n"(2 * kTN) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem_base = tmem_base_s;

  if (warp == 0) {
    if (lane == 0) {
      long long it = 0;
      bool dead = false;
      for (long long t = blockIdx.x; t < tiles && !dead; t += gridDim.x) {
        const int nt = int(t % NT), mt = int((t / NT) % MT), e = int(t / (static_cast<long long>(NT) * MT));
        for (int kb = 0; kb < kKB; kb++, it++) {
          const int s = int(it % kStagesF);
          const uint32_t use = uint32_t(it / kStagesF);
          if (!mbar_wait(empty(s), (use & 1) ^ 1)) { atomicExch(err, 1); dead = true; break; }
          const uint32_t sA = s0 + s * kStageBytesF, sB = sA + kTM * kTKf * 4;
          const int prod = kb >> 2, koff = (kb & 3) * kTKf;      // 0: hi.hi  1: hi.lo  2: lo.hi
          mbar_expect_tx(full(s), kStageBytesF);
          tma_load_2d(sA, prod == 2 ? &tmAlo : &tmAhi, koff, e * M + mt * kTM, full(s));
          tma_load_2d(sB, prod == 1 ? &tmBlo : &tmBhi, koff, e * N2 + nt * kTN, full(s));
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      long long it = 0;
      int k = 0;
      bool dead = false;
      for (long long t = blockIdx.x; t < tiles && !dead; t += gridDim.x, k++) {
        const int acc = k & 1;
        if (!mbar_wait(tempty(acc), ((k >> 1) & 1) ^ 1)) { atomicExch(err, 1); break; }
        for (int kb = 0; kb < kKB; kb++, it++) {
          const int s = int(it % kStagesF);
          const uint32_t use = uint32_t(it / kStagesF);
          if (!mbar_wait(full(s), use & 1)) { atomicExch(err, 1); dead = true; break; }
          asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
          const uint32_t sA = s0 + s * kStageBytesF, sB = sA + kTM * kTKf * 4;
#pragma unroll
          for (int kk = 0; kk < kTKf / 8; kk++)
            umma_tf32(tmem_base + uint32_t(acc * kTN), umma_desc_sw128(sA + kk * 32), umma_desc_sw128(sB + kk * 32), kIdescTF32,
                      (kb | kk) != 0 ? 1u : 0u);
          umma_commit(empty(s));
        }
        if (!dead) umma_commit(tfull(acc));
      }
    }
  } else {
    const int quarter = warp & 3;
    int k = 0;
    for (long long t = blockIdx.x; t < tiles; t += gridDim.x, k++) {
      const int acc = k & 1;
      const int nt = int(t % NT), mt = int((t / NT) % MT), e = int(t / (static_cast<long long>(NT) * MT));
      const bool ok = mbar_wait(tfull(acc), (k >> 1) & 1);
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      if (!ok) { if (lane == 0) atomicExch(err, 1); break; }
      const int n0 = nt * kTN, trow = 32 * quarter + lane;
      if (tid == 64) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");      // the single staging tile is free again
      epi_bar<128>();
#pragma unroll 1
      for (int cb = 0; cb < kTN; cb += 32) {
        uint32_t v[32];
        tmem_ld32(tmem_base + (uint32_t(32 * quarter) << 16) + uint32_t(acc * kTN + cb), v);
        const uint32_t box = stg + uint32_t(cb >> 5) * (kTM * 128) + uint32_t(trow) * 128;
#pragma unroll
        for (int c4 = 0; c4 < 8; c4++) {
          const uint32_t a = box + uint32_t((c4 ^ (trow & 7)) << 4);
          asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(a), "f"(__uint_as_float(v[4 * c4]) * 0.0625f),
                       "f"(__uint_as_float(v[4 * c4 + 1]) * 0.0625f), "f"(__uint_as_float(v[4 * c4 + 2]) * 0.0625f),
                       "f"(__uint_as_float(v[4 * c4 + 3]) * 0.0625f) : "memory");
        }
      }
      asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
      __syncwarp();
      if (lane == 0) mbar_arrive(tempty(acc));
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      epi_bar<128>();
      if (tid == 64) {
#pragma unroll
        for (int b = 0; b < 4; b++)
          if (n0 + 32 * b < N2) tma_store_3d(&tmO, stg + b * (kTM * 128), n0 + 32 * b, mt * kTM, e);
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
      }
    }
    if (tid == 64) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 1) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(2 * kTN) : "memory");
}

// ---- host -------------------------------------------------------------------------------------------------------
// rows x C f16 matrix, row-major; box = 64 columns x 128 rows, 128-byte swizzle
static bool make_map(CUtensorMap* m, const void* base, uint64_t rows, int C) {
  EncodeTiledFn fn = encode_fn();
  if (!fn) return false;
  const cuuint64_t gdim[2] = {cuuint64_t(C), rows};
  const cuuint64_t gstride[1] = {cuuint64_t(C) * 2};
  const cuuint32_t box[2] = {cuuint32_t(kTK), cuuint32_t(kTM)};
  const cuuint32_t estr[2] = {1, 1};
  return fn(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, const_cast<void*>(base), gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
            CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

// fmap [rows = E*C][HW] f16, pixels contiguous: box = 64 pixels x 128 channel rows (one 16 KB MN block of an MN-major operand)
static bool make_map_mn(CUtensorMap* m, const void* base, int HW, uint64_t rows) {
  EncodeTiledFn fn = encode_fn();
  if (!fn) return false;
  const cuuint64_t gdim[2] = {cuuint64_t(HW), rows};
  const cuuint64_t gstride[1] = {cuuint64_t(HW) * 2};
  const cuuint32_t box[2] = {64, cuuint32_t(2 * kTK)};
  const cuuint32_t estr[2] = {1, 1};
  return fn(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 2, const_cast<void*>(base), gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
            CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

// output level [E][M][N2] f16 as a 3-D tensor: box = 64 columns x 128 rows x 1 edge, 128-byte swizzle
static bool make_out_map(CUtensorMap* m, void* base, int N2, int M, int E) {
  EncodeTiledFn fn = encode_fn();
  if (!fn) return false;
  const cuuint64_t gdim[3] = {cuuint64_t(N2), cuuint64_t(M), cuuint64_t(E)};
  const cuuint64_t gstride[2] = {cuuint64_t(N2) * 2, cuuint64_t(M) * N2 * 2};
  const cuuint32_t box[3] = {64, cuuint32_t(kTM), 1};
  const cuuint32_t estr[3] = {1, 1, 1};
  return fn(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 3, base, gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
            CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

// FUSED pyramid: level 1 in tiled slices seen as [2 rows x 8 = 16 halfs][2 row pairs][w1/8 tile columns][h1/4 tile rows][rows of
// the volume]; box = the row pair of two neighbouring tiles for 128 volume rows.  Level 2 (row-major [h2][w2] slices): box = 8
// columns of one row for 128 volume rows.
static bool make_fused_maps(CUtensorMap* m1, CUtensorMap* m2, void* l1, void* l2, int ht, int wd, uint64_t rows) {
  EncodeTiledFn fn = encode_fn();
  if (!fn) return false;
  const int h1 = ht >> 1, w1 = wd >> 1, h2 = ht >> 2, w2 = wd >> 2;
  const cuuint32_t estr[5] = {1, 1, 1, 1, 1};
  if (l1) {
    const cuuint64_t gdim[5] = {16, 2, cuuint64_t(w1 / 8), cuuint64_t(h1 / 4), rows};
    const cuuint64_t gstride[4] = {32, 64, cuuint64_t(w1 / 8) * 64, cuuint64_t(h1) * w1 * 2};
    const cuuint32_t box[5] = {16, 1, 2, 1, cuuint32_t(kTM)};
    if (fn(m1, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 5, l1, gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
           CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
      return false;
  }
  if (l2) {
    const cuuint64_t gdim[3] = {cuuint64_t(w2), cuuint64_t(h2), rows};
    const cuuint64_t gstride[2] = {cuuint64_t(w2) * 2, cuuint64_t(h2) * w2 * 2};
    const cuuint32_t box[3] = {8, 1, cuuint32_t(kTM)};
    if (fn(m2, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 3, l2, gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
           CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS)
      return false;
  }
  return true;
}

size_t corr_tc_workspace_bytes(int N, int C, int ht, int wd, int num_levels) {
  size_t tot = 256 + 256;                                           // status word
  tot += align_up256(size_t(N) * C * ht * wd * 2);                  // f1T
  int h = ht, w = wd;
  for (int l = 0; l < num_levels; l++) { tot += align_up256(size_t(N) * C * h * w * 2); h >>= 1; w >>= 1; }
  return tot;
}

bool corr_tc_supported(int C, int ht, int wd) { return C == 2 * kTK && (ht * wd) % 8 == 0 && encode_fn() != nullptr; }

int corr_volume_pyramid_tc(const __half* fmap1, const __half* fmap2, void* const* levels_out, int N, int C, int ht, int wd,
                           int num_levels, void* workspace, cudaStream_t st, unsigned tiled_mask) {
  char* ws = static_cast<char*>(workspace);
  int* err = reinterpret_cast<int*>(ws);
  ws += 256;
  DB200_CUDA(cudaMemsetAsync(err, 0, sizeof(int), st));
  const int HW = ht * wd;
  // fused pyramid: 64-wide maps whose level 0 (and level 1, if wanted) are tiled; every level from ONE launch, fmap1 read in place
  static const bool no_fused = getenv("DB200_CORR_NO_FUSED_PYRAMID") != nullptr || getenv("DB200_CORR_TC_ONESHOT") != nullptr ||
                               getenv("DB200_CORR_TC_NO_TMA_STORE") != nullptr;
  bool fused = !no_fused && wd == 64 && ht % 8 == 0 && (tiled_mask & 1u) && num_levels <= 4 && (num_levels < 2 || (tiled_mask & 2u)) &&
               (tiled_mask >> 2) == 0 && HW % kTM == 0 && reinterpret_cast<uintptr_t>(fmap1) % 16 == 0;
  for (int q = 0; q < num_levels && fused; q++) fused = reinterpret_cast<uintptr_t>(levels_out[q]) % 16 == 0;
  __half* f1T = reinterpret_cast<__half*>(ws);
  ws += align_up256(size_t(N) * C * HW * 2);
  CUtensorMap tmA;
  if (fused) {                                                       // [C][HW] as it lies in memory: an MN-major operand
    if (!make_map_mn(&tmA, fmap1, HW, uint64_t(N) * C)) return DB200_ERR_CUDA;
  } else {
    dim3 grid(cdiv(HW, 64), cdiv(C, 64), N);
    transpose_cl_kernel<<<grid, 256, 0, st>>>(fmap1, f1T, C, HW);
    DB200_CHECK_LAUNCH();
    if (!make_map(&tmA, f1T, uint64_t(N) * HW, C)) return DB200_ERR_CUDA;
  }
  const size_t smem = size_t(2) * (kTM + kTN) * kTK * 2 + 1024;
  static bool attr = false;
  if (!attr) {
    DB200_CUDA(cudaFuncSetAttribute(corr_gemm_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
    DB200_CUDA(cudaFuncSetAttribute(corr_gemm_tc_pipe_kernel<true, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(kPipeSmemTC)));
    DB200_CUDA(cudaFuncSetAttribute(corr_gemm_tc_pipe_kernel<false, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(kPipeSmemTC)));
    DB200_CUDA(cudaFuncSetAttribute(corr_gemm_tc_pipe_kernel<true, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, int(kPipeSmemFused)));
    attr = true;
  }
  const __half* prev = nullptr;
  bool prev_tiled = false;
  int h = ht, w = wd;
  for (int l = 0; l < num_levels; l++) {
    if (h == 0 || w == 0) return DB200_ERR_ARG;
    const bool tiled = ((tiled_mask >> l) & 1u) != 0;                // the level's [h][w] slices come out in tiled order
    if (tiled && !level_is_tiled(h, w)) return DB200_ERR_ARG;
    __half* f2T = reinterpret_cast<__half*>(ws);
    ws += align_up256(size_t(N) * C * h * w * 2);
    if (l == 0) {
      dim3 grid(cdiv(HW, 64), cdiv(C, 64), N);
      transpose_cl_kernel<<<grid, 256, 0, st>>>(fmap2, f2T, C, HW, tiled ? w : 0);
    } else {
      const size_t cnt = size_t(N) * h * w * (C / 2);
      const int blocks = int((cnt + 255) / 256 < 2368 ? (cnt + 255) / 256 : 2368);
      pool_cl_kernel<<<blocks, 256, 0, st>>>(prev, f2T, N, h << 1, w << 1, C, prev_tiled, tiled);
    }
    DB200_CHECK_LAUNCH();
    const int N2 = h * w;
    CUtensorMap tmB;
    if (!make_map(&tmB, f2T, uint64_t(N) * N2, C)) return DB200_ERR_CUDA;
    static const bool one_shot = getenv("DB200_CORR_TC_ONESHOT") != nullptr;       // previous kernel: one tile per CTA
    if (one_shot) {
      dim3 grid(cdiv(N2, kTN), cdiv(HW, kTM), N);
      corr_gemm_tc_kernel<<<grid, 128, smem, st>>>(tmA, tmB, static_cast<__half*>(levels_out[l]), HW, N2, err);
    } else {
      const long long tiles = static_cast<long long>(N) * cdiv(HW, kTM) * cdiv(N2, kTN);
      const int grid = int(tiles < kNumSMs ? tiles : kNumSMs);
      static const bool no_tma_store = getenv("DB200_CORR_TC_NO_TMA_STORE") != nullptr;
      CUtensorMap tmO = tmB;                                           // placeholder when the direct-store epilogue is used
      const bool tma_store = !no_tma_store && N2 % 64 == 0 && (reinterpret_cast<uintptr_t>(levels_out[l]) % 16 == 0) &&
                             make_out_map(&tmO, levels_out[l], N2, HW, N);
      if (fused && !tma_store) return DB200_ERR_CUDA;
      if (fused) {                                                   // all levels from this one launch
        FusedOut fo{num_levels > 1 ? static_cast<__half*>(levels_out[1]) : nullptr, num_levels > 2 ? static_cast<__half*>(levels_out[2]) : nullptr,
                    num_levels > 3 ? static_cast<__half*>(levels_out[3]) : nullptr};
        CUtensorMap tmL1 = tmB, tmL2 = tmB;                            // placeholders for levels that are not wanted
        if (!make_fused_maps(&tmL1, &tmL2, fo.l1, fo.l2, ht, wd, uint64_t(N) * HW)) return DB200_ERR_CUDA;
        const long long units = tiles / 4;
        const int g4 = int(units < kNumSMs ? units : kNumSMs);
        corr_gemm_tc_pipe_kernel<true, true><<<g4, kPipeThreadsTC, kPipeSmemFused, st>>>(tmA, tmB, tmO, static_cast<__half*>(levels_out[l]), HW, N2, N,
                                                                                         err, fo, tmL1, tmL2);
        DB200_CHECK_LAUNCH();
        return DB200_OK;
      }
      if (tma_store)
        corr_gemm_tc_pipe_kernel<true, false><<<grid, kPipeThreadsTC, kPipeSmemTC, st>>>(tmA, tmB, tmO, static_cast<__half*>(levels_out[l]), HW, N2, N, err, FusedOut{}, tmB, tmB);
      else
        corr_gemm_tc_pipe_kernel<false, false><<<grid, kPipeThreadsTC, kPipeSmemTC, st>>>(tmA, tmB, tmO, static_cast<__half*>(levels_out[l]), HW, N2, N, err, FusedOut{}, tmB, tmB);
    }
    DB200_CHECK_LAUNCH();
    prev = f2T;
    prev_tiled = tiled;
    h >>= 1; w >>= 1;
  }
  return DB200_OK;
}

// ---- fp32 host path ---------------------------------------------------------------------------------------------------
static bool make_map_f32(CUtensorMap* m, const void* base, uint64_t rows, int C) {
  EncodeTiledFn fn = encode_fn();
  if (!fn) return false;
  const cuuint64_t gdim[2] = {cuuint64_t(C), rows};
  const cuuint64_t gstride[1] = {cuuint64_t(C) * 4};
  const cuuint32_t box[2] = {cuuint32_t(kTKf), cuuint32_t(kTM)};
  const cuuint32_t estr[2] = {1, 1};
  return fn(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<void*>(base), gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
            CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}
static bool make_out_map_f32(CUtensorMap* m, void* base, int N2, int M, int E) {
  EncodeTiledFn fn = encode_fn();
  if (!fn) return false;
  const cuuint64_t gdim[3] = {cuuint64_t(N2), cuuint64_t(M), cuuint64_t(E)};
  const cuuint64_t gstride[2] = {cuuint64_t(N2) * 4, cuuint64_t(M) * N2 * 4};
  const cuuint32_t box[3] = {32, cuuint32_t(kTM), 1};
  const cuuint32_t estr[3] = {1, 1, 1};
  return fn(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, base, gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
            CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

// every level needs N2 >= 32 and N2 % 4 == 0 (TMA store box / 16-byte global strides)
bool corr_tc_f32_supported(int C, int ht, int wd, int num_levels) {
  if (C != 128 || encode_fn() == nullptr) return false;
  int h = ht, w = wd;
  for (int l = 0; l < num_levels; l++) {
    const int n2 = h * w;
    if (n2 < 32 || n2 % 4 != 0) return false;
    h >>= 1; w >>= 1;
  }
  return true;
}

size_t corr_tc_f32_workspace_bytes(int N, int C, int ht, int wd, int num_levels) {
  size_t tot = 512;
  tot += 2 * align_up256(size_t(N) * C * ht * wd * 4);               // f1 hi, lo (the full copy reuses f2's level-0 slot order)
  tot += align_up256(size_t(N) * C * ht * wd * 4);                   // f1 full (transposed) before the split
  int h = ht, w = wd;
  for (int l = 0; l < num_levels; l++) { tot += 3 * align_up256(size_t(N) * C * h * w * 4); h >>= 1; w >>= 1; }   // f2 full, hi, lo
  return tot;
}

int corr_volume_pyramid_tc_f32(const float* fmap1, const float* fmap2, void* const* levels_out, int N, int C, int ht, int wd,
                               int num_levels, void* workspace, cudaStream_t st) {
  char* ws = static_cast<char*>(workspace);
  int* err = reinterpret_cast<int*>(ws);
  ws += 256;
  DB200_CUDA(cudaMemsetAsync(err, 0, sizeof(int), st));
  const int HW = ht * wd;
  auto take = [&ws](size_t bytes) { char* p = ws; ws += align_up256(bytes); return reinterpret_cast<float*>(p); };
  const size_t n1 = size_t(N) * C * HW;
  float* f1full = take(n1 * 4); float* f1hi = take(n1 * 4); float* f1lo = take(n1 * 4);
  {
    dim3 grid(cdiv(HW, 32), cdiv(C, 32), N);
    transpose_cl_f32_kernel<<<grid, 256, 0, st>>>(fmap1, f1full, C, HW);
    DB200_CHECK_LAUNCH();
    split_tf32_kernel<<<kNumSMs * 8, 256, 0, st>>>(f1full, f1hi, f1lo, n1);
    DB200_CHECK_LAUNCH();
  }
  CUtensorMap tmAhi, tmAlo;
  if (!make_map_f32(&tmAhi, f1hi, uint64_t(N) * HW, C) || !make_map_f32(&tmAlo, f1lo, uint64_t(N) * HW, C)) return DB200_ERR_CUDA;
  static bool attr = false;
  if (!attr) { DB200_CUDA(cudaFuncSetAttribute(corr_gemm_tc_tf32_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, int(kSmemTF32))); attr = true; }
  const float* prev = nullptr;
  int h = ht, w = wd;
  for (int l = 0; l < num_levels; l++) {
    const size_t n2e = size_t(N) * C * h * w;
    float* full = take(n2e * 4); float* hi = take(n2e * 4); float* lo = take(n2e * 4);
    if (l == 0) {
      dim3 grid(cdiv(HW, 32), cdiv(C, 32), N);
      transpose_cl_f32_kernel<<<grid, 256, 0, st>>>(fmap2, full, C, HW);
    } else {
      const int blocks = int((n2e + 255) / 256 < 2368 ? (n2e + 255) / 256 : 2368);
      pool_cl_f32_kernel<<<blocks, 256, 0, st>>>(prev, full, N, h << 1, w << 1, C);
    }
    DB200_CHECK_LAUNCH();
    split_tf32_kernel<<<kNumSMs * 8, 256, 0, st>>>(full, hi, lo, n2e);
    DB200_CHECK_LAUNCH();
    const int N2 = h * w;
    CUtensorMap tmBhi, tmBlo, tmO;
    if (!make_map_f32(&tmBhi, hi, uint64_t(N) * N2, C) || !make_map_f32(&tmBlo, lo, uint64_t(N) * N2, C) ||
        !make_out_map_f32(&tmO, levels_out[l], N2, HW, N))
      return DB200_ERR_CUDA;
    const long long tiles = static_cast<long long>(N) * cdiv(HW, kTM) * cdiv(N2, kTN);
    const int grid = int(tiles < kNumSMs ? tiles : kNumSMs);
    corr_gemm_tc_tf32_kernel<<<grid, kPipeThreadsTC, kSmemTF32, st>>>(tmAhi, tmAlo, tmBhi, tmBlo, tmO, HW, N2, N, err);
    DB200_CHECK_LAUNCH();
    prev = full;
    h >>= 1; w >>= 1;
  }
  return DB200_OK;
}

}  // namespace db200
