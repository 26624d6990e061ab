// Micro-benchmark of the Cholesky tile primitives and of the instruction latencies they are bound by (1 CTA).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 --expt-relaxed-constexpr -o /tmp/mb_chol tools/mb_chol.cu
#include <cstdio>
#include <vector>
#include <cmath>
#include "../droid-slam_reserch_b200/csrc/ba_solve.cu"
extern "C" void db200_note_cuda_error(int, const char*, int) {}
extern "C" void db200_count_launches(int) {}
using namespace db200;

__global__ void __launch_bounds__(256, 1) mb_kernel(long long* out, const double* spd, const double* rect, int* fail) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  double (*sA)[TP] = reinterpret_cast<double (*)[TP]>(smem_raw);
  double (*sB)[TP] = reinterpret_cast<double (*)[TP]>(smem_raw + sizeof(double) * T * TP);
  double* sv = reinterpret_cast<double*>(smem_raw + 2 * sizeof(double) * T * TP);
  const int tid = threadIdx.x;
  long long t0, t1;
  int slot = 0;
  auto mark = [&](long long v) { if (tid == 0) out[slot] = v; slot++; };
  for (int rep = 0; rep < 3; rep++) {
    __syncthreads(); t0 = clock64();
    load_tile(sA, spd, 64);
    __syncthreads(); t1 = clock64(); mark(t1 - t0);          // 0: load_tile (L2/DRAM)
    double cf[4][4];
    load_frag(cf, spd, 64);
    __syncthreads();
    t0 = clock64();
    factor_tile(cf, sA, 0, 64, fail);
    __syncthreads(); t1 = clock64(); mark(t1 - t0);          // 1: factor_tile
    load_frag(cf, rect, 64);
    if (tid < T) sv[tid] = 1.0 / sA[tid][tid];
    __syncthreads(); t0 = clock64();
    trsm_tile(cf, sB, sA, sv);
    __syncthreads(); t1 = clock64(); mark(t1 - t0);          // 2: trsm_tile
    double acc[4][4];
    t0 = clock64();
    tile_aat_lower(sB, acc);
    __syncthreads(); t1 = clock64(); mark(t1 - t0);          // 3: tile_aat_lower
    double s = 0; for (int a = 0; a < 4; a++) for (int b = 0; b < 4; b++) s += acc[a][b];
    if (s == 1.2345) out[100] = 1;
    t0 = clock64();
    store_tile(const_cast<double*>(rect) + 4096, sB, 64);
    __syncthreads(); t1 = clock64(); mark(t1 - t0);          // 4: store_tile (issue only)
    t0 = clock64();
    if (tid == 0) { __threadfence(); }
    __syncthreads(); t1 = clock64(); mark(t1 - t0);          // 5: threadfence by thread 0 after stores
  }
  // latencies
  double x = spd[tid & 63] + 1.0, y = 1.000001;
  __syncthreads(); t0 = clock64();
#pragma unroll 1
  for (int i = 0; i < 256; i++) { x = fma(x, y, 0.5); x = fma(x, y, 0.5); x = fma(x, y, 0.5); x = fma(x, y, 0.5); }
  t1 = clock64(); mark((t1 - t0));                           // 18: 1024 dependent DFMA (all 8 warps)
  if (x == 1.2345) out[100] = 1;
  x = spd[tid & 63] + 2.0;
  t0 = clock64();
#pragma unroll 1
  for (int i = 0; i < 256; i++) { x = fast_rsqrt(x) + 1.5; }
  t1 = clock64(); mark((t1 - t0));                           // 19: 256 x (fast_rsqrt + add)
  if (x == 1.2345) out[100] = 1;
  x = spd[tid & 63] + 2.0;
  t0 = clock64();
#pragma unroll 1
  for (int i = 0; i < 256; i++) { x = rsqrt(x) + 1.5; }
  t1 = clock64(); mark((t1 - t0));                           // 20: 256 x (rsqrt + add)
  if (x == 1.2345) out[100] = 1;
  x = spd[tid & 63] + 2.0;
  t0 = clock64();
#pragma unroll 1
  for (int i = 0; i < 256; i++) { x = __shfl_sync(0xffffffffu, x, (tid + 1) & 31) + 1.0; }
  t1 = clock64(); mark((t1 - t0));                           // 21: 256 x (double shuffle + add)
  if (x == 1.2345) out[100] = 1;
  // single warp DFMA chain (only warp 0) to see latency without contention
  x = spd[tid & 63] + 1.0;
  __syncthreads(); t0 = clock64();
  if (tid < 32) {
#pragma unroll 1
    for (int i = 0; i < 256; i++) { x = fma(x, y, 0.5); x = fma(x, y, 0.5); x = fma(x, y, 0.5); x = fma(x, y, 0.5); }
  }
  t1 = clock64(); mark((t1 - t0));                           // 22: 1024 dependent DFMA, one warp
  if (x == 1.2345) out[100] = 1;
  // independent DFMA throughput: 8 independent chains per thread, all warps
  double z[8];
  for (int i = 0; i < 8; i++) z[i] = spd[(tid + i) & 63];
  __syncthreads(); t0 = clock64();
#pragma unroll 1
  for (int i = 0; i < 128; i++) {
#pragma unroll
    for (int k = 0; k < 8; k++) z[k] = fma(z[k], y, 0.5);
  }
  __syncthreads(); t1 = clock64(); mark((t1 - t0));          // 23: 1024 DFMA per thread, 8-way ILP, 8 warps
  for (int i = 0; i < 8; i++) if (z[i] == 1.2345) out[100] = 1;
  // smem latency chain
  int idx = tid & 63;
  __syncthreads(); t0 = clock64();
#pragma unroll 1
  for (int i = 0; i < 256; i++) { idx = (int)(sA[idx & 63][(idx >> 1) & 63] * 0.0) + ((idx + 1) & 63); }
  t1 = clock64(); mark((t1 - t0));                           // 24: 256 dependent LDS.64 (+cvt)
  if (idx == 12345) out[100] = 1;
}

int main() {
  std::vector<double> spd(4096), rect(8192);
  for (int r = 0; r < 64; r++) for (int c = 0; c < 64; c++) spd[r * 64 + c] = (r == c ? 80.0 : 0.0) + std::sin(0.1 * (r + 1) * (c + 1) + 0.3 * (r + c)) ;
  for (int r = 0; r < 64; r++) for (int c = 0; c < r; c++) spd[c * 64 + r] = spd[r * 64 + c];
  for (int i = 0; i < 4096; i++) rect[i] = std::cos(0.01 * i);
  double *dspd, *drect; long long* dout; int* dfail;
  cudaMalloc(&dspd, 4096 * 8); cudaMalloc(&drect, 8192 * 8); cudaMalloc(&dout, 128 * 8); cudaMalloc(&dfail, 4);
  cudaMemcpy(dspd, spd.data(), 4096 * 8, cudaMemcpyHostToDevice); cudaMemcpy(drect, rect.data(), 4096 * 8, cudaMemcpyHostToDevice);
  cudaMemset(dout, 0, 128 * 8); cudaMemset(dfail, 0, 4);
  const size_t smem = 3 * sizeof(double) * T * TP + sizeof(double) * 6 * 64;
  cudaFuncSetAttribute(mb_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem));
  mb_kernel<<<1, 256, smem>>>(dout, dspd, drect, dfail);
  cudaError_t e = cudaDeviceSynchronize();
  printf("status %s\n", cudaGetErrorString(e));
  long long out[128]; cudaMemcpy(out, dout, 128 * 8, cudaMemcpyDeviceToHost);
  const char* names[6] = {"load_tile", "factor_tile", "trsm_tile", "tile_abt", "store_tile(issue)", "fence"};
  for (int rep = 0; rep < 3; rep++) for (int k = 0; k < 6; k++) printf("rep%d %-18s %8lld cyc\n", rep, names[k], out[rep * 6 + k]);
  const char* n2[7] = {"1024 dep DFMA (8 warps)", "256 fast_rsqrt+add", "256 rsqrt+add", "256 dshfl+add", "1024 dep DFMA (1 warp)", "1024 DFMA ILP8 (8 warps)", "256 dep LDS64"};
  for (int k = 0; k < 7; k++) printf("%-26s %8lld cyc\n", n2[k], out[18 + k]);
  return 0;
}
