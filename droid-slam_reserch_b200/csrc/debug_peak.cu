// debug_peak.cu — instrumentation only (no reference counterpart): the FP64 rate this GPU sustains on the path the
// Cholesky solver's trailing updates use (mma.sync.m8n8k4.f64 = SASS DMMA), measured in the benchmark process so that
// bench.py's `roofline.solver.peak_gflops` is a number of THIS run, not a literal.  Register-only: 8 independent
// accumulator tiles per warp, 8 warps per CTA, 2 CTAs per SM — nothing but the FP64 pipe can limit it.
#include "common.cuh"

namespace db200 {

__global__ void __launch_bounds__(256, 2) fp64_peak_kernel(double* __restrict__ sink, int iters) {
  double acc[8][2];
#pragma unroll
  for (int i = 0; i < 8; i++) { acc[i][0] = 0.0; acc[i][1] = 0.0; }
  double a = 1.0 + 1e-9 * threadIdx.x, b = 1.0 - 1e-9 * threadIdx.x;
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int i = 0; i < 8; i++)
      asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                   : "+d"(acc[i][0]), "+d"(acc[i][1]) : "d"(a), "d"(b));
  }
  double s = 0.0;
#pragma unroll
  for (int i = 0; i < 8; i++) s += acc[i][0] + acc[i][1];
  if (s == 12345.678) sink[blockIdx.x * blockDim.x + threadIdx.x] = s;     // never true: keeps the loop alive
}

}  // namespace db200

// returns the measured rate in GFLOP/s (2 flops per FMA; one m8n8k4 = 256 FMA per warp), or a negative status
extern "C" double db200_debug_fp64_peak_gflops(db200_stream_t stream) {
  using namespace db200;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  double* sink = nullptr;
  if (cudaMalloc(&sink, sizeof(double) * 2 * kNumSMs * 256) != cudaSuccess) return double(DB200_ERR_CUDA);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  const int iters = 20000, grid = 2 * kNumSMs;
  fp64_peak_kernel<<<grid, 256, 0, st>>>(sink, 2000);                      // warm-up
  double best = 0.0;
  for (int rep = 0; rep < 3; rep++) {
    cudaEventRecord(e0, st);
    fp64_peak_kernel<<<grid, 256, 0, st>>>(sink, iters);
    cudaEventRecord(e1, st);
    if (cudaEventSynchronize(e1) != cudaSuccess) { best = double(DB200_ERR_CUDA); break; }
    float ms = 0.0f;
    cudaEventElapsedTime(&ms, e0, e1);
    const double flops = 2.0 * 256.0 * 8.0 * double(iters) * 8.0 * grid;  // 8 tiles x iters per warp, 8 warps per CTA
    const double g = flops / (double(ms) * 1e-3) / 1e9;
    if (g > best) best = g;
  }
  DB200_COUNT(4);
  cudaEventDestroy(e0); cudaEventDestroy(e1);
  cudaFree(sink);
  return best;
}
