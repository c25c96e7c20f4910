// tools/fp32_peak.cu — measured issue peaks of one B200 for the roofline denominators bench.py reports (SURVEY 8d asks for a
// MEASURED FP32 figure next to MEASURED_PEAKS.json; MEASURED_PEAKS.json itself only has HBM and bf16 tensor numbers).
//   ffma        : FFMA chains, 8 independent accumulators per thread            -> lane-instructions/s on the FMA pipe (x2 = FLOP/s)
//   fadd_fmul   : alternating FADD / FMUL chains (what a --fmad=false build can issue)  -> lane-instructions/s, 1 flop each
//   l2_read     : every SM streams a 48 MB buffer (L2-resident on B200's 126 MB L2) with 128-bit loads -> GB/s
//   hbm_read    : same over a 4 GB buffer (> L2)                                  -> GB/s
// Timed with CUDA events on the launching stream after warm-up, best of 5.  Prints one JSON object.
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/fp32_peak tools/fp32_peak.cu
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { fprintf(stderr, "%s: %s\n", #x, cudaGetErrorString(e_)); return 1; } } while (0)

__global__ void ffmaKernel(float* out, int iters, float a, float b) {
  float x0 = threadIdx.x, x1 = x0 + 1, x2 = x0 + 2, x3 = x0 + 3, x4 = x0 + 4, x5 = x0 + 5, x6 = x0 + 6, x7 = x0 + 7;
#pragma unroll 1
  for (int i = 0; i < iters; ++i) {
#pragma unroll
    for (int k = 0; k < 16; ++k) {
      x0 = __fmaf_rn(x0, a, b); x1 = __fmaf_rn(x1, a, b); x2 = __fmaf_rn(x2, a, b); x3 = __fmaf_rn(x3, a, b);
      x4 = __fmaf_rn(x4, a, b); x5 = __fmaf_rn(x5, a, b); x6 = __fmaf_rn(x6, a, b); x7 = __fmaf_rn(x7, a, b);
    }
  }
  out[blockIdx.x * blockDim.x + threadIdx.x] = x0 + x1 + x2 + x3 + x4 + x5 + x6 + x7;
}
__global__ void faddFmulKernel(float* out, int iters, float a, float b) {
  float x0 = threadIdx.x, x1 = x0 + 1, x2 = x0 + 2, x3 = x0 + 3, x4 = x0 + 4, x5 = x0 + 5, x6 = x0 + 6, x7 = x0 + 7;
#pragma unroll 1
  for (int i = 0; i < iters; ++i) {
#pragma unroll
    for (int k = 0; k < 8; ++k) {
      x0 = __fmul_rn(x0, a); x1 = __fadd_rn(x1, b); x2 = __fmul_rn(x2, a); x3 = __fadd_rn(x3, b);
      x4 = __fmul_rn(x4, a); x5 = __fadd_rn(x5, b); x6 = __fmul_rn(x6, a); x7 = __fadd_rn(x7, b);
      x0 = __fadd_rn(x0, b); x1 = __fmul_rn(x1, a); x2 = __fadd_rn(x2, b); x3 = __fmul_rn(x3, a);
      x4 = __fadd_rn(x4, b); x5 = __fmul_rn(x5, a); x6 = __fadd_rn(x6, b); x7 = __fmul_rn(x7, a);
    }
  }
  out[blockIdx.x * blockDim.x + threadIdx.x] = x0 + x1 + x2 + x3 + x4 + x5 + x6 + x7;
}
__global__ void readKernel(const float4* buf, size_t n4, int passes, float* out) {
  float acc = 0;
  const size_t tid = blockIdx.x * (size_t)blockDim.x + threadIdx.x, nt = (size_t)gridDim.x * blockDim.x;
  for (int p = 0; p < passes; ++p)
    for (size_t i = tid; i < n4; i += nt) { const float4 v = __ldg(buf + i); acc += v.x + v.y + v.z + v.w; }
  if (acc == 123.456f) out[0] = acc;
}

template <class F> static float bestOf(int reps, F f) {
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  float best = 1e30f;
  for (int r = 0; r < reps + 2; ++r) {
    cudaEventRecord(e0); f(); cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    if (r >= 2 && ms < best) best = ms;
  }
  cudaEventDestroy(e0); cudaEventDestroy(e1);
  return best;
}

int main() {
  cudaDeviceProp prop; CK(cudaGetDeviceProperties(&prop, 0));
  const int sms = prop.multiProcessorCount, threads = 256, blocks = sms * 8;
  float* out; CK(cudaMalloc(&out, (size_t)blocks * threads * sizeof(float)));
  const int iters = 4096;
  const float msF = bestOf(5, [&] { ffmaKernel<<<blocks, threads>>>(out, iters, 1.0000001f, 1e-9f); });
  const float msA = bestOf(5, [&] { faddFmulKernel<<<blocks, threads>>>(out, iters, 1.0000001f, 1e-9f); });
  CK(cudaGetLastError());
  const double laneF = (double)blocks * threads * iters * 16 * 8, laneA = (double)blocks * threads * iters * 8 * 16;
  const size_t l2Bytes = 48ull << 20, hbmBytes = 4ull << 30;
  float4 *bl2, *bhbm; CK(cudaMalloc(&bl2, l2Bytes)); CK(cudaMalloc(&bhbm, hbmBytes));
  CK(cudaMemset(bl2, 0, l2Bytes)); CK(cudaMemset(bhbm, 0, hbmBytes));
  const int rblocks = sms * 16, passes = 20;
  readKernel<<<rblocks, 256>>>(bl2, l2Bytes / 16, 2, out);  // warm the L2
  const float msL2 = bestOf(5, [&] { readKernel<<<rblocks, 256>>>(bl2, l2Bytes / 16, passes, out); });
  const float msH = bestOf(5, [&] { readKernel<<<rblocks, 256>>>(bhbm, hbmBytes / 16, 1, out); });
  CK(cudaGetLastError());
  int clk = 0; cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
  printf("{\"gpu\": \"%s\", \"sms\": %d, \"sm_clock_max_mhz\": %.0f, "
         "\"ffma_lane_inst_per_s\": %.4e, \"ffma_tflops\": %.3f, \"fadd_fmul_lane_inst_per_s\": %.4e, \"fadd_fmul_tflops\": %.3f, "
         "\"fp32_warp_inst_per_s\": %.4e, \"l2_read_gbs\": %.1f, \"l2_buffer_mb\": %d, \"hbm_read_gbs\": %.1f, "
         "\"how\": \"tools/fp32_peak.cu: 8 independent chains per thread, %d blocks x %d threads, best of 5, CUDA events\"}\n",
         prop.name, sms, clk / 1e3, laneF / (msF * 1e-3), 2 * laneF / (msF * 1e-3) / 1e12, laneA / (msA * 1e-3), laneA / (msA * 1e-3) / 1e12,
         laneF / 32 / (msF * 1e-3), (double)l2Bytes * passes / (msL2 * 1e-3) / 1e9, (int)(l2Bytes >> 20), (double)hbmBytes / (msH * 1e-3) / 1e9, blocks, threads);
  cudaFree(out); cudaFree(bl2); cudaFree(bhbm);
  return 0;
}
