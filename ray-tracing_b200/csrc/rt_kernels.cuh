// rt_kernels.cuh — the kernels of the B200 path-tracing core and their launchers.
//   renderKernel  : K2, persistent warps; every lane owns one camera path at a time and refills itself from the
//                   warp's chunk of the global path index space as soon as its path ends (path regeneration by
//                   warp ballot + prefix rank); one traversal site serves both the bounce query and the
//                   light-pdf queries of the Lambertian mixture pdf, so lanes in different shading phases still
//                   walk the BVH together.
//   traceKernel   : K1, `hit world ray t_min t_max` for explicit rays with the visit log (parity probe).
//   textureKernel : textureValue probe.   tonemapKernel : albedoToColor.
// Included by rt_launch_exact.cu (--fmad=false, MODE 0), rt_launch_fast.cu (MODE 1) and rt_launch_strict.cu (--fmad=false, MODE 2).
#pragma once
#include <algorithm>
#include <cstdlib>

#include "rt_device.cuh"
#include "rt_traverse.cuh"
#include "rt_launch.h"

namespace rtb {


#ifndef RTB_MIN_BLOCKS
#define RTB_MIN_BLOCKS 6  // resident 128-thread blocks per SM the fp32 render kernels are compiled for (80 registers)
#endif
template <class T> struct MinBlocks { static constexpr int v = RTB_MIN_BLOCKS; };
template <> struct MinBlocks<double> { static constexpr int v = 1; };

#include "rt_render_kernel.inl"
}  // namespace rtb
#include "rt_coop.cuh"
#include "rt_wavefront.cuh"
namespace rtb {

template <class T, int MODE>
__global__ void textureKernel(const __grid_constant__ DScene<T> sc, uint32_t tex, uint32_t n, const double* q, double* out) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  Instr<0> ins;
  const double* a = q + (size_t)i * 5;
  V3<T> c = textureValue<T, 0, MODE>(sc, tex, (T)a[0], (T)a[1], mk<T>((T)a[2], (T)a[3], (T)a[4]), ins);
  out[i * 3] = c.x; out[i * 3 + 1] = c.y; out[i * 3 + 2] = c.z;
}

// ---- launchers -------------------------------------------------------------------------------------
template <class T, class S, int INSTR, int MODE>
static cudaError_t launchRenderT(const DScene<T>& sc, const RenderArgs& a, int stackDepth, int numSMs, cudaStream_t st, int* gridOut) {
  auto kern = renderKernel<T, S, INSTR, MODE>;
  size_t smem = (size_t)columnWords<T>((uint32_t)stackDepth, a.lightCap) * kBlock * sizeof(uint32_t) + (kBlock / 32) * 16;  // columns + per-warp chunk slots
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  int perSM = 0;
  e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSM, kern, kBlock, smem);
  if (e != cudaSuccess) return e;
  if (perSM < 1) perSM = 1;
  if (const char* v = std::getenv("RTB_BLOCKS_PER_SM")) perSM = std::max(1, std::min(perSM, std::atoi(v)));  // tuning experiments only
  int grid = numSMs * perSM;  // persistent: exactly one resident wave
  if (gridOut) *gridOut = grid;
  RenderArgs b = a; b.stackCap = (uint32_t)stackDepth;
  kern<<<grid, kBlock, smem, st>>>(sc, b);
  return cudaGetLastError();
}

template <class T, class S, int MODE>
static cudaError_t launchRenderI(int instr, const DScene<T>& sc, const RenderArgs& a, int stackDepth, int numSMs, cudaStream_t st, int* gridOut) {
  if (instr == 1) return launchRenderT<T, S, 1, MODE>(sc, a, stackDepth, numSMs, st, gridOut);
  if constexpr (MODE != MATH_FAST) { if (instr == 3) return launchRenderT<T, S, 3, MODE>(sc, a, stackDepth, numSMs, st, gridOut); }
  return launchRenderT<T, S, 0, MODE>(sc, a, stackDepth, numSMs, st, gridOut);
}
template <int MODE>
cudaError_t launchRender(int prec, int instr, const void* dscene, const RenderArgs& a, int stackDepth, int numSMs, cudaStream_t st, int* gridOut) {
  if (prec == 0) return launchRenderI<float, float, MODE>(instr, *static_cast<const DScene<float>*>(dscene), a, stackDepth, numSMs, st, gridOut);
  if (prec == 1) return launchRenderI<float, double, MODE>(instr, *static_cast<const DScene<float>*>(dscene), a, stackDepth, numSMs, st, gridOut);
  return launchRenderI<double, double, MODE>(instr, *static_cast<const DScene<double>*>(dscene), a, stackDepth, numSMs, st, gridOut);
}

template <class T, class S, int MODE>
static cudaError_t launchTraceT(const DScene<T>& sc, const TraceArgs& a, int stackDepth, cudaStream_t st) {
  auto kern = traceKernel<T, S, MODE>;
  size_t smem = (size_t)(stackDepth + extraWords<T>()) * kBlock * sizeof(uint32_t);
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  TraceArgs b = a; b.stackCap = (uint32_t)stackDepth;
  kern<<<(a.n + kBlock - 1) / kBlock, kBlock, smem, st>>>(sc, b);
  return cudaGetLastError();
}
template <int MODE>
cudaError_t launchTrace(int prec, const void* dscene, const TraceArgs& a, int stackDepth, cudaStream_t st) {
  if (prec == 0) return launchTraceT<float, float, MODE>(*static_cast<const DScene<float>*>(dscene), a, stackDepth, st);
  if (prec == 1) return launchTraceT<float, double, MODE>(*static_cast<const DScene<float>*>(dscene), a, stackDepth, st);
  return launchTraceT<double, double, MODE>(*static_cast<const DScene<double>*>(dscene), a, stackDepth, st);
}

// ---- scan-stream (warp-cooperative) kernels: persistent warps, no shared memory
template <class T, class S, int INSTR, int MODE>
static cudaError_t launchRenderCoopT(const DScene<T>& sc, const RenderArgs& a, int numSMs, cudaStream_t st, int* gridOut) {
  auto kern = renderKernelCoop<T, S, INSTR, MODE>;
  int perSM = 0;
  cudaError_t e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSM, kern, kBlock, 0);
  if (e != cudaSuccess) return e;
  if (perSM < 1) perSM = 1;
  if (const char* v = std::getenv("RTB_COOP_BLOCKS_PER_SM")) perSM = std::max(1, std::min(perSM, std::atoi(v)));  // tuning experiments only
  // one resident wave, but never more warps than paths (a warp owns a whole path)
  const unsigned long long warpsWanted = a.total_paths;
  int grid = (int)std::min<unsigned long long>((unsigned long long)numSMs * perSM, (warpsWanted + (kBlock / 32) - 1) / (kBlock / 32));
  if (grid < 1) grid = 1;
  if (gridOut) *gridOut = grid;
  kern<<<grid, kBlock, 0, st>>>(sc, a);
  return cudaGetLastError();
}
template <class T, class S, int MODE>
static cudaError_t launchRenderCoopI(int instr, const DScene<T>& sc, const RenderArgs& a, int numSMs, cudaStream_t st, int* gridOut) {
  if (instr == 1) return launchRenderCoopT<T, S, 1, MODE>(sc, a, numSMs, st, gridOut);
  if (instr == 3) return launchRenderCoopT<T, S, 3, MODE>(sc, a, numSMs, st, gridOut);
  return launchRenderCoopT<T, S, 0, MODE>(sc, a, numSMs, st, gridOut);
}
template <int MODE>
cudaError_t launchRenderCoop(int prec, int instr, const void* dscene, const RenderArgs& a, int numSMs, cudaStream_t st, int* gridOut) {
  if (prec == 0) return launchRenderCoopI<float, float, MODE>(instr, *static_cast<const DScene<float>*>(dscene), a, numSMs, st, gridOut);
  if (prec == 1) return launchRenderCoopI<float, double, MODE>(instr, *static_cast<const DScene<float>*>(dscene), a, numSMs, st, gridOut);
  return launchRenderCoopI<double, double, MODE>(instr, *static_cast<const DScene<double>*>(dscene), a, numSMs, st, gridOut);
}
template <class T, class S, int MODE>
static cudaError_t launchTraceCoopT(const DScene<T>& sc, const TraceArgs& a, cudaStream_t st) {
  const unsigned long long threads = (unsigned long long)a.n * 32ull;
  traceKernelCoop<T, S, MODE><<<(unsigned)((threads + kBlock - 1) / kBlock), kBlock, 0, st>>>(sc, a);
  return cudaGetLastError();
}
template <int MODE>
cudaError_t launchTraceCoop(int prec, const void* dscene, const TraceArgs& a, cudaStream_t st) {
  if (prec == 0) return launchTraceCoopT<float, float, MODE>(*static_cast<const DScene<float>*>(dscene), a, st);
  if (prec == 1) return launchTraceCoopT<float, double, MODE>(*static_cast<const DScene<float>*>(dscene), a, st);
  return launchTraceCoopT<double, double, MODE>(*static_cast<const DScene<double>*>(dscene), a, st);
}

template <int MODE>
cudaError_t launchTexture(int prec, const void* dscene, uint32_t tex, uint32_t n, const double* q, double* out, cudaStream_t st) {
  if (prec == 2) textureKernel<double, MODE><<<(n + 127) / 128, 128, 0, st>>>(*static_cast<const DScene<double>*>(dscene), tex, n, q, out);
  else textureKernel<float, MODE><<<(n + 127) / 128, 128, 0, st>>>(*static_cast<const DScene<float>*>(dscene), tex, n, q, out);
  return cudaGetLastError();
}

}  // namespace rtb
