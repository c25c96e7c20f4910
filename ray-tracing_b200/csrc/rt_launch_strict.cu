// Strict-math instantiation of the kernels (RT_FLAG_STRICT_MATH): the literal arithmetic of rt_launch_exact.cu (build with
// --fmad=false, IEEE division / sqrt) with sin / cos / atan2 / asin / log taken from rt_strict_math.h instead of libdevice, so that
// every operation on a path is bit-identical to the test oracle's (which includes the same header).  A parity instrument: same
// kernels, same step and draw order, slower transcendentals.
#include "rt_kernels.cuh"

namespace rtb {
template cudaError_t launchRender<MATH_STRICT>(int, int, const void*, const RenderArgs&, int, int, cudaStream_t, int*);
template cudaError_t launchTrace<MATH_STRICT>(int, const void*, const TraceArgs&, int, cudaStream_t);
template cudaError_t launchRenderCoop<MATH_STRICT>(int, int, const void*, const RenderArgs&, int, cudaStream_t, int*);
template cudaError_t launchTraceCoop<MATH_STRICT>(int, const void*, const TraceArgs&, cudaStream_t);
template cudaError_t launchTexture<MATH_STRICT>(int, const void*, uint32_t, uint32_t, const double*, double*, cudaStream_t);

// device evaluation of the shared functions for the host/device identity test (tests/test_gpu_strict.py)
static __global__ void strictMathKernel(uint32_t fn, uint32_t n, const double* x, const double* y, double* out, int asFloat) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  if (asFloat) {
    const float a = (float)x[i], b = (float)y[i];
    out[i] = fn == 0 ? rtb_strict::sinS(a) : fn == 1 ? rtb_strict::cosS(a) : fn == 2 ? rtb_strict::atan2S(a, b) : fn == 3 ? rtb_strict::asinS(a) : rtb_strict::logS(a);
  } else {
    const double a = x[i], b = y[i];
    out[i] = fn == 0 ? rtb_strict::sinS(a) : fn == 1 ? rtb_strict::cosS(a) : fn == 2 ? rtb_strict::atan2S(a, b) : fn == 3 ? rtb_strict::asinS(a) : rtb_strict::logS(a);
  }
}
cudaError_t launchStrictMathEval(uint32_t fn, uint32_t n, const double* x, const double* y, double* out, int asFloat, cudaStream_t st) {
  strictMathKernel<<<(n + 255) / 256, 256, 0, st>>>(fn, n, x, y, out, asFloat);
  return cudaGetLastError();
}
}  // namespace rtb
