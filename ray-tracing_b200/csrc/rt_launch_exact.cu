// Literal-arithmetic instantiation of the kernels.  Build with --fmad=false (no FMA contraction), default
// IEEE division / sqrt: hit ids and traversal order are bit-exact against the same algorithm on the CPU.
#include "rt_kernels.cuh"

namespace rtb {
// albedoToColor / scaleColor (Lib.hs:287-292) applied to sum / ns
static __global__ void tonemapKernel(const double* sum, uint32_t n, double ns, uint8_t* out) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const double x = sqrt(sum[i] / ns);
  const double c = (x < 0.0) ? 0.0 : ((x > 0.999) ? 0.999 : x);
  const double f = floor(256.0 * c);
  out[i] = (f == f) ? (uint8_t)(int)f : (uint8_t)0;
}

template cudaError_t launchRender<false>(int, int, const void*, const RenderArgs&, int, int, cudaStream_t, int*);
template cudaError_t launchTrace<false>(int, const void*, const TraceArgs&, int, cudaStream_t);

cudaError_t launchTexture(int prec, const void* dscene, uint32_t tex, uint32_t n, const double* q, double* out, cudaStream_t st) {
  if (prec == 2) textureKernel<double><<<(n + 127) / 128, 128, 0, st>>>(*static_cast<const DScene<double>*>(dscene), tex, n, q, out);
  else textureKernel<float><<<(n + 127) / 128, 128, 0, st>>>(*static_cast<const DScene<float>*>(dscene), tex, n, q, out);
  return cudaGetLastError();
}
cudaError_t launchTonemap(const double* sum, uint32_t n, double ns, uint8_t* out, cudaStream_t st) {
  tonemapKernel<<<(n + 255) / 256, 256, 0, st>>>(sum, n, ns, out);
  return cudaGetLastError();
}
}  // namespace rtb
