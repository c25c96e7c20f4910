// rt_reduce.cu — the multi-GPU frame reduction of rt_render: ONE kernel on the device that owns the scene adds the
// other devices' partial radiance sums (double[H*W*3]) onto its own, reading them straight out of the peers' HBM over NVLink
// (peer access, 128-bit loads), in a fixed left-to-right order per element so the result does not depend on timing.
// HBM/NVLink-bound streaming: (n + 1) * 8 bytes read + 8 written per element; grid = a multiple of the SM count.
#include <algorithm>

#include "rt_launch.h"

namespace rtb {
struct PeerSrc { const double* p[15]; int n; };
static __global__ void __launch_bounds__(256) reducePeerKernel(double* __restrict__ dst, const PeerSrc src, size_t elems) {
  const size_t nt = (size_t)gridDim.x * blockDim.x, tid = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  const size_t pairs = elems / 2;
  for (size_t i = tid; i < pairs; i += nt) {
    double2 acc = reinterpret_cast<const double2*>(dst)[i];
#pragma unroll 1
    for (int k = 0; k < src.n; ++k) {
      const double2 v = __ldcs(reinterpret_cast<const double2*>(src.p[k]) + i);  // streamed once: do not keep peer lines in L1/L2
      acc.x += v.x; acc.y += v.y;
    }
    reinterpret_cast<double2*>(dst)[i] = acc;
  }
  if (tid == 0 && (elems & 1)) {
    double a = dst[elems - 1];
    for (int k = 0; k < src.n; ++k) a += src.p[k][elems - 1];
    dst[elems - 1] = a;
  }
}
cudaError_t launchReducePeer(double* dst, const double* const* src, int n, size_t elems, int numSMs, cudaStream_t st) {
  if (n <= 0 || elems == 0) return cudaSuccess;
  if (n > 15) return cudaErrorInvalidValue;
  PeerSrc s{};
  for (int k = 0; k < n; ++k) s.p[k] = src[k];
  s.n = n;
  const size_t pairs = (elems + 1) / 2;
  int blocks = (int)std::min<size_t>((pairs + 255) / 256, (size_t)numSMs * 8);
  if (blocks < 1) blocks = 1;
  reducePeerKernel<<<blocks, 256, 0, st>>>(dst, s, elems);
  return cudaGetLastError();
}
}  // namespace rtb
