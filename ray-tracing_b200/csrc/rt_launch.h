// rt_launch.h — argument blocks and launcher declarations shared by the kernel TUs and the C-ABI host code.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace rtb {

struct RenderArgs {
  uint32_t W, H;                   // full image
  uint32_t s0, ns;                 // sample range [s0, s0+ns)
  uint32_t maxDepth;
  uint32_t row0, row1, col0, col1; // region (rows counted from the top)
  uint32_t chunk;                  // path indices a warp takes per atomic
  uint32_t stackCap;               // traversal stack entries per thread (set by the launcher)
  uint32_t lightCap;               // entries of htblPdfValue's walk stack per thread: levels of the lights BVH, 0 when the light list is no BVH
  uint32_t shadeMin, rectMin, rareMin;  // full vote: run the shade / rect+cuboid / rare class as soon as this many lanes wait for it
  uint32_t keepNode, keepLeaf;     // run a node / sphere step without a full vote when at least this many lanes wait for one
  unsigned long long seed;
  unsigned long long total_paths;  // region pixels * ns
  unsigned long long* counter;     // device: next free path index
  unsigned long long* ctr;         // device: 16 algorithmic-work counters (INSTR=1) or null
  double* out;                     // device: double[H*W*3] radiance sums, accumulated with atomicAdd
};

struct TraceArgs {
  uint32_t n, logStride;
  uint32_t stackCap;
  double tmin, tmax;
  unsigned long long seed;
  const double* rays;  // device [n][7]
  int32_t* hit; double* rec; uint32_t* prim; uint32_t* mat; uint8_t* front;
  unsigned long long* nvisit; unsigned long long* vhash; uint32_t* vlog;
};

// prec: 0 fp32, 1 mixed, 2 fp64 ; dscene points at a host copy of DScene<float> (prec 0/1) or DScene<double> (2)
template <bool FAST> cudaError_t launchRender(int prec, int instr, const void* dscene, const RenderArgs& a, int stackDepth, int numSMs, cudaStream_t st, int* gridOut);
template <bool FAST> cudaError_t launchTrace(int prec, const void* dscene, const TraceArgs& a, int stackDepth, cudaStream_t st);
cudaError_t launchTexture(int prec, const void* dscene, uint32_t tex, uint32_t n, const double* q, double* out, cudaStream_t st);
cudaError_t launchDivSelftest(unsigned long long n, unsigned long long seed, unsigned long long* out2, cudaStream_t st);
cudaError_t launchTonemap(const double* sum, uint32_t n, double ns, uint8_t* out, cudaStream_t st);

}  // namespace rtb
