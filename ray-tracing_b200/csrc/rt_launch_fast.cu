// Fast-math instantiation of the kernels (RT_FLAG_FAST_MATH): FMA contraction allowed, reciprocal slab test.
// Not bit-exact against the fp32 oracle; tests quantify the agreement.
#include "rt_kernels.cuh"

namespace rtb {
template cudaError_t launchRender<MATH_FAST>(int, int, const void*, const RenderArgs&, int, int, cudaStream_t, int*);
template cudaError_t launchTrace<MATH_FAST>(int, const void*, const TraceArgs&, int, cudaStream_t);
}  // namespace rtb
