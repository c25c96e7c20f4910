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

// divByRcp(a, d, rcpRefined(d)) against the IEEE quotient over random operands of the validated range (rt_device.cuh)
static __global__ void divSelftestKernel(unsigned long long n, unsigned long long seed, unsigned long long* out2) {
  const unsigned long long tid = blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x, nthreads = (unsigned long long)gridDim.x * blockDim.x;
  Rng g = streamGen(seed, (uint32_t)tid, (uint32_t)(tid >> 32));
  unsigned long long checked = 0, bad = 0;
  for (unsigned long long i = tid; i < n; i += nthreads) {
    const unsigned long long w = nextWord(g), w2 = nextWord(g);
    // d: sign, exponent 2^-40 .. 2^39, random mantissa; a: sign, exponent 2^-90 .. 2^40, random mantissa
    const uint32_t db = ((uint32_t)(w >> 63) << 31) | ((87u + (uint32_t)((w >> 32) % 80u)) << 23) | ((uint32_t)w & 0x7fffffu);
    const uint32_t ab = ((uint32_t)(w2 >> 63) << 31) | ((37u + (uint32_t)((w2 >> 32) % 131u)) << 23) | ((uint32_t)w2 & 0x7fffffu);
    float d = __uint_as_float(db), a = __uint_as_float(ab);
    if ((i & 7u) == 0u) a = d * (float)(int)((w2 >> 40) & 1023u);  // exact and near-exact quotients
    if (!(fabsf(d) >= kDivLo && fabsf(d) <= kDivHi)) continue;
    const float q = divByRcp(a, d, rcpRefined(d));
    if (!(fabsf(q) >= kDivLo)) continue;  // outside the validated range: the kernels take the generic division
    ++checked;
    if (__float_as_uint(q) != __float_as_uint(__fdiv_rn(a, d))) ++bad;
  }
  atomicAdd(out2, checked); atomicAdd(out2 + 1, bad);
}
cudaError_t launchDivSelftest(unsigned long long n, unsigned long long seed, unsigned long long* out2, cudaStream_t st) {
  divSelftestKernel<<<148 * 8, 256, 0, st>>>(n, seed, out2);
  return cudaGetLastError();
}

// ---- wavefront launcher (literal arithmetic only)
size_t wfLayout(uint32_t P, char* base, WfState* out) {
  size_t off = 0;
  auto take = [&](size_t bytes) { char* p = base ? base + off : nullptr; off += (bytes + 255) & ~(size_t)255; return p; };
  WfState w{};
  w.P = P;
  float** f[] = {&w.ox, &w.oy, &w.oz, &w.dx, &w.dy, &w.dz, &w.time, &w.bx, &w.by, &w.bz, &w.v2, &w.spdf, &w.rT, &w.rW};
  for (float** q : f) *q = reinterpret_cast<float*>(take((size_t)P * 4));
  w.rngS = reinterpret_cast<unsigned long long*>(take((size_t)P * 8)); w.rngG = reinterpret_cast<unsigned long long*>(take((size_t)P * 8));
  uint32_t** u[] = {&w.depth, &w.phase, &w.rPrim, &w.rFlags, &w.accPix, &w.chunkLeft};
  for (uint32_t** q : u) *q = reinterpret_cast<uint32_t*>(take((size_t)P * 4));
  w.rChain = reinterpret_cast<uint32_t*>(take((size_t)P * 4 * kMaxFrames));
  w.acc = reinterpret_cast<double*>(take((size_t)P * 8 * 3));
  w.chunkNext = reinterpret_cast<unsigned long long*>(take((size_t)P * 8));
  w.ctl = reinterpret_cast<unsigned long long*>(take(kWfCtlWords * 8));
  if (out) *out = w;
  return off;
}
template <class S>
static cudaError_t launchWavefrontT(const DScene<float>& sc, const RenderArgs& a, const WfState& w, int stackDepth, int numSMs, int rounds, bool init, cudaStream_t st, int* gridOut) {
  auto trav = wfTraverseKernel<float, S>;
  auto shade = wfShadeKernel<float>;
  const size_t smemT = (size_t)columnWords<float>((uint32_t)stackDepth, a.lightCap) * kBlock * sizeof(uint32_t) + (kBlock / 32) * 16;
  const size_t smemS = ((size_t)extraWords<float>() * kBlock * (kShadeBlock / kBlock) + kShadeBlock + 8) * sizeof(uint32_t);
  cudaError_t e = cudaFuncSetAttribute(trav, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smemT);
  if (e != cudaSuccess) return e;
  e = cudaFuncSetAttribute(shade, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smemS);
  if (e != cudaSuccess) return e;
  int perSM = 0;
  e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSM, trav, kBlock, smemT);
  if (e != cudaSuccess) return e;
  if (perSM < 1) perSM = 1;
  if (const char* v = std::getenv("RTB_WF_BLOCKS_PER_SM")) perSM = std::max(1, std::min(perSM, std::atoi(v)));
  const int grid = std::max(1, std::min(numSMs * perSM, (int)((w.P + kBlock - 1) / kBlock)));
  if (gridOut) *gridOut = grid;
  RenderArgs b = a; b.stackCap = (uint32_t)stackDepth;
  if (init) wfInitKernel<<<(w.P + 255) / 256, 256, 0, st>>>(w);
  for (int r = 0; r < rounds; ++r) {
    shade<<<(w.P + kShadeBlock - 1) / kShadeBlock, kShadeBlock, smemS, st>>>(sc, b, w);
    trav<<<grid, kBlock, smemT, st>>>(sc, b, w);
  }
  return cudaGetLastError();
}
cudaError_t launchWavefront(int prec, const void* dscene, const RenderArgs& a, const WfState& w, int stackDepth, int numSMs, int rounds, bool init, cudaStream_t st, int* gridOut) {
  const auto& sc = *static_cast<const DScene<float>*>(dscene);
  if (prec == 0) return launchWavefrontT<float>(sc, a, w, stackDepth, numSMs, rounds, init, st, gridOut);
  if (prec == 1) return launchWavefrontT<double>(sc, a, w, stackDepth, numSMs, rounds, init, st, gridOut);
  return cudaErrorInvalidValue;
}

template cudaError_t launchRender<MATH_LITERAL>(int, int, const void*, const RenderArgs&, int, int, cudaStream_t, int*);
template cudaError_t launchTrace<MATH_LITERAL>(int, const void*, const TraceArgs&, int, cudaStream_t);
template cudaError_t launchRenderCoop<MATH_LITERAL>(int, int, const void*, const RenderArgs&, int, cudaStream_t, int*);
template cudaError_t launchTraceCoop<MATH_LITERAL>(int, const void*, const TraceArgs&, cudaStream_t);
template cudaError_t launchTexture<MATH_LITERAL>(int, const void*, uint32_t, uint32_t, const double*, double*, cudaStream_t);

cudaError_t launchTonemap(const double* sum, uint32_t n, double ns, uint8_t* out, cudaStream_t st) {
  tonemapKernel<<<(n + 255) / 256, 256, 0, st>>>(sum, n, ns, out);
  return cudaGetLastError();
}
}  // namespace rtb
