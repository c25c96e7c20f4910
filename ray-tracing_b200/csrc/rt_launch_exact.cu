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

// Decision equality of the interval-filtered slab test (slabFiltered, rt_device.cuh) with boxRayIntersect evaluated by three IEEE
// divisions (Lib.hs:798-814), over random (box, ray, interval) triples: a third generic (ray through / past a random box), two thirds
// adversarial -- a box face placed so that its quotient lands within 32 ulps of t_min or t_max, boxes a few ulps thin (a Rect's
// bounding box at Cornell scale), near-zero numerators.  out3 = {triples in the hoisted-reciprocal range, of those decided by the
// filter alone, filter decisions that differ from the IEEE decision}.  The last must be 0.
static __global__ void boxFilterSelftestKernel(unsigned long long n, unsigned long long seed, unsigned long long* out3) {
  const unsigned long long tid = blockIdx.x * (unsigned long long)blockDim.x + threadIdx.x, nthreads = (unsigned long long)gridDim.x * blockDim.x;
  Rng g = streamGen(seed, (uint32_t)tid, (uint32_t)(tid >> 32));
  unsigned long long checked = 0, decided = 0, bad = 0;
  auto unit = [&]() { return (float)(nextWord(g) >> 40) * (1.0f / 16777216.0f); };                      // [0, 1)
  auto nudge = [&](float x, int k) { return __uint_as_float(__float_as_uint(x) + (uint32_t)k); };        // k ulps away from zero / towards it
  for (unsigned long long i = tid; i < n; i += nthreads) {
    const unsigned long long w = nextWord(g);
    const float scale = __uint_as_float((uint32_t)(120u + (w & 15u)) << 23);                              // 2^-7 .. 2^8
    float o[3], d[3], lo[3], hi[3];
    for (int k = 0; k < 3; ++k) {
      o[k] = (unit() - 0.5f) * 2.0f * scale;
      const float mag = __uint_as_float((uint32_t)(112u + ((w >> (8 + 5 * k)) & 15u)) << 23);             // 2^-15 .. 2^0
      d[k] = (unit() + 0.5f) * mag * (((w >> (4 + k)) & 1u) ? -1.0f : 1.0f);
    }
    const float tstar = unit() * 4.0f * scale + 1e-3f;
    float t_min = ((w >> 24) & 1u) ? 1e-4f : 0.0f, t_max = ((w >> 25) & 3u) ? tstar * (0.25f + 2.0f * unit()) : __int_as_float(0x7f800000);
    for (int k = 0; k < 3; ++k) {
      const float c = o[k] + d[k] * tstar * (0.5f + unit()), e = unit() * 0.25f * scale;
      lo[k] = c - e; hi[k] = c + e;
    }
    const uint32_t kind = (uint32_t)((w >> 28) % 6u), ax = (uint32_t)((w >> 32) % 3u);
    const int ulps = (int)((w >> 36) & 63u) - 32;  // +-32 ulps = +-3.8e-6 relative: both sides of the filter's threshold
    if (kind == 1u) { const float q = __fdiv_rn(lo[ax] - o[ax], d[ax]); if (q > 0.0f) t_max = nudge(q, ulps); }          // near face at t_max
    else if (kind == 2u) { const float q = __fdiv_rn(hi[ax] - o[ax], d[ax]); if (q > 0.0f) t_min = nudge(q, ulps); }     // far face at t_min
    else if (kind == 3u) hi[ax] = nudge(lo[ax], (int)((w >> 36) & 7u));                                                   // 0 .. 7 ulps thin
    else if (kind == 4u) { lo[ax] = o[ax]; if ((w >> 40) & 1u) hi[ax] = nudge(o[ax], 1 + (int)((w >> 36) & 3u)); }          // exact-zero numerator
    else if (kind == 5u) { const float q = __fdiv_rn(hi[ax] - o[ax], d[ax]); if (q > 0.0f) t_max = nudge(q, ulps); }     // far face at t_max
    if (hi[ax] < lo[ax]) { const float t = lo[ax]; lo[ax] = hi[ax]; hi[ax] = t; }
    // a quarter of the triples have one or two zero direction components (the reference's (1, 0, 0) light direction): division mode 1
    if (((w >> 44) & 3u) == 0u) { d[(ax + 1u) % 3u] = ((w >> 46) & 1u) ? -0.0f : 0.0f; if ((w >> 47) & 1u) d[(ax + 2u) % 3u] = 0.0f; }
    if (kind == 4u && ((w >> 48) & 1u)) lo[(ax + 1u) % 3u] = o[(ax + 1u) % 3u];   // 0 * inf = NaN on a zero-divisor axis
    for (int k = 0; k < 3; ++k) if (hi[k] < lo[k]) { const float t = lo[k]; lo[k] = hi[k]; hi[k] = t; }
    bool ok = true;
    float rr[3];
    for (int k = 0; k < 3; ++k) { uint32_t mode; axisMode(d[k], true, rr[k], mode); ok = ok && mode != 2u; }
    if (!ok) continue;
    ++checked;
    bool sure = true, pass = true, ref = true;
    for (int k = 0; k < 3; ++k) {
      pass = slabFiltered(lo[k], hi[k], o[k], rr[k], t_min, fminf(t_max, 3.4028234663852886e38f), sure) && pass;  // as stepNode calls it
      ref = slabDecide<float>(__fdiv_rn(lo[k] - o[k], d[k]), __fdiv_rn(hi[k] - o[k], d[k]), t_min, t_max) && ref;
    }
    if (sure) { ++decided; if (pass != ref) ++bad; }
  }
  atomicAdd(out3, checked); atomicAdd(out3 + 1, decided); atomicAdd(out3 + 2, bad);
}
cudaError_t launchBoxFilterSelftest(unsigned long long n, unsigned long long seed, unsigned long long* out3, cudaStream_t st) {
  boxFilterSelftestKernel<<<148 * 8, 256, 0, st>>>(n, seed, out3);
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
