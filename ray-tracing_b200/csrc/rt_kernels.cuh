// rt_kernels.cuh — the kernels of the B200 path-tracing core and their launchers.
//   renderKernel  : K2, persistent warps; every lane owns one camera path at a time and refills itself from the
//                   warp's chunk of the global path index space as soon as its path ends (path regeneration by
//                   warp ballot + prefix rank); one traversal site serves both the bounce query and the
//                   light-pdf queries of the Lambertian mixture pdf, so lanes in different shading phases still
//                   walk the BVH together.
//   traceKernel   : K1, `hit world ray t_min t_max` for explicit rays with the visit log (parity probe).
//   textureKernel : textureValue probe.   tonemapKernel : albedoToColor.
// Included by rt_launch_exact.cu (--fmad=false, FAST=false) and rt_launch_fast.cu (FAST=true).
#pragma once
#include "rt_device.cuh"
#include "rt_launch.h"

namespace rtb {

constexpr int kBlock = 128;

template <class T> struct MinBlocks { static constexpr int v = 4; };
template <> struct MinBlocks<double> { static constexpr int v = 1; };

template <class T, class S, int INSTR, bool FAST>
__global__ void __launch_bounds__(kBlock, MinBlocks<T>::v) renderKernel(const __grid_constant__ DScene<T> sc, const __grid_constant__ RenderArgs a) {
  extern __shared__ uint32_t smem[];
  Stack stk{smem + threadIdx.x, blockDim.x, 0};
  const unsigned lane = threadIdx.x & 31u;
  const unsigned ltMask = (1u << lane) - 1u;
  Instr<INSTR> ins;
  if constexpr (INSTR == 1) ins.init();

  enum { IDLE = 0, MAIN = 1, LPDF = 2 };
  int phase = IDLE;
  V3<T> o = mk<T>(0, 0, 0), d = mk<T>(0, 0, 1), beta = mk<T>(1, 1, 1);
  T time = 0; int depthLeft = 0; Rng rng{0, 1};
  double acc0 = 0, acc1 = 0, acc2 = 0; uint32_t accPix = kNone;
  T l_spdf = 0, l_v2 = 0, l_v1 = 0, curW = 1; uint32_t curLight = 0; int lsp = 0;
  uint32_t lref[kMaxLightDepth]; T lwt[kMaxLightDepth];
  unsigned long long cNext = 0, cEnd = 0, cBase = 0; uint32_t cPix0 = 0, cS0 = 0; bool exhausted = false;
  const uint32_t regionW = a.col1 - a.col0;
  const uint32_t lightsKind = sc.lights_root >> 28;

  auto flush = [&]() {
    if (accPix != kNone) {
      atomicAdd(a.out + (size_t)accPix * 3, acc0); atomicAdd(a.out + (size_t)accPix * 3 + 1, acc1); atomicAdd(a.out + (size_t)accPix * 3 + 2, acc2);
    }
  };
  auto finish = [&](const V3<T>& L) { acc0 += (double)L.x; acc1 += (double)L.y; acc2 += (double)L.z; phase = IDLE; };

  for (;;) {
    // ---- path regeneration: idle lanes take the next path indices of the warp's chunk, in lane order
    unsigned idle = __ballot_sync(0xffffffffu, phase == IDLE);
    while (idle) {
      if (cNext >= cEnd) {
        if (exhausted) break;
        unsigned long long base = 0;
        if (lane == 0) base = atomicAdd(a.counter, (unsigned long long)a.chunk);
        base = __shfl_sync(0xffffffffu, base, 0);
        if (base >= a.total_paths) { exhausted = true; break; }
        cBase = cNext = base;
        cEnd = (base + a.chunk < a.total_paths) ? base + a.chunk : a.total_paths;
        cPix0 = (uint32_t)(base / a.ns); cS0 = (uint32_t)(base % a.ns);
      }
      const unsigned rank = __popc(idle & ltMask);
      if (phase == IDLE && cNext + rank < cEnd) {
        // path index -> (pixel of the region, sample); consecutive indices = consecutive samples of one pixel
        const uint32_t off = cS0 + (uint32_t)(cNext - cBase) + rank;
        const uint32_t pl = cPix0 + off / a.ns, s = a.s0 + off % a.ns;
        const uint32_t row = a.row0 + pl / regionW, col = a.col0 + pl % regionW;
        const uint32_t pixel = row * a.W + col;
        if (pixel != accPix) { flush(); accPix = pixel; acc0 = acc1 = acc2 = 0; }
        rng = streamGen(a.seed, pixel, s);
        ins.count(0);
        // uniformRandomUVs (Lib.hs:1362-1365): u = (x + ru)/W, v = (y + rv)/H with y counted from the bottom row
        const T ru = rnd<T>(rng), rv = rnd<T>(rng);
        const T u = ((T)col + ru) / (T)a.W, v = ((T)(a.H - 1 - row) + rv) / (T)a.H;
        // getRay (Lib.hs:1253-1267): lens sample and shutter time are drawn for every ray
        const V3<T> rd = scale(sc.cam[21], randomInUnitDisk<T>(rng));
        const V3<T> cu = mk<T>(sc.cam[12], sc.cam[13], sc.cam[14]), cv = mk<T>(sc.cam[15], sc.cam[16], sc.cam[17]);
        const V3<T> offset = scale(rd.x, cu) + scale(rd.y, cv);
        const T r = rnd<T>(rng);
        time = sc.cam[22] + (sc.cam[23] - sc.cam[22]) * r;
        const V3<T> org = mk<T>(sc.cam[0], sc.cam[1], sc.cam[2]);
        o = org + offset;
        d = mk<T>(sc.cam[3], sc.cam[4], sc.cam[5]) + scale(u, mk<T>(sc.cam[6], sc.cam[7], sc.cam[8])) + scale(v, mk<T>(sc.cam[9], sc.cam[10], sc.cam[11])) - org - offset;
        beta = mk<T>(1, 1, 1); depthLeft = (int)a.maxDepth;
        phase = MAIN;
        if (depthLeft <= 0) finish(mk<T>(0, 0, 0));  // rayColor with depth <= 0 (Lib.hs:1303-1304)
      }
      const unsigned long long avail = cEnd - cNext; const unsigned n = __popc(idle);
      cNext += (n < avail) ? n : avail;
      idle = __ballot_sync(0xffffffffu, phase == IDLE);
    }
    if (__ballot_sync(0xffffffffu, phase != IDLE) == 0) break;

    if (phase != IDLE) {
      // ---- one query: the bounce's `hit world` (Lib.hs:1309) or one `hit htbl (Ray o v 0)` of htblPdfValue (677)
      HitInfo<T> hi;
      ins.count(13);
      if (phase == MAIN) ins.count(1);
      traverse<T, S, INSTR, FAST>(sc, stk, phase == MAIN ? sc.world_root : curLight, o, d, phase == MAIN ? time : (T)0, kEps<T>(), kInf<T>(), rng, hi, ins);

      bool scattered = false;  // true when the path continues with (o, d) as the next bounce ray
      if (phase == MAIN) {
        if (!hi.hit) {
          finish(beta * mk<T>(sc.bg[0], sc.bg[1], sc.bg[2]));  // Lib.hs:1310-1313
        } else {
          HitRec<T> rec;
          evalHit<T>(sc, o, d, time, hi, rec);
          const Q4<T> m = ldq(sc.mats + rec.mat);
          const uint32_t mkind = bitsOf(m.x), tex = bitsOf(m.y);
          if (mkind == M_LIGHT) {  // scatter = Nothing -> emitted, back face only (Lib.hs:860, 880-885)
            V3<T> e = mk<T>(0, 0, 0);
            if (!rec.front) e = textureValue<T, INSTR>(sc, tex, rec.u, rec.v, rec.p, ins);
            finish(beta * e);
          } else if (mkind == M_LAMBERTIAN) {  // Lib.hs:823-836 with pdfGenerate/pdfValue of the mixture (367-382)
            const V3<T> att = textureValue<T, INSTR>(sc, tex, rec.u, rec.v, rec.p, ins);
            const Onb<T> uvw = onbFromW(rec.n);
            const T rd = rnd<T>(rng);
            V3<T> pdfD;
            if (rd < (T)0.5) pdfD = htblRandom<T>(sc, sc.lights_root, rec.p, rng);
            else pdfD = onbLocal(uvw, randomCosineDirection<T>(rng));
            const V3<T> dir = unit(pdfD);
            const T cosine = dot(unit(dir), uvw.w);
            l_v2 = (cosine <= 0) ? (T)0 : cosine / kPi<T>();
            const T c2 = dot(rec.n, dir);  // scatteringPdf (Lib.hs:874-878)
            l_spdf = (c2 < 0) ? (T)0 : c2 / kPi<T>();
            beta = beta * att;
            o = rec.p; d = dir;
            if (lightsKind == K_UNHITTABLE) {  // htblPdfValue Unhittable = 0 without any traversal (Lib.hs:677-680, 1109)
              ins.count(13); ins.visit(sc.lights_root);
              const T pdfVal = (T)0.5 * ((T)0 + l_v2);
              beta = scale(l_spdf / pdfVal, beta);
              scattered = true;
            } else {
              phase = LPDF; l_v1 = 0; lsp = 0; curLight = sc.lights_root; curW = 1;
            }
          } else if (mkind == M_METAL) {  // Lib.hs:837-841: never absorbs, fuzz unclamped
            const V3<T> rUnit = randomUnitVector<T>(rng);
            const V3<T> reflected = reflectV(unit(d), rec.n);
            beta = beta * textureValue<T, INSTR>(sc, tex, rec.u, rec.v, rec.p, ins);
            o = rec.p; d = reflected + scale(m.z, rUnit);
            scattered = true;
          } else if (mkind == M_DIELECTRIC) {  // Lib.hs:842-859: xi is always drawn
            const T eta = rec.front ? (T)1.0 / m.z : m.z;
            const V3<T> ud = unit(d);
            const T cosTheta = hmin(dot(neg(ud), rec.n), (T)1.0);
            const T sinTheta = msqrt((T)1.0 - cosTheta * cosTheta);
            const T rd = rnd<T>(rng);
            if ((eta * sinTheta > (T)1.0) || rd < schlick<T>(cosTheta, eta)) d = reflectV(ud, rec.n);
            else d = refractV(ud, rec.n, eta);
            o = rec.p;
            scattered = true;
          } else {  // Isotropic (Lib.hs:861-865) weighted by the generic scatteringPdf / 1.0 (867-878, 1329-1333)
            const V3<T> randDr = randomInUnitSphere<T>(rng);
            const V3<T> att = textureValue<T, INSTR>(sc, tex, rec.u, rec.v, rec.p, ins);
            const T c2 = dot(rec.n, randDr);
            const T spdf = (c2 < 0) ? (T)0 : c2 / kPi<T>();
            beta = scale(spdf / (T)1.0, beta * att);
            o = rec.p; d = randDr;
            scattered = true;
          }
        }
      } else {  // LPDF: htblPdfValue (Lib.hs:673-705) for `curLight`, walked depth-first, left first
        if (hi.hit) {
          const uint32_t lk = curLight >> 28, li = curLight & 0x0fffffffu;
          if (lk == K_RECT_XZ) {
            const Q4<T> ra = ldq(sc.rects + 2 * (size_t)li);
            const T area = (ra.y - ra.x) * (ra.w - ra.z);
            const T distanceSquared = hi.t * hi.t * sqlen(d);
            const T cosine = mabs(d.y / len(d));  // |v . (0,+-1,0)| / |v|
            l_v1 = l_v1 + curW * (distanceSquared / (cosine * area));
          } else if (lk == K_SPHERE) {
            const Q4<T> q = ldq(sc.spheres + li);
            const T cosThetaMax = msqrt(1 - q.w * q.w / sqlen(mk<T>(q.x, q.y, q.z) - o));
            const T solidAngle = 2 * kPi<T>() * (1 - cosThetaMax);
            l_v1 = l_v1 + curW * (1 / solidAngle);
          } else if (lk == K_BVH) {
            const uint4 sz = __ldg(sc.node_sz + li);
            const uint32_t lc = bitsOf(ldq(sc.nodes + 2 * (size_t)li).w), rc = bitsOf(ldq(sc.nodes + 2 * (size_t)li + 1).w);
            if (lsp + 2 <= kMaxLightDepth) {
              lref[lsp] = rc; lwt[lsp] = curW * ((T)sz.z / (T)sz.x); ++lsp;
              lref[lsp] = lc; lwt[lsp] = curW * ((T)sz.y / (T)sz.x); ++lsp;
            }
          }
        }
        if (lsp > 0) { --lsp; curLight = lref[lsp]; curW = lwt[lsp]; }
        else {
          const T pdfVal = (T)0.5 * (l_v1 + l_v2);
          beta = scale(l_spdf / pdfVal, beta);
          phase = MAIN; scattered = true;
        }
      }
      if (scattered) {  // rayColorHelp newRay (d - 1) (Lib.hs:1321-1333)
        --depthLeft;
        if (depthLeft <= 0) finish(beta * mk<T>(0, 0, 0));
      }
    }
  }
  flush();
  if constexpr (INSTR == 1) {
    for (int i = 0; i < 16; ++i) if (ins.c[i]) atomicAdd(a.ctr + i, (unsigned long long)ins.c[i]);
  }
}

// K1
template <class T, class S, bool FAST>
__global__ void __launch_bounds__(kBlock) traceKernel(const __grid_constant__ DScene<T> sc, const __grid_constant__ TraceArgs a) {
  extern __shared__ uint32_t smem[];
  Stack stk{smem + threadIdx.x, blockDim.x, 0};
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= a.n) return;
  const double* r = a.rays + (size_t)i * 7;
  const V3<T> o = mk<T>((T)r[0], (T)r[1], (T)r[2]), d = mk<T>((T)r[3], (T)r[4], (T)r[5]);
  const T time = (T)r[6];
  Rng rng = streamGen(a.seed, i, 0);
  Instr<2> ins; ins.init(a.vlog + (size_t)i * a.logStride, a.logStride);
  for (uint32_t k = 0; k < a.logStride; ++k) a.vlog[(size_t)i * a.logStride + k] = kNone;
  HitInfo<T> hi;
  traverse<T, S, 2, FAST>(sc, stk, sc.world_root, o, d, time, (T)a.tmin, (T)a.tmax, rng, hi, ins);
  a.nvisit[i] = ins.n; a.vhash[i] = ins.h;
  a.hit[i] = hi.hit ? 1 : 0;
  double* out = a.rec + (size_t)i * 9;
  if (hi.hit) {
    HitRec<T> rec;
    evalHit<T>(sc, o, d, time, hi, rec);
    out[0] = rec.t; out[1] = rec.p.x; out[2] = rec.p.y; out[3] = rec.p.z; out[4] = rec.n.x; out[5] = rec.n.y; out[6] = rec.n.z; out[7] = rec.u; out[8] = rec.v;
    a.prim[i] = hi.prim; a.mat[i] = rec.mat; a.front[i] = rec.front ? 1 : 0;
  } else {
    for (int k = 0; k < 9; ++k) out[k] = 0;
    a.prim[i] = 0; a.mat[i] = kNone; a.front[i] = 0;
  }
}

template <class T>
__global__ void textureKernel(const __grid_constant__ DScene<T> sc, uint32_t tex, uint32_t n, const double* q, double* out) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  Instr<0> ins;
  const double* a = q + (size_t)i * 5;
  V3<T> c = textureValue<T, 0>(sc, tex, (T)a[0], (T)a[1], mk<T>((T)a[2], (T)a[3], (T)a[4]), ins);
  out[i * 3] = c.x; out[i * 3 + 1] = c.y; out[i * 3 + 2] = c.z;
}

// ---- launchers -------------------------------------------------------------------------------------
template <class T, class S, int INSTR, bool FAST>
static cudaError_t launchRenderT(const DScene<T>& sc, const RenderArgs& a, int stackDepth, int numSMs, cudaStream_t st, int* gridOut) {
  auto kern = renderKernel<T, S, INSTR, FAST>;
  size_t smem = (size_t)stackDepth * kBlock * sizeof(uint32_t);
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  int perSM = 0;
  e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&perSM, kern, kBlock, smem);
  if (e != cudaSuccess) return e;
  if (perSM < 1) perSM = 1;
  int grid = numSMs * perSM;  // persistent: exactly one resident wave
  if (gridOut) *gridOut = grid;
  kern<<<grid, kBlock, smem, st>>>(sc, a);
  return cudaGetLastError();
}

template <bool FAST>
cudaError_t launchRender(int prec, int instr, const void* dscene, const RenderArgs& a, int stackDepth, int numSMs, cudaStream_t st, int* gridOut) {
  if (prec == 0) {
    const auto& sc = *static_cast<const DScene<float>*>(dscene);
    return instr ? launchRenderT<float, float, 1, FAST>(sc, a, stackDepth, numSMs, st, gridOut) : launchRenderT<float, float, 0, FAST>(sc, a, stackDepth, numSMs, st, gridOut);
  } else if (prec == 1) {
    const auto& sc = *static_cast<const DScene<float>*>(dscene);
    return instr ? launchRenderT<float, double, 1, FAST>(sc, a, stackDepth, numSMs, st, gridOut) : launchRenderT<float, double, 0, FAST>(sc, a, stackDepth, numSMs, st, gridOut);
  } else {
    const auto& sc = *static_cast<const DScene<double>*>(dscene);
    return instr ? launchRenderT<double, double, 1, FAST>(sc, a, stackDepth, numSMs, st, gridOut) : launchRenderT<double, double, 0, FAST>(sc, a, stackDepth, numSMs, st, gridOut);
  }
}

template <class T, class S, bool FAST>
static cudaError_t launchTraceT(const DScene<T>& sc, const TraceArgs& a, int stackDepth, cudaStream_t st) {
  auto kern = traceKernel<T, S, FAST>;
  size_t smem = (size_t)stackDepth * kBlock * sizeof(uint32_t);
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  kern<<<(a.n + kBlock - 1) / kBlock, kBlock, smem, st>>>(sc, a);
  return cudaGetLastError();
}
template <bool FAST>
cudaError_t launchTrace(int prec, const void* dscene, const TraceArgs& a, int stackDepth, cudaStream_t st) {
  if (prec == 0) return launchTraceT<float, float, FAST>(*static_cast<const DScene<float>*>(dscene), a, stackDepth, st);
  if (prec == 1) return launchTraceT<float, double, FAST>(*static_cast<const DScene<float>*>(dscene), a, stackDepth, st);
  return launchTraceT<double, double, FAST>(*static_cast<const DScene<double>*>(dscene), a, stackDepth, st);
}

}  // namespace rtb
