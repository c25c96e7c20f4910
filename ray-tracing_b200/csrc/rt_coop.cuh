// rt_coop.cuh — warp-cooperative traversal of ONE path (cfg5: a BVH of ~10^6 spheres).  Included by rt_kernels.cuh in namespace rtb.
//
// Why: the reference's boxRayIntersect tests the three slabs independently (Lib.hs:798-814), so almost every box in
// front of a ray passes and reference-order traversal visits O(N) nodes and leaves per query (~1.8 M `hit` calls per camera sample
// at 10^6 spheres).  Visit order is part of the parity contract, so the work cannot be pruned — but it can be spread over a warp.
// With one path per lane (renderKernel) a path is a chain of ~10^6 dependent steps: a bounded sample of the frame is bound by the
// latency of its slowest paths, not by throughput.  Here a warp owns ONE path and its 32 lanes test 32 consecutive records of the
// BVH laid out in reference (pre-order, left first) visit order:
//     record i of the scan stream = a BVHNode (box + index of the first record after its subtree) or a leaf (sphere / rect);
//     sequential semantics (Lib.hs:971-988): node box passes [t_min, t_best] -> next record; fails -> jump past the subtree;
//     leaf accepted -> t_best := t, record replaces the hit (any later accepted leaf replaces it: reference order, not nearest).
// All 32 lanes evaluate their record against the CURRENT t_best; the records up to and including the first lane whose outcome
// changes the sequential state (a failing box, an accepted leaf) are committed in order, a failing box whose subtree ends inside
// the window just kills the lanes of that subtree and the scan goes on in the same window; an accepted leaf restarts the window
// behind it with the new t_best.  The committed records are exactly the reference's visits, in its order (tests: visit count,
// hash and log bit-identical to the oracle's and to the one-path-per-lane kernels').
// Loads are one contiguous 1 KB line group per window (32 x 32 B records, 128-bit loads): no stack, no shared memory.
// Shading runs uniformly on all lanes (identical state, identical draws): a bounce's shading is ~10^3 instructions against
// ~10^5..10^6 for its traversal.  Eligible scenes: world = a BVH whose leaves are spheres / rects / Unhittable only, light list =
// one rect, one sphere or Unhittable (rt_abi.cu decides and builds the stream at upload).
#pragma once

namespace rtb {

constexpr unsigned kFullMask = 0xffffffffu;
RTB_D unsigned laneRange(uint32_t a, uint32_t b) {  // bits a..b inclusive, a <= b <= 31
  return (kFullMask >> (31u - b)) & (kFullMask << a);
}

template <class T> struct CoopQuery {
  V3<T> o, d, inv; T tmin, tbest; uint32_t flags, bprim; bool hit;
};

template <class T, int MODE> RTB_D void coopBegin(const DScene<T>& sc, CoopQuery<T>& q, const V3<T>& o, const V3<T>& d, T t_min, T t_max) {
  q.o = o; q.d = d; q.tmin = t_min; q.tbest = t_max; q.bprim = 0; q.hit = false;
  q.flags = (t_max == t_max) ? 0u : F_NANT;
  q.inv = mk<T>(0, 0, 0);
  if constexpr (MODE == MATH_FAST) {
    q.inv = mk<T>((T)1 / d.x, (T)1 / d.y, (T)1 / d.z);
  } else if constexpr (sizeof(T) == 4) {  // the division modes of the literal slab test (rt_traverse.cuh setInv)
    const bool oSafe = !sc.bigCoords && !(q.flags & F_NANT) && fabsf(o.x) <= kDivHi && fabsf(o.y) <= kDivHi && fabsf(o.z) <= kDivHi;
    uint32_t mx, my, mz;
    axisMode(d.x, oSafe, q.inv.x, mx); axisMode(d.y, oSafe, q.inv.y, my); axisMode(d.z, oSafe, q.inv.z, mz);
    q.flags |= (mx | (my << 2) | (mz << 4)) << F_MODE_SHIFT;
  }
}

template <class T, int MODE> RTB_D bool coopBoxPass(const Q4<T>& lo, const Q4<T>& hi, const CoopQuery<T>& q) {
  if constexpr (MODE != MATH_FAST && sizeof(T) == 4) {
    bool ok = (q.flags & F_MODE_GENERIC) == 0u;
    const bool bx = slabHoisted(lo.x, hi.x, q.o.x, q.d.x, q.inv.x, q.tmin, q.tbest, ok);
    const bool by = slabHoisted(lo.y, hi.y, q.o.y, q.d.y, q.inv.y, q.tmin, q.tbest, ok);
    const bool bz = slabHoisted(lo.z, hi.z, q.o.z, q.d.z, q.inv.z, q.tmin, q.tbest, ok);
    if (!ok) return boxFallback(lo.x, lo.y, lo.z, hi.x, hi.y, hi.z, q.o.x, q.o.y, q.o.z, q.d.x, q.d.y, q.d.z, q.inv.x, q.inv.y, q.inv.z,
                                (q.flags & F_MODE_GENERIC) != 0u, q.tmin, q.tbest);
    return bx && by && bz;
  } else {
    const bool bx = slab<T, MODE == MATH_FAST>(lo.x, hi.x, q.o.x, q.d.x, q.inv.x, q.tmin, q.tbest);
    const bool by = slab<T, MODE == MATH_FAST>(lo.y, hi.y, q.o.y, q.d.y, q.inv.y, q.tmin, q.tbest);
    const bool bz = slab<T, MODE == MATH_FAST>(lo.z, hi.z, q.o.z, q.d.z, q.inv.z, q.tmin, q.tbest);
    return bx && by && bz;
  }
}

// instrumentation of the committed lanes of a window, in lane (= visit) order
template <int INSTR> RTB_D void coopCommit(Instr<INSTR>& ins, unsigned committed, uint32_t ref, unsigned nodeMask) {
  if constexpr (INSTR == 1) {
    // every lane holds the same counters (uniform execution); lane 0 publishes them at the end of the kernel
    const uint32_t k = ref >> 28;
    ins.c[14] += __popc(committed);
    ins.c[2] += __popc(committed & nodeMask);
    ins.c[3] += __popc(committed & __ballot_sync(kFullMask, k == K_SPHERE));
    ins.c[5] += __popc(committed & __ballot_sync(kFullMask, k >= K_RECT_XY && k <= K_RECT_YZ));
    ins.c[15] += __popc(committed & __ballot_sync(kFullMask, k == K_UNHITTABLE)) ;
  } else if constexpr (INSTR == 2) {
    for (unsigned m = committed; m; m &= m - 1u) ins.visit(__shfl_sync(kFullMask, ref, __ffs((int)m) - 1));
  }
}

#ifndef RTB_COOP_PF
#define RTB_COOP_PF 0  /* measured (profiles/r2_coop_prefetch_ab.log): 1 / 2 / 4 windows ahead lose 5 / 6 / 7 % at 90 k spheres, 2-4 % at 1 M */
#endif
constexpr uint32_t kCoopPrefetch = RTB_COOP_PF;  // windows of 32 records
// `hit world ray t_min t_max` over the scan stream; the whole warp calls it with identical arguments
template <class T, class S, int INSTR, int MODE>
RTB_D void coopTraverse(const DScene<T>& sc, CoopQuery<T>& q, Instr<INSTR>& ins) {
  const uint32_t lane = threadIdx.x & 31u;
  const uint32_t n = sc.coop_n;
  uint32_t c = 0;
  while (c < n) {
    const uint32_t i = c + lane;
    const bool valid = i < n;
    Q4<T> r0{0, 0, 0, 0}, r1{0, 0, 0, 0};
    if (valid) { r0 = ldq(sc.coop + 2 * (size_t)i); r1 = ldq(sc.coop + 2 * (size_t)i + 1); }
    // the scan is sequential unless a box fails or a leaf is accepted: pull this lane's record of the window kCoopPrefetch ahead into L1
    // (one 32-byte sector = both loads) while this window is tested -- the kernel is bound by the L2 latency of these loads
    if constexpr (kCoopPrefetch > 0) {
      if (i + 32u * kCoopPrefetch < n) asm volatile("prefetch.global.L1 [%0];" ::"l"(sc.coop + 2 * ((size_t)i + 32u * kCoopPrefetch)));
    }
    const uint32_t ref = valid ? bitsOf(r1.w) : (K_UNHITTABLE << 28);
    const uint32_t kind = ref >> 28;
    bool ev = false; T tHit = 0; uint32_t skip = 0;
    if (valid) {
      if (kind == K_BVH) { skip = bitsOf(r0.w); ev = !coopBoxPass<T, MODE>(r0, r1, q); }
      else if (kind == K_SPHERE) ev = sphereT<T, S>(mk<T>(r0.x, r0.y, r0.z), r0.w, q.o, q.d, q.tmin, q.tbest, tHit);
      else if (kind >= K_RECT_XY && kind <= K_RECT_YZ) ev = rectT<T>(kind - K_RECT_XY, r0.x, r0.y, r0.z, r0.w, r1.x, q.o, q.d, q.tmin, q.tbest, tHit);
    }
    unsigned alive = __ballot_sync(kFullMask, valid);
    const unsigned mev = __ballot_sync(kFullMask, ev);
    const unsigned nodeMask = __ballot_sync(kFullMask, valid && kind == K_BVH);
    unsigned committed = 0;
    uint32_t pos = 0;
    for (;;) {
      const unsigned m = mev & alive & (kFullMask << pos);
      if (!m) { committed |= alive & (kFullMask << pos); c += 32u; break; }
      const uint32_t f = (uint32_t)__ffs((int)m) - 1u;
      committed |= alive & laneRange(pos, f);
      if ((nodeMask >> f) & 1u) {  // a failing box: skip its subtree
        const uint32_t s = __shfl_sync(kFullMask, skip, f);
        if (s >= c + 32u) { c = s; break; }
        pos = s - c;  // > f; lanes f+1 .. pos-1 belong to the skipped subtree
        if (pos > f + 1u) alive &= ~laneRange(f + 1u, pos - 1u);
      } else {                     // an accepted leaf: it replaces the hit, everything behind it sees the new t_best
        q.tbest = __shfl_sync(kFullMask, tHit, f);
        q.bprim = __shfl_sync(kFullMask, ref, f);
        q.hit = true;
        if constexpr (INSTR == 1) { if ((q.bprim >> 28) == K_SPHERE) ins.sphereHit(); }
        if (!(q.tbest == q.tbest)) q.flags |= F_NANT | F_MODE_GENERIC;  // a NaN t_max: only the literal slab test is right from here on
        c += f + 1u;
        break;
      }
    }
    coopCommit<INSTR>(ins, committed, ref, nodeMask);
  }
}

// htblPdfValue (Lib.hs:673-705) of a light list that is ONE primitive (rect / sphere): `hit light (Ray o v 0) eps inf`, then the formula
template <class T, class S, int INSTR, int MODE>
RTB_D T coopLightPdf(const DScene<T>& sc, const V3<T>& o, const V3<T>& v, Instr<INSTR>& ins) {
  const uint32_t root = sc.lights_root, kind = root >> 28, idx = root & kIdxMask;
  ins.count(13); ins.visit(root);
  T t;
  if (kind == K_RECT_XZ || kind == K_RECT_XY || kind == K_RECT_YZ) {
    const Q4<T> a = ldq(sc.rects + 2 * (size_t)idx), b = ldq(sc.rects + 2 * (size_t)idx + 1);
    if (!rectT<T>(kind - K_RECT_XY, a.x, a.y, a.z, a.w, b.x, o, v, kEps<T>(), kInf<T>(), t)) return (T)0;
    if (kind != K_RECT_XZ) return (T)0;  // only XZ rects are real lights (Lib.hs:684)
    ins.count(22);
    const T area = (a.y - a.x) * (a.w - a.z);
    const T distanceSquared = t * t * sqlen(v);
    const T cosine = mabs(v.y / len(v));
    return (T)0 + (T)1 * (distanceSquared / (cosine * area));
  }
  if (kind == K_SPHERE) {
    const Q4<T> s = ldq(sc.spheres + idx);
    if (!sphereT<T, S>(mk<T>(s.x, s.y, s.z), s.w, o, v, kEps<T>(), kInf<T>(), t)) return (T)0;
    ins.sphereHit(); ins.count(22);
    const T cosThetaMax = msqrt(1 - s.w * s.w / sqlen(mk<T>(s.x, s.y, s.z) - o));
    const T solidAngle = 2 * kPi<T>() * (1 - cosThetaMax);
    return (T)0 + (T)1 * (1 / solidAngle);
  }
  return (T)0;
}

// rayColor (Lib.hs:1298-1333) for one camera path, the warp executing it uniformly; returns the path's radiance
template <class T, class S, int INSTR, int MODE>
__device__ __noinline__ V3<T> coopPath(const DScene<T>& sc, V3<T> qo, V3<T> qd, T time, int depth, Rng& rng, Instr<INSTR>& ins) {
  V3<T> beta = mk<T>(1, 1, 1);
  for (;;) {
    if (depth <= 0) return beta * mk<T>(0, 0, 0);
    ins.count(13); ins.count(1);
    CoopQuery<T> q;
    coopBegin<T, MODE>(sc, q, qo, qd, kEps<T>(), kInf<T>());
    coopTraverse<T, S, INSTR, MODE>(sc, q, ins);
    if (!q.hit) { ins.bounce(kNone, 0.0); return beta * mk<T>(sc.bg[0], sc.bg[1], sc.bg[2]); }
    ins.bounce(q.bprim, (double)q.tbest);
    HitRec<T> rec;
    evalHit<T, MODE>(sc, qo, qd, time, q.tbest, q.bprim, 0u, 0, nullptr, 0u, rec);
    const Q4<T> m = ldq(sc.mats + rec.mat);
    const uint32_t mkind = bitsOf(m.x), tex = bitsOf(m.y);
    V3<T> tv = mk<T>(0, 0, 0);
    if (mkind != M_DIELECTRIC && !(mkind == M_LIGHT && rec.front)) tv = textureValue<T, INSTR, MODE>(sc, tex, rec.u, rec.v, rec.p, ins);
    if (mkind == M_LIGHT) { ins.count(20); return beta * tv; }
    if (mkind == M_LAMBERTIAN) {  // Lib.hs:823-836, 367-382, 1329-1333 — same operation order as shadeStep / lpdfAdvance
      ins.count(16);
      const Onb<T> uvw = onbFromW(rec.n);
      const T rd = rnd<T>(rng);
      V3<T> pdfD;
      if (rd < (T)0.5) { ins.count(21); pdfD = htblRandom<T, MODE>(sc, sc.lights_root, rec.p, rng); }
      else pdfD = onbLocal(uvw, randomCosineDirection<T, MODE>(rng));
      const V3<T> dir = unit(pdfD);
      const T cosine = dot(unit(dir), uvw.w);
      const T v2 = (cosine <= 0) ? (T)0 : cosine / kPi<T>();
      const T c2 = dot(rec.n, dir);
      const T spdf = (c2 < 0) ? (T)0 : c2 / kPi<T>();
      beta = beta * tv;
      T v1 = (T)0;
      if ((sc.lights_root >> 28) == K_UNHITTABLE) { ins.count(13); ins.visit(sc.lights_root); }
      else v1 = coopLightPdf<T, S, INSTR, MODE>(sc, rec.p, dir, ins);
      const T pdfVal = (T)0.5 * (v1 + v2);
      beta = scale(spdf / pdfVal, beta);
      qo = rec.p; qd = dir;
    } else if (mkind == M_METAL) {
      ins.count(17);
      const V3<T> rUnit = randomUnitVector<T, MODE>(rng);
      const V3<T> reflected = reflectV(unit(qd), rec.n);
      beta = beta * tv;
      qo = rec.p; qd = reflected + scale(m.z, rUnit);
    } else if (mkind == M_DIELECTRIC) {
      ins.count(18);
      const T eta = rec.front ? (T)1.0 / m.z : m.z;
      const V3<T> ud = unit(qd);
      const T cosTheta = hmin(dot(neg(ud), rec.n), (T)1.0);
      const T sinTheta = msqrt((T)1.0 - cosTheta * cosTheta);
      const T rd = rnd<T>(rng);
      if ((eta * sinTheta > (T)1.0) || rd < schlick<T>(cosTheta, eta)) qd = reflectV(ud, rec.n);
      else qd = refractV(ud, rec.n, eta);
      qo = rec.p;
    } else {  // Isotropic
      ins.count(19);
      const V3<T> randDr = randomInUnitSphere<T>(rng);
      const T c2 = dot(rec.n, randDr);
      const T spdf = (c2 < 0) ? (T)0 : c2 / kPi<T>();
      beta = scale(spdf / (T)1.0, beta * tv);
      qo = rec.p; qd = randDr;
    }
    --depth;
  }
}

// K2 for scan-stream scenes: persistent warps, one camera path per WARP at a time
template <class T, class S, int INSTR, int MODE>
__global__ void __launch_bounds__(kBlock) renderKernelCoop(const __grid_constant__ DScene<T> sc, const __grid_constant__ RenderArgs a) {
  const unsigned lane = threadIdx.x & 31u;
  Instr<INSTR> ins;
  if constexpr (INSTR == 1) ins.init();
  if constexpr (INSTR == 3) ins.init(a.plogPrim, a.plogT, a.plogStride);
  const uint32_t regionW = a.col1 - a.col0;
  uint32_t accPix = kNone; double acc0 = 0, acc1 = 0, acc2 = 0;
  for (;;) {
    unsigned long long idx = 0;
    if (lane == 0) idx = atomicAdd(a.counter, 1ull);
    idx = __shfl_sync(kFullMask, idx, 0);
    if (idx >= a.total_paths) {
      if (lane == 0) atomicCAS(a.counter + kCtrTailBase, 0ull, globalTimerNs());
      break;
    }
    const uint32_t pl = (uint32_t)(idx / a.ns), s = a.s0 + (uint32_t)(idx % a.ns);
    const uint32_t row = a.row0 + (pl / regionW) * a.rowStep, col = a.col0 + pl % regionW;
    const uint32_t pixel = row * a.W + col;
    if (pixel != accPix) {
      if (lane == 0 && accPix != kNone) { atomicAdd(a.out + (size_t)accPix * 3, acc0); atomicAdd(a.out + (size_t)accPix * 3 + 1, acc1); atomicAdd(a.out + (size_t)accPix * 3 + 2, acc2); }
      accPix = pixel; acc0 = acc1 = acc2 = 0;
    }
    Rng rng = streamGen(a.seed, pixel, s);
    ins.count(0);
    ins.beginPath(idx);
    const T ru = rnd<T>(rng), rv = rnd<T>(rng);
    const T u = ((T)col + ru) / (T)a.W, v = ((T)(a.H - 1 - row) + rv) / (T)a.H;
    const V3<T> rd = scale(sc.cam[21], randomInUnitDisk<T>(rng));
    const V3<T> cu = mk<T>(sc.cam[12], sc.cam[13], sc.cam[14]), cv = mk<T>(sc.cam[15], sc.cam[16], sc.cam[17]);
    const V3<T> offset = scale(rd.x, cu) + scale(rd.y, cv);
    const T r = rnd<T>(rng);
    const T time = sc.cam[22] + (sc.cam[23] - sc.cam[22]) * r;
    const V3<T> org = mk<T>(sc.cam[0], sc.cam[1], sc.cam[2]);
    const V3<T> qo = org + offset;
    const V3<T> qd = mk<T>(sc.cam[3], sc.cam[4], sc.cam[5]) + scale(u, mk<T>(sc.cam[6], sc.cam[7], sc.cam[8])) + scale(v, mk<T>(sc.cam[9], sc.cam[10], sc.cam[11])) - org - offset;
    if ((int)a.maxDepth > 0) {  // rayColor with depth <= 0 is black and adds nothing (as renderKernel)
      const V3<T> L = coopPath<T, S, INSTR, MODE>(sc, qo, qd, time, (int)a.maxDepth, rng, ins);
      acc0 += (double)L.x; acc1 += (double)L.y; acc2 += (double)L.z;
    }
  }
  if (lane == 0 && accPix != kNone) { atomicAdd(a.out + (size_t)accPix * 3, acc0); atomicAdd(a.out + (size_t)accPix * 3 + 1, acc1); atomicAdd(a.out + (size_t)accPix * 3 + 2, acc2); }
  if (lane == 0) atomicMax(a.counter + kCtrTailBase + 1, globalTimerNs());
  if constexpr (INSTR == 1) {
    if (lane == 0) for (int i = 0; i < kWorkCounters; ++i) if (ins.c[i]) atomicAdd(a.ctr + i, (unsigned long long)ins.c[i]);
  }
}

// K1 for scan-stream scenes: one WARP per ray
template <class T, class S, int MODE>
__global__ void __launch_bounds__(kBlock) traceKernelCoop(const __grid_constant__ DScene<T> sc, const __grid_constant__ TraceArgs a) {
  const uint32_t lane = threadIdx.x & 31u;
  const uint32_t i = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (i >= a.n) return;
  const double* r = a.rays + (size_t)i * 7;
  const V3<T> o = mk<T>((T)r[0], (T)r[1], (T)r[2]), d = mk<T>((T)r[3], (T)r[4], (T)r[5]);
  const T time = (T)r[6];
  Instr<2> ins; ins.init(a.vlog + (size_t)i * a.logStride, a.logStride);
  for (uint32_t k = lane; k < a.logStride; k += 32u) a.vlog[(size_t)i * a.logStride + k] = kNone;
  __syncwarp();
  CoopQuery<T> q;
  coopBegin<T, MODE>(sc, q, o, d, (T)a.tmin, (T)a.tmax);
  coopTraverse<T, S, 2, MODE>(sc, q, ins);
  if (lane != 0) return;
  a.nvisit[i] = ins.n; a.vhash[i] = ins.h;
  a.hit[i] = q.hit ? 1 : 0;
  double* out = a.rec + (size_t)i * 9;
  if (q.hit) {
    HitRec<T> rec;
    evalHit<T, MODE>(sc, o, d, time, q.tbest, q.bprim, 0u, 0, nullptr, 0u, rec);
    out[0] = rec.t; out[1] = rec.p.x; out[2] = rec.p.y; out[3] = rec.p.z; out[4] = rec.n.x; out[5] = rec.n.y; out[6] = rec.n.z; out[7] = rec.u; out[8] = rec.v;
    a.prim[i] = q.bprim; a.mat[i] = rec.mat; a.front[i] = rec.front ? 1 : 0;
  } else {
    for (int k = 0; k < 9; ++k) out[k] = 0;
    a.prim[i] = 0; a.mat[i] = kNone; a.front[i] = 0;
  }
}

}  // namespace rtb
