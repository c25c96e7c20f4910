// rt_traverse.cuh — `hit root ray t_min t_max` (Lib.hs:970-1109) as a resumable explicit-stack machine.
//
// The traversal state of one query lives in registers (Trav: scalars only, never address-taken) plus a per-thread column of
// shared memory (stack + rarely used frame / medium / light-pdf / accumulator state), and advances by ONE step at a time:
//     stepNode   : a BVHNode — the three independent slab tests (Lib.hs:798-814), then left child next, right child stacked
//     stepSphere : a Sphere or MovingSphere leaf (one code path for both)
//     stepRect   : a Rect or Cuboid leaf
//     stepRare   : Translate / Rotate frames and their exit markers, ConstantMedium and its two boundary sub-queries,
//                  Unhittable
// so that the render kernel can run, per warp iteration, the step class most lanes are waiting for (rt_render_kernel.inl).
// Per lane the sequence of steps is exactly the reference's evaluation order:
//  * BVHNode (971-988): left first, then right with t_max := running closest t; any later accepted hit replaces the earlier
//    one — a DFS with a running closest t in which every accepted primitive overwrites the record.
//  * Translate / Rotate (1029-1052) open a frame: the ray is transformed, a frame-exit marker is stacked under the child;
//    t is frame-invariant so the running closest t carries through.  The record transform is deferred: the chain of open
//    frames is stored with the hit and replayed once by evalHit().
//  * ConstantMedium (1053-1080) runs two sub-queries on its boundary with their own [t_min,t_max] and only needs their t;
//    the outer query state is parked in shared memory meanwhile (one level of nesting, validated on upload).
#pragma once
#include "rt_device.cuh"

namespace rtb {

constexpr uint32_t kEmpty = 0xffffffffu;  // Trav::cur when the query has no step left (kind 15)
constexpr uint32_t kDead = 0xe0000000u;   // Trav::cur of a lane with no work left at all (kind 14)

constexpr int kBlock = 128;               // threads per block of every kernel that owns shared-memory columns

// Shared-memory column of one thread: word e of thread t is col[e * kBlock] (compile-time stride and offsets).
//   words 0..5                      radiance accumulator of the current pixel (3 doubles), word 6 its pixel index
//   words 8 + slot * w              "T slots" (w = sizeof(T)/4 words each; integers use the first word):
//   words extraWords<T>() + [0, stackCap)   traversal stack
//   words extraWords<T>() + stackCap + i * (1 + w)   light-pdf walk stack entry i: ref, weight (render kernel only)
constexpr int XW_ACC = 0, XW_ACCPIX = 6, XW_TSLOTS = 8;
constexpr int X_CHAIN = 0;      // ..3   open frames (transform refs)
constexpr int X_BCHAIN = 4;     // ..7   frames of the best hit
constexpr int X_STMIN = 8, X_STBEST = 9, X_H1T = 10;   // outer query parked during a medium sub-query; first boundary hit
constexpr int X_WORLD = 11;     // ..16  world ray o,d saved while frames are open
constexpr int X_LSPDF = 17, X_LV2 = 18, X_LV1 = 19, X_CURW = 20;  // Lambertian mixture-pdf bookkeeping (Lib.hs:367-382, 823-836)
constexpr int X_CURLIGHT = 21, X_LSP = 22;
constexpr int kTSlots = 23;
// htblPdfValue's walk stack over the lights BVH (ref + weight per entry) sits BEHIND the traversal stack and is sized by the scene's
// light list (RenderArgs::lightCap entries; none when the light list is not a BVH).  Columns are shared memory that the L1 does not
// get: 16 words less per thread are 48 KB more L1 per SM at 6 blocks (cfg3: L1-bound node / leaf fetches, +15 % samples/s).
template <class T> __host__ __device__ constexpr int extraWords() { return XW_TSLOTS + kTSlots * (int)(sizeof(T) / 4); }
// words of one render-kernel column: fixed slots + traversal stack + light-pdf walk stack
template <class T> __host__ __device__ constexpr uint32_t columnWords(uint32_t stackCap, uint32_t lightCap) { return (uint32_t)extraWords<T>() + stackCap + lightCap * (1u + (uint32_t)(sizeof(T) / 4)); }

// Trav::flags bit fields
constexpr uint32_t F_INSUB = 1u, F_SQHIT = 2u, F_QHIT = 4u, F_NANT = 8u;  // F_NANT: a NaN t was accepted (rect with d_k = 0, o_k = k)
constexpr int F_FD_SHIFT = 4, F_BDEPTH_SHIFT = 8, F_BAUX_SHIFT = 12;
constexpr int F_MODE_SHIFT = 16;  // 2 bits per axis: division mode of the slab test (rt_device.cuh slabMode), literal fp32 build only
constexpr uint32_t F_MODE_MASK = 63u << F_MODE_SHIFT, F_MODE_GENERIC = (2u | (2u << 2) | (2u << 4)) << F_MODE_SHIFT;

template <class T> struct Trav {
  V3<T> o, d;       // ray in the current frame
  V3<T> inv;        // FAST: 1/d; literal fp32: the refined reciprocal (or +-inf) of the axis' division mode
  T tmin, tbest;    // the query's t_min and its running closest t (t_max)
  uint32_t cur;     // next hittable ref / marker to process, kEmpty when done
  uint32_t bprim;   // primitive of the best hit (valid when qhit, outside sub-queries)
  uint32_t flags;   // F_INSUB | F_SQHIT | F_QHIT | F_NANT | fd << 4 | bdepth << 8 | baux << 12 | axis modes << 16
  int sp;           // stack entries in use
  uint32_t* col;    // this thread's shared-memory column
  static constexpr uint32_t stride = kBlock;
  static constexpr int kStackBase = XW_TSLOTS + kTSlots * (int)(sizeof(T) / 4);

  RTB_D void push(uint32_t v) { col[(kStackBase + sp) * stride] = v; ++sp; }
  RTB_D void next() { if (sp > 0) { --sp; cur = col[(kStackBase + sp) * stride]; } else cur = kEmpty; }
  RTB_D uint32_t& xw(int word) { return col[word * stride]; }                                            // raw extra word
  RTB_D uint32_t& xu(int slot) { return col[(XW_TSLOTS + slot * (int)(sizeof(T) / 4)) * stride]; }       // integer T slot
  RTB_D bool qhit() const { return (flags & F_QHIT) != 0; }
  RTB_D int fd() const { return (int)((flags >> F_FD_SHIFT) & 7u); }
  RTB_D int bdepth() const { return (int)((flags >> F_BDEPTH_SHIFT) & 7u); }
  RTB_D uint32_t baux() const { return (flags >> F_BAUX_SHIFT) & 7u; }
  RTB_D void setFd(int v) { flags = (flags & ~(7u << F_FD_SHIFT)) | ((uint32_t)v << F_FD_SHIFT); }
};

RTB_D void stT(uint32_t* p, uint32_t stride, float v) { p[0] = __float_as_uint(v); }
RTB_D void stT(uint32_t* p, uint32_t stride, double v) { unsigned long long b = (unsigned long long)__double_as_longlong(v); p[0] = (uint32_t)b; p[stride] = (uint32_t)(b >> 32); }
RTB_D void ldT(const uint32_t* p, uint32_t stride, float& v) { v = __uint_as_float(p[0]); }
RTB_D void ldT(const uint32_t* p, uint32_t stride, double& v) { v = __longlong_as_double((long long)((unsigned long long)p[0] | ((unsigned long long)p[stride] << 32))); }
template <class T> RTB_D void xst(Trav<T>& tr, int slot, T v) { stT(&tr.xu(slot), tr.stride, v); }
template <class T> RTB_D T xld(Trav<T>& tr, int slot) { T v; ldT(&tr.xu(slot), tr.stride, v); return v; }
// entry i of the light-pdf walk stack; lbase = extraWords<T>() + stackCap (words)
template <class T> RTB_D uint32_t* lightEntry(Trav<T>& tr, uint32_t lbase, uint32_t i) { return tr.col + (size_t)(lbase + i * (1u + (uint32_t)(sizeof(T) / 4))) * tr.stride; }

RTB_D void axisMode(float d, bool oSafe, float& r, uint32_t& mode) {
  const float ad = fabsf(d);
  if (oSafe && ad >= kDivLo && ad <= kDivHi) { r = rcpRefined(d); mode = 0u; }
  else if (d == 0.0f) { r = copysignf(__int_as_float(0x7f800000), d); mode = 1u; }
  else { r = 0.0f; mode = 2u; }
}
// per-ray (per frame) setup of the slab test's divisions
template <class T, int MODE> RTB_D void setInv(const DScene<T>& sc, Trav<T>& tr) {
  if constexpr (MODE == MATH_FAST) {
    tr.inv = mk<T>((T)1 / tr.d.x, (T)1 / tr.d.y, (T)1 / tr.d.z);
  } else if constexpr (sizeof(T) == 4) {
    const bool oSafe = !sc.bigCoords && !(tr.flags & F_NANT) && fabsf(tr.o.x) <= kDivHi && fabsf(tr.o.y) <= kDivHi && fabsf(tr.o.z) <= kDivHi;
    uint32_t mx, my, mz;
    axisMode(tr.d.x, oSafe, tr.inv.x, mx); axisMode(tr.d.y, oSafe, tr.inv.y, my); axisMode(tr.d.z, oSafe, tr.inv.z, mz);
    tr.flags = (tr.flags & ~F_MODE_MASK) | ((mx | (my << 2) | (mz << 4)) << F_MODE_SHIFT);
  }
}

// start `hit root (Ray o d _) t_min t_max`
template <class T, int MODE>
RTB_D void travBegin(const DScene<T>& sc, Trav<T>& tr, uint32_t root, const V3<T>& o, const V3<T>& d, T t_min, T t_max) {
  tr.o = o; tr.d = d; tr.flags = (t_max == t_max) ? 0u : F_NANT;
  setInv<T, MODE>(sc, tr);
  tr.tmin = t_min; tr.tbest = t_max; tr.cur = root; tr.bprim = 0; tr.sp = 0;
}

// an accepted primitive hit replaces the record (Lib.hs:979-987)
template <class T> RTB_D void travAccept(Trav<T>& tr, T t, uint32_t ref, uint32_t aux) {
  tr.tbest = t;
  if (!(t == t)) tr.flags |= F_NANT | F_MODE_GENERIC;  // a NaN t_max: only the literal slab test is right from here on
  if (!(tr.flags & F_INSUB)) {
    const uint32_t fd = (tr.flags >> F_FD_SHIFT) & 7u;
    tr.bprim = ref;
    tr.flags = (tr.flags & 0xffff00ffu) | F_QHIT | (fd << F_BDEPTH_SHIFT) | (aux << F_BAUX_SHIFT);
    for (uint32_t i = 0; i < fd; ++i) tr.xu(X_BCHAIN + (int)i) = tr.xu(X_CHAIN + (int)i);
  } else {
    tr.flags |= F_QHIT;
  }
}

template <class T, int INSTR, int MODE>
RTB_D void stepNode(const DScene<T>& sc, Trav<T>& tr, Instr<INSTR>& ins) {
  const uint32_t idx = tr.cur & kIdxMask;
  ins.visit(tr.cur); ins.boxTest();
  // byte offset of the node record = index * 2 * sizeof(Q4<T>): the shift drops the kind nibble and the (clear) pair bit with it
  const Q4<T>* np = reinterpret_cast<const Q4<T>*>(reinterpret_cast<const char*>(sc.nodes) + ((size_t)(tr.cur << 5) * (sizeof(T) / 4)));
  const Q4<T> lo = ldq(np), hi = ldq(np + 1);
  // pass: go left, right child stacked -- or, when both children are leaves of one class, one pair step and nothing stacked
  auto descend = [&]() {
    const uint32_t lc = bitsOf(lo.w);
    if (lc & kPairBit) tr.cur = (lc & 0xf0000000u) | kPairBit | idx;
    else { tr.push(bitsOf(hi.w)); tr.cur = lc; }
  };
  bool pass;
  if constexpr (MODE != MATH_FAST && sizeof(T) == 4) {
    bool ok = (tr.flags & F_MODE_GENERIC) == 0u;  // no axis in the generic division mode
    if constexpr (kSlabFilter) {
      // interval filter: the approximate quotients decide unless a comparison margin is inside their error bound (rt_device.cuh)
      bool sure = ok;  // no axis in the generic division mode; zero divisors are handled by the filter itself
      // t_max = +inf (no hit yet) enters as FLT_MAX: same decisions (finite quotients are < 2^81, infinite ones compare alike), and a
      // zero-divisor axis that misses the box (both quotients +inf) gives m = -inf, a certain "no", instead of inf - inf
      const float tF = fminf(tr.tbest, 3.4028234663852886e38f);
      const bool fx = slabFiltered(lo.x, hi.x, tr.o.x, tr.inv.x, tr.tmin, tF, sure);
      const bool fy = slabFiltered(lo.y, hi.y, tr.o.y, tr.inv.y, tr.tmin, tF, sure);
      const bool fz = slabFiltered(lo.z, hi.z, tr.o.z, tr.inv.z, tr.tmin, tF, sure);
#ifdef RTB_DEBUG_UNSURE
      if (!sure) ins.count(23);
      if ((tr.flags & F_MODE_MASK) != 0u) ins.count(22);
#endif
      if (sure) {
        if (fx & fy & fz) descend(); else tr.next();
        return;
      }
      if (boxRefined(lo.x, lo.y, lo.z, hi.x, hi.y, hi.z, tr.o.x, tr.o.y, tr.o.z, tr.d.x, tr.d.y, tr.d.z, tr.inv.x, tr.inv.y, tr.inv.z,
                     (tr.flags & F_MODE_GENERIC) != 0u, tr.tmin, tr.tbest)) descend();
      else tr.next();
      return;
    }
    const bool bx = slabHoisted(lo.x, hi.x, tr.o.x, tr.d.x, tr.inv.x, tr.tmin, tr.tbest, ok);
    const bool by = slabHoisted(lo.y, hi.y, tr.o.y, tr.d.y, tr.inv.y, tr.tmin, tr.tbest, ok);
    const bool bz = slabHoisted(lo.z, hi.z, tr.o.z, tr.d.z, tr.inv.z, tr.tmin, tr.tbest, ok);
    pass = bx && by && bz;
    if (!ok) {
      // a quotient left the validated range of the hoisted reciprocal (e.g. an exact 0), or generic mode: the out-of-line test
      // decides.  Its own copy of the two outcomes keeps `pass` a predicate (not a register merged after the call) in the hot path.
      if (boxFallback(lo.x, lo.y, lo.z, hi.x, hi.y, hi.z, tr.o.x, tr.o.y, tr.o.z, tr.d.x, tr.d.y, tr.d.z, tr.inv.x, tr.inv.y, tr.inv.z,
                      (tr.flags & F_MODE_GENERIC) != 0u, tr.tmin, tr.tbest)) descend();
      else tr.next();
      return;
    }
  } else {
    const bool bx = slab<T, MODE == MATH_FAST>(lo.x, hi.x, tr.o.x, tr.d.x, tr.inv.x, tr.tmin, tr.tbest);
    const bool by = slab<T, MODE == MATH_FAST>(lo.y, hi.y, tr.o.y, tr.d.y, tr.inv.y, tr.tmin, tr.tbest);
    const bool bz = slab<T, MODE == MATH_FAST>(lo.z, hi.z, tr.o.z, tr.d.z, tr.inv.z, tr.tmin, tr.tbest);
    pass = bx && by && bz;
  }
  if (pass) descend(); else tr.next();
}

template <class T, class S, int INSTR>
RTB_D void stepSphere(const DScene<T>& sc, Trav<T>& tr, T time, Instr<INSTR>& ins) {
  uint32_t ref = tr.cur, ref2 = kNone;
  if (ref & kPairBit) {  // the two leaves of a bottom node, left then right (Lib.hs:971-988)
    const uint32_t n = ref & kIdxMask;
    ref = bitsOf(ldq(sc.nodes + 2 * (size_t)n).w) & ~kPairBit; ref2 = bitsOf(ldq(sc.nodes + 2 * (size_t)n + 1).w);
  }
#pragma unroll 1
  for (;;) {
    const uint32_t idx = ref & kIdxMask;
    ins.visit(ref);
    V3<T> ctr; T rad;
    if ((ref >> 28) == K_SPHERE) {
      const Q4<T> q = ldq(sc.spheres + idx);
      ctr = mk<T>(q.x, q.y, q.z); rad = q.w;
    } else {  // MovingSphere (Lib.hs:1106-1108)
      const Q4<T> a = ldq(sc.moving + 3 * (size_t)idx), b = ldq(sc.moving + 3 * (size_t)idx + 1), c = ldq(sc.moving + 3 * (size_t)idx + 2);
      const V3<T> c0 = mk<T>(a.x, a.y, a.z), c1 = mk<T>(b.x, b.y, b.z);
      ctr = c0 + scale(movingFrac<T>(time, b.w, c.x, c.z), c1 - c0); rad = a.w;
    }
    T t;
    if (sphereT<T, S>(ctr, rad, tr.o, tr.d, tr.tmin, tr.tbest, t)) { ins.sphereHit(); travAccept<T>(tr, t, ref, 0); }
    if (ref2 == kNone) break;
    ref = ref2; ref2 = kNone;
  }
  tr.next();
}

// Rect and Cuboid leaves
template <class T, int INSTR, int MODE = MATH_LITERAL>
RTB_D void stepRect(const DScene<T>& sc, Trav<T>& tr, Instr<INSTR>& ins) {
  // plane quotients through the ray's refined reciprocal in the fp32 literal / strict builds (rt_device.cuh rectTHoisted)
  auto rectHit = [&](uint32_t plane, T i0, T i1, T j0, T j1, T k, T& t) -> bool {
    if constexpr (MODE != MATH_FAST && sizeof(T) == 4 && kPlaneHoist && kRectHoist) return rectTHoisted(plane, i0, i1, j0, j1, k, tr.o, tr.d, tr.inv, tr.flags >> F_MODE_SHIFT, tr.tmin, tr.tbest, t);
    else return rectT<T>(plane, i0, i1, j0, j1, k, tr.o, tr.d, tr.tmin, tr.tbest, t);
  };
  uint32_t ref = tr.cur, ref2 = kNone;
  if (ref & kPairBit) {  // the two leaves of a bottom node, left then right
    const uint32_t n = ref & kIdxMask;
    ref = bitsOf(ldq(sc.nodes + 2 * (size_t)n).w) & ~kPairBit; ref2 = bitsOf(ldq(sc.nodes + 2 * (size_t)n + 1).w);
  }
#pragma unroll 1
  for (;;) {
    const uint32_t kind = ref >> 28, idx = ref & kIdxMask;
    ins.visit(ref);
    if (kind != K_CUBOID) {
      const Q4<T> a = ldq(sc.rects + 2 * (size_t)idx), b = ldq(sc.rects + 2 * (size_t)idx + 1);
      T t;
      if (rectHit(kind - K_RECT_XY, a.x, a.y, a.z, a.w, b.x, t)) travAccept<T>(tr, t, ref, 0);
    } else {  // Lib.hs:989-1004: foldr over the six rects => last first; strict `<` replaces
      const Q4<T> mn = ldq(sc.cuboids + 2 * (size_t)idx), mx = ldq(sc.cuboids + 2 * (size_t)idx + 1);
      bool acc; T tNew; uint32_t aux;
      if constexpr (MODE != MATH_FAST && sizeof(T) == 4 && kPlaneHoist) {
        const int h = cuboidFacesHoisted(mn, mx, tr.o, tr.d, tr.inv, tr.flags >> F_MODE_SHIFT, tr.tmin, tr.tbest, tNew, aux);
        acc = h < 0 ? cuboidFacesRolled(mn, mx, tr.o, tr.d, tr.tmin, tr.tbest, tNew, aux) : h > 0;
      } else {
        acc = cuboidFaces<T>(mn, mx, tr.o, tr.d, tr.tmin, tr.tbest, tNew, aux);
      }
      if (acc) travAccept<T>(tr, tNew, ref, aux);
    }
    if (ref2 == kNone) break;
    ref = ref2; ref2 = kNone;
  }
  tr.next();
}

// restore the current-frame ray after a frame exit: the saved world ray, then replay the chain of still-open frames
template <class T, int MODE> RTB_D void travReframe(const DScene<T>& sc, Trav<T>& tr) {
  const int fd = tr.fd();
  tr.o = mk<T>(xld<T>(tr, X_WORLD + 0), xld<T>(tr, X_WORLD + 1), xld<T>(tr, X_WORLD + 2));
  tr.d = mk<T>(xld<T>(tr, X_WORLD + 3), xld<T>(tr, X_WORLD + 4), xld<T>(tr, X_WORLD + 5));
  for (int i = 0; i < fd; ++i) applyXform<T>(sc, tr.xu(X_CHAIN + i), tr.o, tr.d);
  setInv<T, MODE>(sc, tr);
}

// the rare kinds: frames, markers, media, Unhittable
template <class T, int INSTR, int MODE>
RTB_D void stepRare(const DScene<T>& sc, Trav<T>& tr, Rng& rng, Instr<INSTR>& ins) {
  const uint32_t ref = tr.cur, kind = ref >> 28, idx = ref & kIdxMask;
  if (kind == K_TRANSLATE || kind == K_ROTATE) {
    ins.visit(ref);
    const uint32_t child = (kind == K_TRANSLATE) ? bitsOf(ldq(sc.translates + idx).w) : bitsOf(ldq(sc.rotates + idx).w);
    const int fd = tr.fd();
    if (fd == 0) {  // park the world ray: restoring it must be exact, not (o - off) + off
      xst<T>(tr, X_WORLD + 0, tr.o.x); xst<T>(tr, X_WORLD + 1, tr.o.y); xst<T>(tr, X_WORLD + 2, tr.o.z);
      xst<T>(tr, X_WORLD + 3, tr.d.x); xst<T>(tr, X_WORLD + 4, tr.d.y); xst<T>(tr, X_WORLD + 5, tr.d.z);
    }
    tr.xu(X_CHAIN + fd) = ref;
    tr.setFd(fd + 1);
    applyXform<T>(sc, ref, tr.o, tr.d);
    setInv<T, MODE>(sc, tr);
    tr.push(K_MARK_FRAME << 28);
    tr.cur = child;
  } else if (kind == K_MARK_FRAME) {
    tr.setFd(tr.fd() - 1);
    travReframe<T, MODE>(sc, tr);
    tr.next();
  } else if (kind == K_MEDIUM) {  // Lib.hs:1054: hit boundary ray (-inf) inf
    ins.visit(ref);
    xst<T>(tr, X_STMIN, tr.tmin); xst<T>(tr, X_STBEST, tr.tbest);
    tr.flags = (tr.flags & ~(F_SQHIT | F_QHIT)) | F_INSUB | ((tr.flags & F_QHIT) ? F_SQHIT : 0u);
    tr.tmin = -kInf<T>(); tr.tbest = kInf<T>();
    tr.push((K_MARK_MED1 << 28) | idx);
    tr.cur = bitsOf(ldq(sc.media + idx).z);
  } else if (kind == K_MARK_MED1) {
    if (!tr.qhit()) {
      tr.tmin = xld<T>(tr, X_STMIN); tr.tbest = xld<T>(tr, X_STBEST);
      tr.flags = (tr.flags & ~(F_INSUB | F_QHIT | F_SQHIT)) | ((tr.flags & F_SQHIT) ? F_QHIT : 0u);
      tr.next();
    } else {  // Lib.hs:1057: hit boundary ray (h1t + epsilon) inf
      xst<T>(tr, X_H1T, tr.tbest);
      tr.tmin = tr.tbest + kEps<T>(); tr.tbest = kInf<T>(); tr.flags &= ~F_QHIT;
      tr.push((K_MARK_MED2 << 28) | idx);
      tr.cur = bitsOf(ldq(sc.media + idx).z);
    }
  } else if (kind == K_MARK_MED2) {
    const bool h2 = tr.qhit(); const T h2t = tr.tbest;
    tr.tmin = xld<T>(tr, X_STMIN); tr.tbest = xld<T>(tr, X_STBEST);
    tr.flags = (tr.flags & ~(F_INSUB | F_QHIT | F_SQHIT)) | ((tr.flags & F_SQHIT) ? F_QHIT : 0u);
    if (h2) {  // Lib.hs:1060-1079
      const T h1t = xld<T>(tr, X_H1T);
      const T rec1t_ = hmax(tr.tmin, h1t);
      const T rec2t = hmin(tr.tbest, h2t);
      if (!(rec1t_ >= rec2t)) {
        const T rec1t = (rec1t_ < 0) ? (T)0 : rec1t_;
        const T rayLength = len(tr.d);
        const T distInside = (rec2t - rec1t) * rayLength;
        const T xi = rnd<T>(rng);
        const T hitDist = ldq(sc.media + idx).x * mlog<MODE>(xi);
        if (!(hitDist > distInside)) travAccept<T>(tr, rec1t + (hitDist / rayLength), (K_MEDIUM << 28) | idx, 0);
      }
    }
    tr.next();
  } else {  // Unhittable (Lib.hs:1109)
    ins.visit(ref);
    tr.next();
  }
}

// step class of a lane's pending work: 0 node, 1 sphere-like leaf, 2 rect / cuboid leaf, 3 rare, 4 query finished (kEmpty),
// 5 dead lane (kDead); one nibble per kind
RTB_D int stepClass(uint32_t cur) {
  return (int)((0x4533333302222113ULL >> ((cur >> 28) * 4u)) & 0xfULL);
}

// run a whole query to completion (K1 and the texture-free probes): same steps, no scheduling
template <class T, class S, int INSTR, int MODE>
RTB_D void travRun(const DScene<T>& sc, Trav<T>& tr, T time, Rng& rng, Instr<INSTR>& ins) {
  while (tr.cur != kEmpty) {
    const int c = stepClass(tr.cur);
    if (c == 0) stepNode<T, INSTR, MODE>(sc, tr, ins);
    else if (c == 1) stepSphere<T, S, INSTR>(sc, tr, time, ins);
    else if (c == 2) stepRect<T, INSTR, MODE>(sc, tr, ins);
    else stepRare<T, INSTR, MODE>(sc, tr, rng, ins);
  }
}

}  // namespace rtb
