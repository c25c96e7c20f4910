// oracle/oracle_capi.cpp — C entry points over rt_oracle.hpp / rt_scenes.hpp for ctypes (tests/, smoke(), bench.py
// cpu_baseline and --impl reference).  TEST INFRASTRUCTURE ONLY; the product never links this.
#include "rt_scenes.hpp"

#include <cstdio>

using namespace orc;

namespace {
struct Handle { Scene scene; SceneArgs args; std::string name; };
thread_local std::string g_err;

Camera cameraFrom24(const double* c) {
  Camera k;
  k.origin = V3d(c[0], c[1], c[2]); k.llc = V3d(c[3], c[4], c[5]); k.horiz = V3d(c[6], c[7], c[8]); k.vert = V3d(c[9], c[10], c[11]);
  k.u = V3d(c[12], c[13], c[14]); k.v = V3d(c[15], c[16], c[17]); k.w = V3d(c[18], c[19], c[20]);
  k.lensRadius = c[21]; k.t0 = c[22]; k.t1 = c[23];
  return k;
}
void cameraTo24(const Camera& k, double* c) {
  const V3d* vs[7] = {&k.origin, &k.llc, &k.horiz, &k.vert, &k.u, &k.v, &k.w};
  for (int i = 0; i < 7; ++i) { c[i * 3] = vs[i]->x; c[i * 3 + 1] = vs[i]->y; c[i * 3 + 2] = vs[i]->z; }
  c[21] = k.lensRadius; c[22] = k.t0; c[23] = k.t1;
}
void statsOut(const Trace& t, uint64_t* s) {
  if (!s) return;
  s[0] = t.paths; s[1] = t.bounces; s[2] = t.boxTests; s[3] = t.prim[Hittable::Sphere]; s[4] = t.prim[Hittable::MovingSphere];
  s[5] = t.prim[Hittable::RectXY] + t.prim[Hittable::RectXZ] + t.prim[Hittable::RectYZ]; s[6] = t.prim[Hittable::Cuboid];
  s[7] = t.prim[Hittable::Translate]; s[8] = t.prim[Hittable::Rotate]; s[9] = t.prim[Hittable::ConstantMedium];
  s[10] = t.sphereHits; s[11] = t.noiseCalls; s[12] = t.texels; s[13] = t.queries; s[14] = t.nVisit; s[15] = t.prim[Hittable::Unhittable];
  for (int i = 0; i < 8; ++i) s[16 + i] = t.shade[i];
}

template <class T, class S>
void traceImpl(const Scene& sc, int n, const double* rays, double tmin, double tmax, uint64_t seed, int logStride, int32_t* hitFlag,
               double* rec, uint32_t* prim, uint32_t* mat, uint8_t* front, uint64_t* nvisit, uint64_t* vhash, uint32_t* vlog, bool strict) {
  std::atomic<int> next{0};
  auto work = [&]() {
  for (;;) {
    const int i = next.fetch_add(1);
    if (i >= n) break;
    const double* r = rays + (size_t)i * 7;
    Ray<T> ray{Vec3<T>((T)r[0], (T)r[1], (T)r[2]), Vec3<T>((T)r[3], (T)r[4], (T)r[5]), (T)r[6]};
    SMGen g = streamGen(seed, (uint32_t)i, 0);
    Trace tr; tr.logVisits = logStride > 0;
    Tracer<T, S> t(sc, g, &tr, strict);
    Hit<T> h{};
    bool ok = t.hit(sc.world.get(), ray, (T)tmin, (T)tmax, h);
    hitFlag[i] = ok ? 1 : 0;
    double* o = rec + (size_t)i * 9;
    if (ok) {
      o[0] = h.t; o[1] = h.p.x; o[2] = h.p.y; o[3] = h.p.z; o[4] = h.normal.x; o[5] = h.normal.y; o[6] = h.normal.z; o[7] = h.u; o[8] = h.v;
      prim[i] = h.prim; mat[i] = (uint32_t)h.mat->id; front[i] = h.frontFace ? 1 : 0;
    } else {
      for (int k = 0; k < 9; ++k) o[k] = 0;
      prim[i] = 0; mat[i] = 0xffffffffu; front[i] = 0;
    }
    nvisit[i] = tr.nVisit; vhash[i] = tr.visitHash;
    if (logStride > 0) {
      for (int k = 0; k < logStride; ++k) vlog[(size_t)i * logStride + k] = k < (int)tr.visits.size() ? tr.visits[k] : 0xffffffffu;
    }
  }
  };
  const int nThreads = (int)std::max(1u, std::min(std::thread::hardware_concurrency(), (unsigned)((n + 63) / 64)));
  std::vector<std::thread> th;
  for (int k = 1; k < nThreads; ++k) th.emplace_back(work);
  work();
  for (auto& t : th) t.join();
}
}  // namespace

extern "C" {

const char* orc_last_error() { return g_err.c_str(); }

void* orc_scene_create(const char* name, int64_t seed, double t0, double t1, const uint8_t* earth, int ew, int eh, int lightsMode, int synthN) {
  try {
    auto* h = new Handle();
    h->name = name; h->args.t0 = t0; h->args.t1 = t1; h->args.lightsMode = lightsMode; h->args.synthN = synthN;
    if (earth && ew > 0 && eh > 0) {
      auto im = std::make_shared<ImageData>(); im->w = ew; im->h = eh; im->rgb.assign(earth, earth + (size_t)ew * eh * 3);
      h->args.earth = im; h->args.earthW = ew; h->args.earthH = eh;
    }
    h->scene = makeSceneByName(h->name, h->args, seed);
    return h;
  } catch (const std::exception& e) { g_err = e.what(); return nullptr; }
}
void orc_scene_destroy(void* p) { delete (Handle*)p; }

// counts[0..10] = objects per kind slot (rects all under slot 3), [11]=materials [12]=textures [13]=perlin tables [14]=images
void orc_scene_counts(void* p, uint32_t* counts) {
  Scene& s = ((Handle*)p)->scene;
  for (int k = 0; k < Hittable::NKinds; ++k) counts[k] = (uint32_t)s.byKind[k].size();
  counts[11] = (uint32_t)s.mats.size(); counts[12] = (uint32_t)s.texs.size(); counts[13] = (uint32_t)s.perlins.size(); counts[14] = (uint32_t)s.images.size();
}
void orc_scene_roots(void* p, uint32_t* roots, double* bg, uint64_t* g1) {
  Scene& s = ((Handle*)p)->scene;
  roots[0] = refOf(s.world.get()); roots[1] = refOf(s.lights.get());
  bg[0] = s.background[0]; bg[1] = s.background[1]; bg[2] = s.background[2];
  g1[0] = s.g1.seed; g1[1] = s.g1.gamma;
}
void orc_camera(const char* name, int W, int H, int synthN, double* out24) { cameraTo24(cameraByName(name, W, H, synthN), out24); }

// Canonical per-kind dumps (layout mirrors include/rt_abi.h's arrays so tests can compare with the product flattener):
//  what = kind slot 1..10:  d = per-object doubles, u = per-object uint32s
//   1 Sphere:   d[4] = cx cy cz r                 u[1] = mat
//   2 Moving:   d[9] = c0 c1 t0 dur r             u[1] = mat
//   3 Rect:     d[5] = a0 a1 b0 b1 k              u[2] = plane(0 XY,1 XZ,2 YZ) mat
//   6 Cuboid:   d[6] = min max                    u[1] = mat
//   7 BVHNode:  d[6] = box min max                u[3] = left right size
//   8 Translate d[3] = offset                     u[1] = child
//   9 Rotate    d[2] = sin cos                    u[2] = axis child
//  10 Medium    d[1] = negInvDensity              u[2] = mat boundary
//  11 Material  d[1] = param                      u[2] = kind tex(0xffffffff if none)
//  12 Texture   d[4] = r g b scale                u[3] = kind a b  (checker: odd even; perlin: table; image: image id or 0xffffffff, b unused)
//  13 Perlin    d[768] = ranvec                   u[768] = permX permY permZ
int orc_export(void* p, int what, double* d, uint32_t* u) {
  Scene& s = ((Handle*)p)->scene;
  auto none = 0xffffffffu;
  if (what >= 1 && what <= 10) {
    size_t i = 0;
    for (const Hittable* h : s.byKind[what]) {
      switch (what) {
        case 1: d[i * 4] = h->c0.x; d[i * 4 + 1] = h->c0.y; d[i * 4 + 2] = h->c0.z; d[i * 4 + 3] = h->r; u[i] = h->mat->id; break;
        case 2: { double* q = d + i * 9; q[0] = h->c0.x; q[1] = h->c0.y; q[2] = h->c0.z; q[3] = h->c1.x; q[4] = h->c1.y; q[5] = h->c1.z; q[6] = h->t0; q[7] = h->dur; q[8] = h->r; u[i] = h->mat->id; break; }
        case 3: { double* q = d + i * 5; q[0] = h->a0; q[1] = h->a1; q[2] = h->b0; q[3] = h->b1; q[4] = h->k; u[i * 2] = h->kind - Hittable::RectXY; u[i * 2 + 1] = h->mat->id; break; }
        case 6: { double* q = d + i * 6; q[0] = h->bmin.x; q[1] = h->bmin.y; q[2] = h->bmin.z; q[3] = h->bmax.x; q[4] = h->bmax.y; q[5] = h->bmax.z; u[i] = h->mat->id; break; }
        case 7: { double* q = d + i * 6; q[0] = h->box.mn.x; q[1] = h->box.mn.y; q[2] = h->box.mn.z; q[3] = h->box.mx.x; q[4] = h->box.mx.y; q[5] = h->box.mx.z;
                  u[i * 3] = refOf(h->left.get()); u[i * 3 + 1] = refOf(h->right.get()); u[i * 3 + 2] = (uint32_t)h->size; break; }
        case 8: d[i * 3] = h->off.x; d[i * 3 + 1] = h->off.y; d[i * 3 + 2] = h->off.z; u[i] = refOf(h->left.get()); break;
        case 9: d[i * 2] = h->sinT; d[i * 2 + 1] = h->cosT; u[i * 2] = (uint32_t)h->axis; u[i * 2 + 1] = refOf(h->left.get()); break;
        case 10: d[i] = h->nInvD; u[i * 2] = h->mat->id; u[i * 2 + 1] = refOf(h->left.get()); break;
        default: return -1;
      }
      ++i;
    }
    return (int)i;
  }
  if (what == 11) { size_t i = 0; for (const Material* m : s.mats) { d[i] = m->param; u[i * 2] = m->kind; u[i * 2 + 1] = m->tex ? (uint32_t)m->tex->id : none; ++i; } return (int)i; }
  if (what == 12) {
    size_t i = 0;
    for (const Texture* t : s.texs) {
      d[i * 4] = t->rgb[0]; d[i * 4 + 1] = t->rgb[1]; d[i * 4 + 2] = t->rgb[2]; d[i * 4 + 3] = t->scale;
      u[i * 3] = t->kind; u[i * 3 + 1] = none; u[i * 3 + 2] = none;
      if (t->kind == Texture::Checker) { u[i * 3 + 1] = t->odd->id; u[i * 3 + 2] = t->even->id; }
      if (t->kind == Texture::Perlin) u[i * 3 + 1] = t->perlin->id;
      if (t->kind == Texture::Image && t->image) u[i * 3 + 1] = t->image->id;
      ++i;
    }
    return (int)i;
  }
  if (what == 13) {
    size_t i = 0;
    for (const PerlinTable* t : s.perlins) {
      for (int k = 0; k < 256; ++k) for (int c = 0; c < 3; ++c) d[i * 768 + k * 3 + c] = t->ranvec[k][c];
      for (int a = 0; a < 3; ++a) for (int k = 0; k < 256; ++k) u[i * 768 + a * 256 + k] = (uint32_t)t->perm[a][k];
      ++i;
    }
    return (int)i;
  }
  return -1;
}

// K1 oracle: hit(world, ray, tmin, tmax) per ray; RNG for media = streamGen(seed, rayIndex, 0).
// prec: 0 = <float,float>, 1 = <float,double> (mixed), 2 = <double,double> (the reference).
// flags: bit 0 = strict math (rt_strict_math.h), bit 1 = beta-order throughput (render only)
int orc_trace(void* p, int prec, int flags, int n, const double* rays7, double tmin, double tmax, uint64_t seed, int logStride, int32_t* hitFlag,
              double* rec9, uint32_t* prim, uint32_t* mat, uint8_t* front, uint64_t* nvisit, uint64_t* vhash, uint32_t* vlog) {
  try {
    Scene& s = ((Handle*)p)->scene;
    const bool strict = (flags & 1) != 0;
    if (prec == 0) traceImpl<float, float>(s, n, rays7, tmin, tmax, seed, logStride, hitFlag, rec9, prim, mat, front, nvisit, vhash, vlog, strict);
    else if (prec == 1) traceImpl<float, double>(s, n, rays7, tmin, tmax, seed, logStride, hitFlag, rec9, prim, mat, front, nvisit, vhash, vlog, strict);
    else traceImpl<double, double>(s, n, rays7, tmin, tmax, seed, logStride, hitFlag, rec9, prim, mat, front, nvisit, vhash, vlog, strict);
    return 0;
  } catch (const std::exception& e) { g_err = e.what(); return -1; }
}

// mode 0: reference RNG plumbing (runRender, per-column SplitMix; `seed` stands in for Main's entropy);
// mode 1: per-(pixel,sample) streams (the product's RNG), region/sample-range restricted.
// out: double[H*W*3] radiance SUMS, row 0 = top.  stats24 may be null.
// flags / path log (mode 1 only): see orc_trace; pathPrim / pathT: [paths][pathStride] or null
int orc_render(void* p, const double* cam24, int prec, int W, int H, int s0, int ns, int maxDepth, int mode, uint64_t seed, int row0, int row1,
               int col0, int col1, int nThreads, double* out, uint64_t* stats24, int flags, uint32_t* pathPrim, double* pathT, uint32_t pathStride) {
  try {
    Scene& s = ((Handle*)p)->scene;
    Camera cam = cameraFrom24(cam24);
    bool st = stats24 != nullptr;
    RenderOut r;
    if (mode == 0) {
      if (prec == 0) r = renderReference<float, float>(s, cam, W, H, ns, maxDepth, seed, nThreads, st);
      else if (prec == 1) r = renderReference<float, double>(s, cam, W, H, ns, maxDepth, seed, nThreads, st);
      else r = renderReference<double, double>(s, cam, W, H, ns, maxDepth, seed, nThreads, st);
    } else {
      const bool strict = (flags & 1) != 0, beta = (flags & 2) != 0;
      if (prec == 0) r = renderStreams<float, float>(s, cam, W, H, s0, ns, maxDepth, seed, row0, row1, col0, col1, nThreads, st, strict, beta, pathPrim, pathT, pathStride);
      else if (prec == 1) r = renderStreams<float, double>(s, cam, W, H, s0, ns, maxDepth, seed, row0, row1, col0, col1, nThreads, st, strict, beta, pathPrim, pathT, pathStride);
      else r = renderStreams<double, double>(s, cam, W, H, s0, ns, maxDepth, seed, row0, row1, col0, col1, nThreads, st, strict, beta, pathPrim, pathT, pathStride);
    }
    std::memcpy(out, r.sum.data(), r.sum.size() * sizeof(double));
    statsOut(r.stats, stats24);
    return 0;
  } catch (const std::exception& e) { g_err = e.what(); return -1; }
}

// textureValue probe (Lib.hs:496-513): q = n x (u, v, px, py, pz) -> out n x 3
int orc_texture_value(void* p, int texId, int prec, int flags, int n, const double* q, double* out) {
  const bool strict = (flags & 1) != 0;
  Scene& s = ((Handle*)p)->scene;
  if (texId < 0 || texId >= (int)s.texs.size()) { g_err = "bad texture id"; return -1; }
  SMGen g{0, 1};
  for (int i = 0; i < n; ++i) {
    const double* a = q + (size_t)i * 5;
    if (prec == 2) {
      Tracer<double, double> t(s, g, nullptr, strict);
      auto c = t.textureValue(s.texs[texId], a[0], a[1], Vec3<double>(a[2], a[3], a[4]));
      out[i * 3] = c.x; out[i * 3 + 1] = c.y; out[i * 3 + 2] = c.z;
    } else {
      Tracer<float, float> t(s, g, nullptr, strict);
      auto c = t.textureValue(s.texs[texId], (float)a[0], (float)a[1], Vec3<float>((float)a[2], (float)a[3], (float)a[4]));
      out[i * 3] = c.x; out[i * 3 + 1] = c.y; out[i * 3 + 2] = c.z;
    }
  }
  return 0;
}

// the shared strict-math functions on the host (fn: 0 sin 1 cos 2 atan2(x, y) 3 asin 4 log) for the host/device identity test
void orc_strict_math_eval(int fn, int n, const double* x, const double* y, double* out, int asFloat) {
  for (int i = 0; i < n; ++i) {
    if (asFloat) {
      const float a = (float)x[i], b = (float)y[i];
      out[i] = fn == 0 ? rtb_strict::sinS(a) : fn == 1 ? rtb_strict::cosS(a) : fn == 2 ? rtb_strict::atan2S(a, b) : fn == 3 ? rtb_strict::asinS(a) : rtb_strict::logS(a);
    } else {
      const double a = x[i], b = y[i];
      out[i] = fn == 0 ? rtb_strict::sinS(a) : fn == 1 ? rtb_strict::cosS(a) : fn == 2 ? rtb_strict::atan2S(a, b) : fn == 3 ? rtb_strict::asinS(a) : rtb_strict::logS(a);
    }
  }
}

// RNG known-answer helpers
void orc_splitmix(int64_t seed, int n, uint64_t* words, double* doubles) {
  SMGen a = randGen(seed), b = a;
  for (int i = 0; i < n; ++i) { words[i] = nextWord64(a); doubles[i] = nextReal<double>(b); }
}
void orc_stream(uint64_t seed, uint32_t pixel, uint32_t sample, int n, uint64_t* state2, double* doubles, float* floats) {
  SMGen a = streamGen(seed, pixel, sample), b = a;
  state2[0] = a.seed; state2[1] = a.gamma;
  for (int i = 0; i < n; ++i) { doubles[i] = nextReal<double>(a); floats[i] = nextReal<float>(b); }
}
// albedoToColor on means (Lib.hs:287-292)
void orc_albedo_to_color(int n, const double* mean, uint8_t* out) { for (int i = 0; i < n; ++i) out[i] = scaleColor(mean[i]); }

}  // extern "C"
