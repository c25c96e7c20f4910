// oracle/rt_oracle.hpp — CPU restatement of shaunplee/ray-tracing @ HEAD (hot path + input producers).
//
// TEST INFRASTRUCTURE ONLY.  Nothing under ray-tracing_b200/ may include, link or call this file; only
// tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs use it, as the checker
// or as the timed CPU baseline.
//
// PARITY UNPINNED: the reference has no tests or golden vectors (test/Spec.hs:1-2), and no Haskell toolchain
// exists in this image, so this restatement could not be checked against the reference's own execution.  The
// only reference artefact it is compared with is cornellBox1000.png (statistically; tests/test_oracle.py).
//
// Every function cites the reference file:line it follows (paths relative to /root/reference).  The code is a
// deliberately literal, recursive transcription: same operation order, same comparison directions, Haskell
// min/max NaN behaviour, same RNG draw order.  It is templated on
//     T = general arithmetic type (double = the reference; float = the device's fp32 arithmetic),
//     S = type used for the sphere quadratic only (S = T is literal; S = double with T = float restates the
//         device's optional "mixed" mode).
// Scene construction is always done in double (it stays on the host in the product, too); the hot path casts a
// stored field to T where it is used, which is identical to pre-rounding the flattened scene to T.
#pragma once
#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <limits>
#include <memory>
#include <stdexcept>
#include <string>
#include <thread>
#include <vector>

// the product's shared strict-math header (test infrastructure may depend on the product, never the other way round)
#include "../ray-tracing_b200/csrc/rt_strict_math.h"

namespace orc {

// =====================================================================================================
// RNG: splitmix-0.1 SMGen + random-1.2.0 Double  (Random.hs:11-29; stack.yaml:42-44 pins the versions;
// neither package is vendored under /root/reference, the algorithm below restates their published source)
// =====================================================================================================
struct SMGen { uint64_t seed, gamma; };
static constexpr uint64_t kGoldenGamma = 0x9e3779b97f4a7c15ULL;

inline uint64_t shiftXor(int n, uint64_t w) { return w ^ (w >> n); }
inline uint64_t shiftXorMultiply(int n, uint64_t k, uint64_t w) { return shiftXor(n, w) * k; }
// splitmix-0.1 System.Random.SplitMix.mix64 (MurmurHash3 finaliser)
inline uint64_t mix64(uint64_t z) {
  z = shiftXorMultiply(33, 0xff51afd7ed558ccdULL, z);
  z = shiftXorMultiply(33, 0xc4ceb9fe1a85ec53ULL, z);
  return shiftXor(33, z);
}
// splitmix-0.1 mix64variant13 (Stafford mix13), used only by mixGamma
inline uint64_t mix64variant13(uint64_t z) {
  z = shiftXorMultiply(30, 0xbf58476d1ce4e5b9ULL, z);
  z = shiftXorMultiply(27, 0x94d049bb133111ebULL, z);
  return shiftXor(31, z);
}
inline uint64_t mixGamma(uint64_t z) {
  uint64_t z1 = mix64variant13(z) | 1ULL;
  int n = __builtin_popcountll(z1 ^ (z1 >> 1));
  return n >= 24 ? z1 : (z1 ^ 0xaaaaaaaaaaaaaaaaULL);
}
inline SMGen mkSMGen(uint64_t s) { return SMGen{mix64(s), mixGamma(s + kGoldenGamma)}; }
inline uint64_t nextWord64(SMGen& g) { g.seed += g.gamma; return mix64(g.seed); }
// Random.hs:20-21  randGen s = mkStdGen s  (random-1.2.0: StdGen . mkSMGen . fromIntegral)
inline SMGen randGen(int64_t s) { return mkSMGen((uint64_t)s); }

// The product's counter-based stream for camera sample (pixel, sample): an SMGen seeded the mkSMGen way from
// a hashed stream id.  NOT in the reference (its per-column sequential streams are Lib.hs:1491-1523); restated
// here so that oracle and device can be compared sample-for-sample (DESIGN.md "RNG").
inline SMGen streamGen(uint64_t seed, uint32_t pixel, uint32_t sample) {
  uint64_t u = ((uint64_t)pixel << 32) | (uint64_t)sample;
  return mkSMGen(mix64variant13(seed + kGoldenGamma * (u + 1ULL)));
}

// Random.hs:23-25 randomDouble = System.Random.random :: Double.  random-1.2.0: uniformRM (0,1) =
// (h-l)*word64ToDoubleUpper w + l with word64ToDoubleUpper w = castWord64ToDouble((w>>12)|0x3ff0..0) - 1.
template <class T> inline T nextReal(SMGen& g);
template <> inline double nextReal<double>(SMGen& g) {
  uint64_t w = nextWord64(g);
  uint64_t bits = (w >> 12) | 0x3ff0000000000000ULL;
  double d; std::memcpy(&d, &bits, 8);
  return (1.0 - 0.0) * (d - 1.0) + 0.0;
}
// fp32 arithmetic: same word, top 23 bits -> [1,2) - 1  (the double value truncated to float precision)
template <> inline float nextReal<float>(SMGen& g) {
  uint64_t w = nextWord64(g);
  uint32_t bits = (uint32_t)(w >> 41) | 0x3f800000u;
  float f; std::memcpy(&f, &bits, 4);
  return f - 1.0f;
}

// Haskell Prelude min/max on Double: max x y = if x <= y then y else x ; min x y = if x <= y then x else y
template <class T> inline T hmax(T x, T y) { return (x <= y) ? y : x; }
template <class T> inline T hmin(T x, T y) { return (x <= y) ? x : y; }

template <class T> struct K {  // constants in T
  static constexpr T epsilon() { return (T)0.0001; }          // Lib.hs:76-77
  static constexpr T infinity() { return std::numeric_limits<T>::infinity(); }  // Lib.hs:79-80
  static constexpr T pi() { return (T)3.14159265358979323846; }
};

// =====================================================================================================
// Vec3 (Lib.hs:200-261), ONB (263-279), clamp (281-285)
// =====================================================================================================
template <class T> struct Vec3 {
  T x, y, z;
  Vec3() : x(0), y(0), z(0) {}
  Vec3(T a, T b, T c) : x(a), y(b), z(c) {}
  template <class U> explicit Vec3(const Vec3<U>& o) : x((T)o.x), y((T)o.y), z((T)o.z) {}
};
template <class T> inline Vec3<T> operator*(const Vec3<T>& a, const Vec3<T>& b) { return {a.x * b.x, a.y * b.y, a.z * b.z}; }  // |*| 224
template <class T> inline Vec3<T> operator+(const Vec3<T>& a, const Vec3<T>& b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }  // |+| 231
template <class T> inline Vec3<T> operator-(const Vec3<T>& a, const Vec3<T>& b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }  // |-| 235
template <class T> inline Vec3<T> vecNegate(const Vec3<T>& a) { return {-a.x, -a.y, -a.z}; }                                    // 238
template <class T> inline T squaredLength(const Vec3<T>& a) { return (a.x * a.x) + (a.y * a.y) + (a.z * a.z); }                // 244
template <class T> inline T length(const Vec3<T>& a) { return std::sqrt(squaredLength(a)); }                                   // 241
template <class T> inline Vec3<T> scale(T k, const Vec3<T>& a) { return {a.x * k, a.y * k, a.z * k}; }                         // 250
template <class T> inline Vec3<T> divide(const Vec3<T>& a, T k) { return {a.x / k, a.y / k, a.z / k}; }                        // 253
template <class T> inline Vec3<T> makeUnitVector(const Vec3<T>& a) { return divide(a, length(a)); }                            // 247
template <class T> inline T dot(const Vec3<T>& a, const Vec3<T>& b) { return a.x * b.x + a.y * b.y + a.z * b.z; }              // 256
template <class T> inline Vec3<T> cross(const Vec3<T>& a, const Vec3<T>& b) {                                                  // 259
  return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x};
}
template <class T> struct ONB { Vec3<T> u, v, w; };
template <class T> inline Vec3<T> onbLocalV(const ONB<T>& o, const Vec3<T>& a) {  // 269-270
  return scale(a.x, o.u) + scale(a.y, o.v) + scale(a.z, o.w);
}
template <class T> inline ONB<T> onbFromW(const Vec3<T>& n) {  // 272-279
  Vec3<T> w = makeUnitVector(n);
  Vec3<T> a = (std::fabs(w.x) > (T)0.9) ? Vec3<T>(0, 1, 0) : Vec3<T>(1, 0, 0);
  Vec3<T> v = makeUnitVector(cross(w, a));
  Vec3<T> u = cross(w, v);
  return ONB<T>{u, v, w};
}
template <class T> inline T clampH(T mn, T mx, T x) { return (x < mn) ? mn : ((x > mx) ? mx : x); }  // 281-285

using V3d = Vec3<double>;

// =====================================================================================================
// Scene ADTs (Lib.hs:339-351, 394-419, 515-660, 793-796, 1230-1251).  Fields are stored in double.
// `id` is the canonical index of the object within its kind (assigned by Scene::number()).
// =====================================================================================================
struct Box { V3d mn, mx; };

struct PerlinTable {  // Lib.hs:401-410 (ranvec, permX, permY, permZ)
  double ranvec[256][3];
  int perm[3][256];
  int id = -1;
};
struct ImageData {  // Lib.hs:384-389: JP.Image PixelRGB8
  int w = 0, h = 0;
  std::vector<uint8_t> rgb;
  int id = -1;
};
struct Texture {
  enum Kind { Constant = 0, Checker = 1, Perlin = 2, Image = 3 } kind = Constant;
  double rgb[3] = {0, 0, 0};
  std::shared_ptr<Texture> odd, even;
  std::shared_ptr<PerlinTable> perlin;
  double scale = 1.0;
  std::shared_ptr<ImageData> image;  // null == ImageTexture Nothing
  int nx = 0, ny = 0;
  int id = -1;
};
using TexP = std::shared_ptr<Texture>;
struct Material {
  enum Kind { Lambertian = 0, Metal = 1, Dielectric = 2, DiffuseLight = 3, Isotropic = 4 } kind = Lambertian;
  TexP tex;
  double param = 0;  // Fuzz / RefractiveIdx
  int id = -1;
};
using MatP = std::shared_ptr<Material>;

struct Hittable;
using HitP = std::shared_ptr<Hittable>;
struct Hittable {
  enum Kind { Unhittable = 0, Sphere = 1, MovingSphere = 2, RectXY = 3, RectXZ = 4, RectYZ = 5, Cuboid = 6,
              BVHNode = 7, Translate = 8, Rotate = 9, ConstantMedium = 10, NKinds = 11 } kind = Unhittable;
  // Sphere: c0, r ; MovingSphere: c0, c1, t0, t1, dur, r
  V3d c0, c1; double t0 = 0, t1 = 0, dur = 0, r = 0;
  // Rect: a0 a1 b0 b1 k  (XY: x0 x1 y0 y1 ; XZ: x0 x1 z0 z1 ; YZ: y0 y1 z0 z1)
  double a0 = 0, a1 = 0, b0 = 0, b1 = 0, k = 0;
  // Cuboid: bmin bmax (six Rects are derived, Lib.hs:594-605)
  V3d bmin, bmax;
  // BVHNode: left right box size ; Translate: off child(left) ; Rotate: axis sin cos box child(left)
  // ConstantMedium: nInvD mat boundary(left)
  HitP left, right; Box box; int size = 0;
  V3d off; int axis = 0; double sinT = 0, cosT = 0; double nInvD = 0;
  MatP mat;
  int id = -1;
};
inline uint32_t refOf(const Hittable* h) { return ((uint32_t)h->kind << 28) | (uint32_t)(h->id & 0x0fffffff); }
// the three Rect kinds share one index space (one flattened rect array; the ref's kind carries the plane)
inline int slotOf(int kind) { return (kind == Hittable::RectXZ || kind == Hittable::RectYZ) ? (int)Hittable::RectXY : kind; }

// ---- smart constructors (Lib.hs:587-660, 726-761, 789-791) ----
inline TexP constantColor(double r, double g, double b) { auto t = std::make_shared<Texture>(); t->kind = Texture::Constant; t->rgb[0] = r; t->rgb[1] = g; t->rgb[2] = b; return t; }
inline TexP checkerTexture(TexP odd, TexP even) { auto t = std::make_shared<Texture>(); t->kind = Texture::Checker; t->odd = odd; t->even = even; return t; }
inline TexP imageTexture(std::shared_ptr<ImageData> im, int w, int h) { auto t = std::make_shared<Texture>(); t->kind = Texture::Image; t->image = im; t->nx = w; t->ny = h; return t; }
inline MatP mkMat(Material::Kind k, TexP tex, double p = 0) { auto m = std::make_shared<Material>(); m->kind = k; m->tex = tex; m->param = p; return m; }
inline MatP lambertian(TexP t) { return mkMat(Material::Lambertian, t); }
inline MatP metal(TexP t, double fuzz) { return mkMat(Material::Metal, t, fuzz); }
inline MatP dielectric(double ri) { return mkMat(Material::Dielectric, nullptr, ri); }
inline MatP diffuseLight(TexP t) { return mkMat(Material::DiffuseLight, t); }
inline MatP isotropic(TexP t) { return mkMat(Material::Isotropic, t); }

inline HitP unhittable() { return std::make_shared<Hittable>(); }
inline HitP sphere(V3d c, double r, MatP m) {  // 587-588
  auto h = std::make_shared<Hittable>(); h->kind = Hittable::Sphere; h->c0 = c; h->r = r; h->mat = m; return h;
}
inline HitP movingSphere(V3d c0, V3d c1, double t0, double t1, double r, MatP m) {  // 590-592
  auto h = std::make_shared<Hittable>(); h->kind = Hittable::MovingSphere; h->c0 = c0; h->c1 = c1; h->t0 = t0; h->t1 = t1;
  h->dur = t1 - t0; h->r = r; h->mat = m; return h;
}
enum Plane { XYPlane = 0, XZPlane = 1, YZPlane = 2 };
inline HitP rect(Plane p, double a0, double a1, double b0, double b1, double k, MatP m) {  // 649-660
  auto h = std::make_shared<Hittable>();
  h->kind = p == XYPlane ? Hittable::RectXY : (p == XZPlane ? Hittable::RectXZ : Hittable::RectYZ);
  h->a0 = a0; h->a1 = a1; h->b0 = b0; h->b1 = b1; h->k = k; h->mat = m; return h;
}
inline HitP cuboid(V3d mn, V3d mx, MatP m) {  // 594-605
  auto h = std::make_shared<Hittable>(); h->kind = Hittable::Cuboid; h->bmin = mn; h->bmax = mx; h->mat = m; return h;
}
inline HitP translate(V3d off, HitP c) {  // 726-727
  auto h = std::make_shared<Hittable>(); h->kind = Hittable::Translate; h->off = off; h->left = c; return h;
}
enum Axis { XAxis = 0, YAxis = 1, ZAxis = 2 };

template <class T> inline Vec3<T> rotatePoint(int axis, T s, T c, const Vec3<T>& p) {  // 763-774
  switch (axis) {
    case XAxis: return {p.x, c * p.y - s * p.z, s * p.y + c * p.z};
    case YAxis: return {c * p.x + s * p.z, p.y, -s * p.x + c * p.z};
    default:    return {c * p.x - s * p.y, s * p.x + c * p.y, p.z};
  }
}
template <class T> inline Vec3<T> unRotatePoint(int axis, T s, T c, const Vec3<T>& p) {  // 776-787
  switch (axis) {
    case XAxis: return {p.x, c * p.y + s * p.z, -s * p.y + c * p.z};
    case YAxis: return {c * p.x - s * p.z, p.y, s * p.x + c * p.z};
    default:    return {c * p.x + s * p.y, -s * p.x + c * p.y, p.z};
  }
}

inline Box surroundingBox(const Box& a, const Box& b) {  // 929-939
  return Box{V3d(hmin(a.mn.x, b.mn.x), hmin(a.mn.y, b.mn.y), hmin(a.mn.z, b.mn.z)),
             V3d(hmax(a.mx.x, b.mx.x), hmax(a.mx.y, b.mx.y), hmax(a.mx.z, b.mx.z))};
}
inline Box boundingBox(const Hittable* h) {  // 905-927 (the Maybe (Time,Time) argument is never inspected)
  const double eps = 0.0001;
  switch (h->kind) {
    case Hittable::Sphere: { V3d rad(h->r, h->r, h->r); return Box{h->c0 - rad, h->c0 + rad}; }
    case Hittable::MovingSphere: {
      V3d rad(h->r, h->r, h->r);
      return surroundingBox(Box{h->c0 - rad, h->c0 + rad}, Box{h->c1 - rad, h->c1 + rad});
    }
    case Hittable::RectXY: return Box{V3d(h->a0, h->b0, h->k - eps), V3d(h->a1, h->b1, h->k + eps)};
    case Hittable::RectXZ: return Box{V3d(h->a0, h->k - eps, h->b0), V3d(h->a1, h->k + eps, h->b1)};
    case Hittable::RectYZ: return Box{V3d(h->k - eps, h->a0, h->b0), V3d(h->k + eps, h->a1, h->b1)};
    case Hittable::BVHNode: return h->box;
    case Hittable::Cuboid: return Box{h->bmin, h->bmax};
    case Hittable::Translate: { Box b = boundingBox(h->left.get()); return Box{b.mn + h->off, b.mx + h->off}; }
    case Hittable::Rotate: return h->box;
    case Hittable::ConstantMedium: return boundingBox(h->left.get());
    default: throw std::runtime_error("Should not be trying to bound an Unhittable");  // 927
  }
}
inline HitP rotate(Axis axis, double angle, HitP c) {  // 729-761
  auto h = std::make_shared<Hittable>(); h->kind = Hittable::Rotate; h->axis = axis; h->left = c;
  double rad = angle * 3.14159265358979323846 / 180.0;
  h->sinT = std::sin(rad); h->cosT = std::cos(rad);
  Box bb = boundingBox(c.get());
  const double inf = std::numeric_limits<double>::infinity();
  V3d mn(inf, inf, inf), mx(-inf, -inf, -inf);
  // foldr over [(i,j,k) | i<-[0,1,2], j<-[0,1,2], k<-[0,1,2]] : last tuple first; min/max are order-free here
  for (int i = 2; i >= 0; --i) for (int j = 2; j >= 0; --j) for (int k = 2; k >= 0; --k) {
    double x = i * bb.mx.x + (1 - i) * bb.mn.x;
    double y = j * bb.mx.y + (1 - j) * bb.mn.y;
    double z = k * bb.mx.z + (1 - k) * bb.mn.z;
    V3d n = rotatePoint<double>(axis, h->sinT, h->cosT, V3d(x, y, z));
    mn = V3d(hmin(n.x, mn.x), hmin(n.y, mn.y), hmin(n.z, mn.z));
    mx = V3d(hmax(n.x, mx.x), hmax(n.y, mx.y), hmax(n.z, mx.z));
  }
  h->box = Box{mn, mx};
  return h;
}
inline HitP constantMedium(double density, TexP tex, HitP boundary) {  // 789-791
  auto h = std::make_shared<Hittable>(); h->kind = Hittable::ConstantMedium;
  h->nInvD = -1 / density; h->mat = isotropic(tex); h->left = boundary; return h;
}
inline int htblSize(const Hittable* h) {  // 662-671
  switch (h->kind) {
    case Hittable::Unhittable: return 0;
    case Hittable::Translate: case Hittable::Rotate: return htblSize(h->left.get());
    case Hittable::BVHNode: return h->size;
    default: return 1;
  }
}

// ---- host-side random helpers (Lib.hs:1119-1150) on an SMGen, double precision ----
inline double randomDoubleM(SMGen& g) { return nextReal<double>(g); }                                        // 1119-1125
inline double randomDoubleRM(SMGen& g, double mn, double mx) { double rd = randomDoubleM(g); return mn + (mx - mn) * rd; }  // 1127-1130
inline int randomIntRM(SMGen& g, int lo, int hi) { return (int)std::floor(randomDoubleRM(g, (double)lo, (double)hi)); }    // 1132-1133
inline V3d randomVec3DoubleM(SMGen& g) { double x = randomDoubleM(g), y = randomDoubleM(g), z = randomDoubleM(g); return V3d(x, y, z); }  // 1135-1143
inline V3d randomVec3DoubleRM(SMGen& g, double mn, double mx) {                                              // 1145-1150
  double x = randomDoubleRM(g, mn, mx), y = randomDoubleRM(g, mn, mx), z = randomDoubleRM(g, mn, mx); return V3d(x, y, z);
}

// makeBVH (Lib.hs:941-961) + boxCompare (963-968)
inline HitP makeBVH(SMGen& g, std::vector<HitP> htbls) {
  int axis = (int)std::floor(randomDoubleRM(g, 0, 3));
  auto key = [axis](const HitP& h) { Box b = boundingBox(h.get()); return axis == 0 ? b.mn.x : (axis == 1 ? b.mn.y : b.mn.z); };
  size_t span = htbls.size();
  HitP lt, rt;
  if (span == 0) throw std::runtime_error("makeBVH: empty sequence");
  if (span == 1) { lt = htbls[0]; rt = htbls[0]; }
  else if (span == 2) {
    if (key(htbls[0]) < key(htbls[1])) { lt = htbls[0]; rt = htbls[1]; } else { lt = htbls[1]; rt = htbls[0]; }
  } else {
    std::vector<std::pair<double, HitP>> keyed; keyed.reserve(span);
    for (auto& h : htbls) keyed.emplace_back(key(h), h);
    std::stable_sort(keyed.begin(), keyed.end(), [](const auto& a, const auto& b) { return a.first < b.first; });  // S.sortBy is stable
    size_t half = span / 2;
    std::vector<HitP> l, r; l.reserve(half); r.reserve(span - half);
    for (size_t i = 0; i < span; ++i) (i < half ? l : r).push_back(keyed[i].second);
    keyed.clear();
    lt = makeBVH(g, std::move(l));
    rt = makeBVH(g, std::move(r));
  }
  auto n = std::make_shared<Hittable>(); n->kind = Hittable::BVHNode; n->left = lt; n->right = rt;
  n->box = surroundingBox(boundingBox(lt.get()), boundingBox(rt.get())); n->size = (int)span;
  return n;
}

// makePerlin (Lib.hs:421-439)
inline TexP makePerlin(SMGen& g, double sc) {
  auto tab = std::make_shared<PerlinTable>();
  for (int i = 0; i < 256; ++i) { V3d v = randomVec3DoubleRM(g, -1.0, 1.0); tab->ranvec[i][0] = v.x; tab->ranvec[i][1] = v.y; tab->ranvec[i][2] = v.z; }
  for (int a = 0; a < 3; ++a) {
    for (int i = 0; i < 256; ++i) tab->perm[a][i] = i;
    for (int i = 255; i >= 1; --i) { int target = randomIntRM(g, 0, i); std::swap(tab->perm[a][i], tab->perm[a][target]); }
  }
  auto t = std::make_shared<Texture>(); t->kind = Texture::Perlin; t->perlin = tab; t->scale = sc; return t;
}

// Camera (Lib.hs:1230-1251) + newCamera (1269-1295)
struct Camera { V3d origin, llc, horiz, vert, u, v, w; double lensRadius, t0, t1; };
inline Camera newCamera(V3d lookfrom, V3d lookat, V3d vup, double vfov, double aspect, double aperture, double focusDist, double t0, double t1) {
  const double pi = 3.14159265358979323846;
  double lensRadius = aperture / 2.0;
  double theta = vfov * pi / 180.0;
  double halfHeight = std::tan(theta / 2.0);
  double halfWidth = aspect * halfHeight;
  V3d origin = lookfrom;
  V3d w = makeUnitVector(lookfrom - lookat);
  V3d u = makeUnitVector(cross(vup, w));
  V3d v = cross(w, u);
  V3d llc = origin - scale(halfWidth * focusDist, u) - scale(halfHeight * focusDist, v) - scale(focusDist, w);
  V3d horizontal = scale(2 * halfWidth * focusDist, u);
  V3d vertical = scale(2 * halfHeight * focusDist, v);
  return Camera{origin, llc, horizontal, vertical, u, v, w, lensRadius, t0, t1};
}

struct Scene {  // Lib.hs:84  type Scene = (Hittable, Hittable, Albedo)
  HitP world, lights; double background[3] = {0, 0, 0};
  SMGen g1{0, 0};  // generator state returned with the scene (Scenes.hs: (Scene, RandGen))
  // canonical numbering (test convention, DESIGN.md "ids"): DFS pre-order from world then lights,
  // left before right, dedup by object identity, one counter per kind; materials/textures/tables likewise.
  std::vector<const Hittable*> byKind[Hittable::NKinds];
  std::vector<const Material*> mats; std::vector<const Texture*> texs;
  std::vector<const PerlinTable*> perlins; std::vector<const ImageData*> images;
  void numberTex(Texture* t) {
    if (!t || t->id >= 0) return;
    t->id = (int)texs.size(); texs.push_back(t);
    if (t->kind == Texture::Checker) { numberTex(t->odd.get()); numberTex(t->even.get()); }
    if (t->kind == Texture::Perlin && t->perlin->id < 0) { t->perlin->id = (int)perlins.size(); perlins.push_back(t->perlin.get()); }
    if (t->kind == Texture::Image && t->image && t->image->id < 0) { t->image->id = (int)images.size(); images.push_back(t->image.get()); }
  }
  void numberMat(Material* m) {
    if (!m || m->id >= 0) return;
    m->id = (int)mats.size(); mats.push_back(m); numberTex(m->tex.get());
  }
  void numberH(Hittable* h) {
    if (h->id >= 0) return;
    h->id = (int)byKind[slotOf(h->kind)].size(); byKind[slotOf(h->kind)].push_back(h);
    numberMat(h->mat.get());
    if (h->left) numberH(h->left.get());
    if (h->right) numberH(h->right.get());
  }
  void number() { numberH(world.get()); numberH(lights.get()); }
};

// =====================================================================================================
// Hot path
// =====================================================================================================
template <class T> struct Ray { Vec3<T> o, d; T time; };                     // Lib.hs:307-315
template <class T> inline Vec3<T> at(const Ray<T>& r, T t) { return r.o + scale(t, r.d); }  // 317-318
template <class T> struct Hit {                                              // 320-337
  T t; Vec3<T> p, normal; T u, v; bool frontFace; const Material* mat; uint32_t prim;
};

// Instrumentation (not in the reference): visit log in `hit`-call order + algorithmic-work counters
struct Trace {
  bool logVisits = false;
  std::vector<uint32_t> visits;
  uint64_t nVisit = 0, visitHash = 0xcbf29ce484222325ULL;
  // counters (BASELINE.md "algorithmic work"): box tests, primitive tests by kind, sphere hit records, bounces
  uint64_t boxTests = 0, prim[Hittable::NKinds] = {0}, sphereHits = 0, bounces = 0, paths = 0, noiseCalls = 0, texels = 0, queries = 0;
  // shading events: [0..3] scatter Lambertian / Metal / Dielectric / Isotropic, [4] emitted, [5] htblRandom on the light list,
  // [6] htblPdfValue leaf formulas evaluated, [7] checker levels evaluated
  uint64_t shade[8] = {0};
  // per-path log of the main queries (rayColor's `hit world`), bounce order: (prim ref | 0xffffffff = miss, t) — same definition as
  // the device's rt_render_pathlog; filled when pathPrim != nullptr
  uint32_t* pathPrim = nullptr; double* pathT = nullptr; uint32_t pathStride = 0, pathBounce = 0;
  inline void bounce(uint32_t prim, double t) { if (pathPrim) { if (pathBounce < pathStride) { pathPrim[pathBounce] = prim; pathT[pathBounce] = t; } ++pathBounce; } }
  inline void visit(const Hittable* h) {
    uint32_t r = refOf(h);
    ++nVisit; visitHash = (visitHash ^ r) * 0x100000001b3ULL;
    if (logVisits) visits.push_back(r);
    ++prim[h->kind];
  }
  void add(const Trace& o) {
    boxTests += o.boxTests; sphereHits += o.sphereHits; bounces += o.bounces; paths += o.paths; noiseCalls += o.noiseCalls;
    texels += o.texels; queries += o.queries; nVisit += o.nVisit;
    for (int i = 0; i < Hittable::NKinds; ++i) prim[i] += o.prim[i];
    for (int i = 0; i < 8; ++i) shade[i] += o.shade[i];
  }
};

template <class T> inline void faceNormal(const Ray<T>& r, const Vec3<T>& outward, bool& front, Vec3<T>& n) {  // 1111-1117
  front = dot(r.d, outward) < 0;
  n = front ? outward : vecNegate(outward);
}

template <class T> inline bool boxRayIntersect(const Box& b, const Ray<T>& r, T t_min, T t_max) {  // 798-814
  const T ro[3] = {r.o.x, r.o.y, r.o.z}, rd[3] = {r.d.x, r.d.y, r.d.z};
  const T mn[3] = {(T)b.mn.x, (T)b.mn.y, (T)b.mn.z}, mx[3] = {(T)b.mx.x, (T)b.mx.y, (T)b.mx.z};
  for (int a = 0; a < 3; ++a) {  // `all` short-circuits in x,y,z order
    T ta = (mn[a] - ro[a]) / rd[a];
    T tb = (mx[a] - ro[a]) / rd[a];
    T t0, t1;
    if (ta < tb) { t0 = ta; t1 = tb; } else { t0 = tb; t1 = ta; }
    T tmin = hmax(t0, t_min);
    T tmax = hmin(t1, t_max);
    if (!(tmax > tmin)) return false;
  }
  return true;
}

template <class T, class S> struct Tracer {
  const Scene& sc; SMGen& g; Trace* tr;
  // strict: sin / cos / atan2 / asin / log from the shared plain-arithmetic header (bit-identical to the device under
  // RT_FLAG_STRICT_MATH) and schlick's fifth power as products; betaOrder: throughput accumulated forward (the device's order)
  // instead of the reference's multiply-on-return chain (Lib.hs:1321-1333).  Both default to the reference's literal form.
  bool strict = false, betaOrder = false;
  Tracer(const Scene& s, SMGen& gen, Trace* t, bool strictMath = false, bool beta = false) : sc(s), g(gen), tr(t), strict(strictMath), betaOrder(beta) {}
  T msin(T x) const { return strict ? rtb_strict::sinS(x) : std::sin(x); }
  T mcos(T x) const { return strict ? rtb_strict::cosS(x) : std::cos(x); }
  T matan2(T y, T x) const { return strict ? rtb_strict::atan2S(y, x) : std::atan2(y, x); }
  T masin(T x) const { return strict ? rtb_strict::asinS(x) : std::asin(x); }
  T mlog(T x) const { return strict ? rtb_strict::logS(x) : std::log(x); }

  // ---- rectHit (Lib.hs:1005-1028) ; plane: 0 XY (i=x,j=y,k=z) 1 XZ (i=x,j=z,k=y) 2 YZ (i=y,j=z,k=x)
  bool rectHit(int plane, T i0, T i1, T j0, T j1, T k, const Material* mat, uint32_t prim, const Ray<T>& r, T t_min, T t_max, Hit<T>& out) {
    auto sel = [](const Vec3<T>& v, int c) { return c == 0 ? v.x : (c == 1 ? v.y : v.z); };
    int ci = plane == 2 ? 1 : 0, cj = plane == 0 ? 1 : 2, ck = plane == 0 ? 2 : (plane == 1 ? 1 : 0);
    T t = (k - sel(r.o, ck)) / sel(r.d, ck);
    if ((t < t_min) || (t > t_max)) return false;
    T i = sel(r.o, ci) + t * sel(r.d, ci);
    T j = sel(r.o, cj) + t * sel(r.d, cj);
    if ((i < i0) || (i > i1) || (j < j0) || (j > j1)) return false;
    out.u = (i - i0) / (i1 - i0);
    out.v = (j - j0) / (j1 - j0);
    out.t = t; out.p = at(r, t);
    Vec3<T> outward = plane == 0 ? Vec3<T>(0, 0, 1) : (plane == 1 ? Vec3<T>(0, 1, 0) : Vec3<T>(1, 0, 0));
    faceNormal(r, outward, out.frontFace, out.normal);
    out.mat = mat; out.prim = prim;
    return true;
  }

  // ---- sphere (Lib.hs:1081-1105); quadratic evaluated in S
  bool sphereHit(const Vec3<T>& scT, T srT, const Material* mat, uint32_t prim, const Ray<T>& r, T t_min, T t_max, Hit<T>& out) {
    Vec3<S> ro(r.o), rd(r.d), scS(scT);
    S sr = (S)srT;
    Vec3<S> oc = ro - scS;
    S a = dot(rd, rd);
    S b = dot(oc, rd);
    S c = dot(oc, oc) - (sr * sr);
    S disc = b * b - a * c;
    if (disc > 0) {
      S sd = std::sqrt(disc);
      S temp1 = ((-b) - sd) / a;
      S temp2 = ((-b) + sd) / a;
      S temp;
      if ((S)t_min < temp1 && temp1 < (S)t_max) temp = temp1;
      else if ((S)t_min < temp2 && temp2 < (S)t_max) temp = temp2;
      else return false;
      if (tr) ++tr->sphereHits;
      T t = (T)temp;
      Vec3<T> p = at(r, t);
      Vec3<T> pShift = p - scT;
      Vec3<T> outward = divide(pShift, srT);
      faceNormal(r, outward, out.frontFace, out.normal);
      T phi = matan2(outward.z, outward.x);
      T theta = masin(outward.y);
      out.u = (T)1.0 - ((phi + K<T>::pi()) / ((T)2 * K<T>::pi()));
      out.v = (theta + (K<T>::pi() / (T)2)) / K<T>::pi();
      out.t = t; out.p = p; out.mat = mat; out.prim = prim;
      return true;
    }
    return false;
  }

  // ---- hit (Lib.hs:970-1109)
  bool hit(const Hittable* h, const Ray<T>& r, T t_min, T t_max, Hit<T>& out) {
    if (tr) tr->visit(h);
    switch (h->kind) {
      case Hittable::BVHNode: {  // 971-988
        if (tr) ++tr->boxTests;
        if (!boxRayIntersect<T>(h->box, r, t_min, t_max)) return false;
        Hit<T> hl;
        if (!hit(h->left.get(), r, t_min, t_max, hl)) return hit(h->right.get(), r, t_min, t_max, out);
        Hit<T> hr;
        if (hit(h->right.get(), r, t_min, hl.t, hr)) { out = hr; return true; }
        out = hl; return true;
      }
      case Hittable::Cuboid: {  // 989-1004 ; list = [XY z1, XY z0, XZ y1, XZ y0, YZ x1, YZ x0], foldr => last first
        const V3d &mn = h->bmin, &mx = h->bmax;
        uint32_t prim = refOf(h);
        bool have = false; Hit<T> acc{};
        struct F { int plane; double i0, i1, j0, j1, k; };
        const F faces[6] = {{0, mn.x, mx.x, mn.y, mx.y, mx.z}, {0, mn.x, mx.x, mn.y, mx.y, mn.z},
                            {1, mn.x, mx.x, mn.z, mx.z, mx.y}, {1, mn.x, mx.x, mn.z, mx.z, mn.y},
                            {2, mn.y, mx.y, mn.z, mx.z, mx.x}, {2, mn.y, mx.y, mn.z, mx.z, mn.x}};
        for (int f = 5; f >= 0; --f) {
          Hit<T> h1;
          if (rectHit(faces[f].plane, (T)faces[f].i0, (T)faces[f].i1, (T)faces[f].j0, (T)faces[f].j1, (T)faces[f].k, h->mat.get(), prim, r, t_min, t_max, h1)) {
            if (!have) { acc = h1; have = true; }
            else if (h1.t < acc.t) acc = h1;  // closerHit: new wins only if strictly closer (1002)
          }
        }
        if (have) out = acc;
        return have;
      }
      case Hittable::RectXY: case Hittable::RectXZ: case Hittable::RectYZ:  // 1005-1012
        return rectHit(h->kind - Hittable::RectXY, (T)h->a0, (T)h->a1, (T)h->b0, (T)h->b1, (T)h->k, h->mat.get(), refOf(h), r, t_min, t_max, out);
      case Hittable::Translate: {  // 1029-1037
        Vec3<T> off(h->off);
        Ray<T> m_r{r.o - off, r.d, r.time};
        Hit<T> c;
        if (!hit(h->left.get(), m_r, t_min, t_max, c)) return false;
        out = c; out.p = c.p + off;
        faceNormal(m_r, c.normal, out.frontFace, out.normal);
        return true;
      }
      case Hittable::Rotate: {  // 1038-1052
        T s = (T)h->sinT, co = (T)h->cosT;
        Ray<T> rr{unRotatePoint<T>(h->axis, s, co, r.o), unRotatePoint<T>(h->axis, s, co, r.d), r.time};
        Hit<T> c;
        if (!hit(h->left.get(), rr, t_min, t_max, c)) return false;
        out = c;
        out.p = rotatePoint<T>(h->axis, s, co, c.p);
        Vec3<T> rotN = rotatePoint<T>(h->axis, s, co, c.normal);
        faceNormal(rr, rotN, out.frontFace, out.normal);
        return true;
      }
      case Hittable::ConstantMedium: {  // 1053-1080
        Hit<T> h1, h2;
        if (!hit(h->left.get(), r, -K<T>::infinity(), K<T>::infinity(), h1)) return false;
        if (!hit(h->left.get(), r, h1.t + K<T>::epsilon(), K<T>::infinity(), h2)) return false;
        T rec1t_ = hmax(t_min, h1.t);
        T rec2t = hmin(t_max, h2.t);
        if (rec1t_ >= rec2t) return false;
        T rec1t = (rec1t_ < 0) ? (T)0 : rec1t_;
        T rayLength = length(r.d);
        T distInsideBound = (rec2t - rec1t) * rayLength;
        T rnd = nextReal<T>(g);
        T hitDist = (T)h->nInvD * mlog(rnd);
        if (hitDist > distInsideBound) return false;
        T newt = rec1t + (hitDist / rayLength);
        out.t = newt; out.p = at(r, newt); out.normal = Vec3<T>(1, 0, 0); out.u = 0; out.v = 0; out.frontFace = true;
        out.mat = h->mat.get(); out.prim = refOf(h);
        return true;
      }
      case Hittable::Sphere:  // 1081
        return sphereHit(Vec3<T>(h->c0), (T)h->r, h->mat.get(), refOf(h), r, t_min, t_max, out);
      case Hittable::MovingSphere: {  // 1106-1108
        Vec3<T> c0(h->c0), c1(h->c1);
        Vec3<T> scv = c0 + scale((r.time - (T)h->t0) / (T)h->dur, c1 - c0);
        return sphereHit(scv, (T)h->r, h->mat.get(), refOf(h), r, t_min, t_max, out);
      }
      default: return false;  // Unhittable 1109
    }
  }

  // ---- samplers (Lib.hs:1152-1228)
  T rnd() { return nextReal<T>(g); }
  Vec3<T> randomInUnitSphere() {  // 1160-1168
    for (;;) { T x = rnd(), y = rnd(), z = rnd(); Vec3<T> p = scale((T)2.0, Vec3<T>(x, y, z)) - Vec3<T>(1, 1, 1); if (squaredLength(p) < (T)1.0) return p; }
  }
  Vec3<T> randomInUnitDisk() {  // 1178-1185
    for (;;) { T x = rnd(), y = rnd(); Vec3<T> p = scale((T)2.0, Vec3<T>(x, y, 0)) - Vec3<T>(1, 1, 0); if (squaredLength(p) < (T)1.0) return p; }
  }
  Vec3<T> randomUnitVector() {  // 1187-1197
    T aa = rnd(); T a = aa * 2 * K<T>::pi(); T zz = rnd(); T z = (zz * 2) - 1; T r = std::sqrt(1 - z * z);
    return Vec3<T>(r * mcos(a), r * msin(a), z);
  }
  Vec3<T> randomCosineDirection() {  // 1206-1217
    T r1 = rnd(), r2 = rnd(); T z = std::sqrt(1 - r2); T phi = 2 * K<T>::pi() * r1;
    T x = mcos(phi) * std::sqrt(r2); T y = msin(phi) * std::sqrt(r2);
    return Vec3<T>(x, y, z);
  }
  Vec3<T> randomToSphere(T radius, T distSquared) {  // 1219-1228
    T r1 = rnd(), r2 = rnd();
    T z = 1 + r2 * (std::sqrt(1 - radius * radius / distSquared) - 1);
    T phi = 2 * K<T>::pi() * r1;
    T om = std::sqrt(1 - z * z);
    return Vec3<T>(mcos(phi) * om, msin(phi) * om, z);
  }

  // ---- textures (Lib.hs:441-513)
  T noise(const Texture* tex, const Vec3<T>& p) {  // 441-461 + perlinInterp 466-484
    if (tr) ++tr->noiseCalls;
    const PerlinTable* tab = tex->perlin.get();
    Vec3<T> q = scale((T)tex->scale, p);
    T fi = std::floor(q.x), fj = std::floor(q.y), fk = std::floor(q.z);
    long long i = (long long)fi, j = (long long)fj, k = (long long)fk;
    T u = q.x - fi, v = q.y - fj, w = q.z - fk;
    auto hermite = [](T z) { return z * z * (3 - 2 * z); };
    T uu = hermite(u), vv = hermite(v), ww = hermite(w);
    auto md = [](long long a) { long long m = a % 256; return (int)(m < 0 ? m + 256 : m); };  // Haskell `mod`
    T acc = 0;  // foldr: last element of ds first, i.e. (1,1,1) ... (0,0,0)
    for (int di = 1; di >= 0; --di) for (int dj = 1; dj >= 0; --dj) for (int dk = 1; dk >= 0; --dk) {
      int idx = tab->perm[0][md(i + di)] ^ tab->perm[1][md(j + dj)] ^ tab->perm[2][md(k + dk)];
      Vec3<T> val((T)tab->ranvec[idx][0], (T)tab->ranvec[idx][1], (T)tab->ranvec[idx][2]);
      T I = (T)di, J = (T)dj, Kk = (T)dk;
      acc = acc + ((I * uu + (1 - I) * (1 - uu)) * (J * vv + (1 - J) * (1 - vv)) * (Kk * ww + (1 - Kk) * (1 - ww)) *
                   dot(val, Vec3<T>(u - I, v - J, w - Kk)));
    }
    return acc;
  }
  T turb(const Texture* tex, const Vec3<T>& p, int depth) {  // 486-494
    T acc = 0; Vec3<T> tp = p; T weight = 1;
    for (int d = 0; d < depth; ++d) { acc = acc + weight * noise(tex, tp); tp = scale((T)2.0, tp); weight = weight * (T)0.5; }
    return std::fabs(acc);
  }
  Vec3<T> textureValue(const Texture* tex, T u, T v, const Vec3<T>& p) {  // 496-513
    switch (tex->kind) {
      case Texture::Constant: return Vec3<T>((T)tex->rgb[0], (T)tex->rgb[1], (T)tex->rgb[2]);
      case Texture::Checker:
        if (tr) ++tr->shade[7];
        if (msin(10 * p.x) * msin(10 * p.y) * msin(10 * p.z) < 0) return textureValue(tex->odd.get(), u, v, p);
        return textureValue(tex->even.get(), u, v, p);
      case Texture::Perlin: {
        T m = (T)0.5 * ((T)1.0 + msin(p.z + 10 * turb(tex, p, 7)));  // marbleTexture 512-513
        return scale(m, Vec3<T>(1, 1, 1));
      }
      default: {
        if (!tex->image) return Vec3<T>(0, 1, 1);  // 510
        if (tr) ++tr->texels;
        T nxd = (T)tex->nx, nyd = (T)tex->ny;
        long long i = (long long)std::floor(clampH<T>(0, nxd - K<T>::epsilon(), u * nxd));
        long long j = (long long)std::floor(clampH<T>(0, nyd - K<T>::epsilon(), ((T)1.0 - v) * nyd - K<T>::epsilon()));
        const uint8_t* px = &tex->image->rgb[(size_t)(j * tex->image->w + i) * 3];
        return Vec3<T>((T)px[0] / (T)255.0, (T)px[1] / (T)255.0, (T)px[2] / (T)255.0);  // colorToAlbedo 294-297
      }
    }
  }

  // ---- PDFs over hittables (Lib.hs:662-724)
  T htblPdfValue(const Hittable* h, const Vec3<T>& origin, const Vec3<T>& v) {  // 673-705
    Hit<T> rec;
    if (tr) ++tr->queries;
    if (!hit(h, Ray<T>{origin, v, (T)0.0}, K<T>::epsilon(), K<T>::infinity(), rec)) return 0;
    switch (h->kind) {
      case Hittable::RectXZ: {
        if (tr) ++tr->shade[6];
        T area = ((T)h->a1 - (T)h->a0) * ((T)h->b1 - (T)h->b0);
        T distanceSquared = rec.t * rec.t * squaredLength(v);
        T cosine = std::fabs(dot(v, rec.normal) / length(v));
        return distanceSquared / (cosine * area);
      }
      case Hittable::Sphere: {
        if (tr) ++tr->shade[6];
        Vec3<T> center(h->c0); T radius = (T)h->r;
        T cosThetaMax = std::sqrt(1 - radius * radius / squaredLength(center - origin));
        T solidAngle = 2 * K<T>::pi() * (1 - cosThetaMax);
        return 1 / solidAngle;
      }
      case Hittable::BVHNode: {
        T leftPdf = htblPdfValue(h->left.get(), origin, v) + (T)0;
        T leftWeight = (T)htblSize(h->left.get()) / (T)h->size;
        T rightPdf = htblPdfValue(h->right.get(), origin, v) + (T)0;
        T rightWeight = (T)htblSize(h->right.get()) / (T)h->size;
        return leftWeight * leftPdf + rightWeight * rightPdf;
      }
      default: return 0;
    }
  }
  Vec3<T> htblRandom(const Hittable* h, const Vec3<T>& o) {  // 707-724
    switch (h->kind) {
      case Hittable::RectXZ: {
        T x0 = (T)h->a0, x1 = (T)h->a1, z0 = (T)h->b0, z1 = (T)h->b1;
        T r1 = rnd(); T rx = x0 + (x1 - x0) * r1;
        T r2 = rnd(); T rz = z0 + (z1 - z0) * r2;
        return Vec3<T>(rx, (T)h->k, rz) - o;
      }
      case Hittable::Sphere: {
        Vec3<T> dir = Vec3<T>(h->c0) - o;
        T distSquared = squaredLength(dir);
        ONB<T> uvw = onbFromW(dir);
        Vec3<T> rts = randomToSphere((T)h->r, distSquared);
        return onbLocalV(uvw, rts);
      }
      case Hittable::BVHNode: {
        T rd = rnd();
        if (rd < (T)htblSize(h->left.get()) / (T)h->size) return htblRandom(h->left.get(), o);
        return htblRandom(h->right.get(), o);
      }
      default: return Vec3<T>(1, 0, 0);
    }
  }

  // ---- materials (Lib.hs:819-903)
  static Vec3<T> reflect(const Vec3<T>& v, const Vec3<T>& n) { return v - scale((T)2.0 * dot(v, n), n); }  // 887-888
  static Vec3<T> refract(const Vec3<T>& v, const Vec3<T>& n, T eta) {  // 890-896
    Vec3<T> uv = makeUnitVector(v);
    T cosTheta = dot(vecNegate(uv), n);
    Vec3<T> rOutParallel = scale(eta, uv + scale(cosTheta, n));
    Vec3<T> rOutPerp = scale(-std::sqrt((T)1.0 - squaredLength(rOutParallel)), n);
    return rOutParallel + rOutPerp;
  }
  T schlick(T cosine, T ref_idx) const {  // 899-903
    T r0 = ((T)1.0 - ref_idx) / ((T)1.0 + ref_idx);
    T r1 = r0 * r0;
    if (strict) { T x = (T)1 - cosine; T x2 = x * x; return r1 + ((T)1.0 - r1) * ((x2 * x2) * x); }  // the device's form of `**5`
    return r1 + ((T)1.0 - r1) * std::pow((T)1 - cosine, (T)5);
  }
  struct ScatterRecord { Ray<T> ray; bool specular; Vec3<T> att; T pdf; };  // 819
  bool scatter(const Material* m, const Ray<T>& r, const Hit<T>& h, ScatterRecord& s) {  // 822-865
    switch (m->kind) {
      case Material::Lambertian: {
        if (tr) ++tr->shade[0];
        Vec3<T> att = textureValue(m->tex.get(), h.u, h.v, h.p);
        const Hittable* lights = sc.lights.get();
        ONB<T> uvw = onbFromW(h.normal);
        // pdfGenerate (MixturePdf lightsPdf cosinePdf)  367-372
        T rd = rnd();
        if (tr && rd < (T)0.5) ++tr->shade[5];
        Vec3<T> pdfD = (rd < (T)0.5) ? htblRandom(lights, h.p) : onbLocalV(uvw, randomCosineDirection());
        s.ray = Ray<T>{h.p, makeUnitVector(pdfD), r.time};
        // pdfValue mixPdf (makeUnitVector pdfD)  374-382
        Vec3<T> dir = makeUnitVector(pdfD);
        T v1 = htblPdfValue(lights, h.p, dir);
        T cosine = dot(makeUnitVector(dir), uvw.w);
        T v2 = (cosine <= 0) ? (T)0 : cosine / K<T>::pi();
        s.specular = false; s.att = att; s.pdf = (T)0.5 * (v1 + v2);
        return true;
      }
      case Material::Metal: {
        if (tr) ++tr->shade[1];
        Vec3<T> rUnit = randomUnitVector();
        Vec3<T> reflected = reflect(makeUnitVector(r.d), h.normal);
        s.ray = Ray<T>{h.p, reflected + scale((T)m->param, rUnit), r.time};
        s.specular = true; s.att = textureValue(m->tex.get(), h.u, h.v, h.p); s.pdf = 0;
        return true;
      }
      case Material::Dielectric: {
        if (tr) ++tr->shade[2];
        T ref_idx = (T)m->param;
        T etaiOverEtat = h.frontFace ? (T)1.0 / ref_idx : ref_idx;
        Vec3<T> unitDirection = makeUnitVector(r.d);
        T cosTheta = hmin(dot(vecNegate(unitDirection), h.normal), (T)1.0);
        T sinTheta = std::sqrt((T)1.0 - cosTheta * cosTheta);
        T rd = rnd();
        s.specular = true; s.att = Vec3<T>(1, 1, 1); s.pdf = 1;
        if ((etaiOverEtat * sinTheta > (T)1.0) || rd < schlick(cosTheta, etaiOverEtat)) s.ray = Ray<T>{h.p, reflect(unitDirection, h.normal), r.time};
        else s.ray = Ray<T>{h.p, refract(unitDirection, h.normal, etaiOverEtat), r.time};
        return true;
      }
      case Material::DiffuseLight: if (tr) ++tr->shade[4]; return false;
      default: {  // Isotropic 861-865
        if (tr) ++tr->shade[3];
        Vec3<T> randDr = randomInUnitSphere();
        s.ray = Ray<T>{h.p, randDr, r.time};
        s.att = textureValue(m->tex.get(), h.u, h.v, h.p);
        s.specular = false; s.pdf = 1;
        return true;
      }
    }
  }
  static T scatteringPdf(const Hit<T>& h, const Ray<T>& scattered) {  // 867-878
    T cosine = dot(h.normal, scattered.d);
    return (cosine < 0) ? (T)0 : cosine / K<T>::pi();
  }
  Vec3<T> emitted(const Material* m, const Hit<T>& h) {  // 880-885
    if (m->kind == Material::DiffuseLight && !h.frontFace) return textureValue(m->tex.get(), h.u, h.v, h.p);
    return Vec3<T>(0, 0, 0);
  }

  // ---- rayColor (Lib.hs:1298-1333): plain recursion == the CPS closure chain, same draw and multiply order
  Vec3<T> rayColor(const Ray<T>& r, int d) {
    if (betaOrder) return rayColorBeta(r, d);
    if (d <= 0) return Vec3<T>(0, 0, 0);
    Hit<T> h;
    if (tr) { ++tr->queries; ++tr->bounces; }
    if (!hit(sc.world.get(), r, K<T>::epsilon(), K<T>::infinity(), h)) {
      if (tr) tr->bounce(0xffffffffu, 0.0);
      return Vec3<T>((T)sc.background[0], (T)sc.background[1], (T)sc.background[2]);
    }
    if (tr) tr->bounce(h.prim, (double)h.t);
    ScatterRecord s;
    if (!scatter(h.mat, r, h, s)) return emitted(h.mat, h);
    if (s.specular) return s.att * rayColor(s.ray, d - 1);
    T spdf = scatteringPdf(h, s.ray);
    Vec3<T> nw = rayColor(s.ray, d - 1);
    return s.att * scale(spdf, divide(nw, s.pdf));
  }

  // The same path with the throughput accumulated FORWARD, beta <- beta (x) att [* spdf / pdf], L = beta (x) terminal: the order the
  // device multiplies in (it cannot hold 50 pending frames per lane).  Identical draws, hits and decisions; only the rounding of the
  // final product differs from rayColor (tests/test_oracle.py bounds the difference).  Not in the reference.
  Vec3<T> rayColorBeta(Ray<T> r, int d) {
    Vec3<T> beta(1, 1, 1);
    for (;;) {
      if (d <= 0) return beta * Vec3<T>(0, 0, 0);
      Hit<T> h;
      if (tr) { ++tr->queries; ++tr->bounces; }
      if (!hit(sc.world.get(), r, K<T>::epsilon(), K<T>::infinity(), h)) {
        if (tr) tr->bounce(0xffffffffu, 0.0);
        return beta * Vec3<T>((T)sc.background[0], (T)sc.background[1], (T)sc.background[2]);
      }
      if (tr) tr->bounce(h.prim, (double)h.t);
      ScatterRecord s;
      if (!scatter(h.mat, r, h, s)) return beta * emitted(h.mat, h);
      if (s.specular) { if (h.mat->kind != Material::Dielectric) beta = beta * s.att; }  // Dielectric: att = (1,1,1)
      else { T spdf = scatteringPdf(h, s.ray); beta = scale(spdf / s.pdf, beta * s.att); }
      r = s.ray; --d;
    }
  }

  // ---- getRay (Lib.hs:1253-1267)
  Ray<T> getRay(const Camera& c, T s, T t) {
    Vec3<T> rd = scale((T)c.lensRadius, randomInUnitDisk());
    Vec3<T> offset = scale(rd.x, Vec3<T>(c.u)) + scale(rd.y, Vec3<T>(c.v));
    T r = rnd(); T tm = (T)c.t0 + ((T)c.t1 - (T)c.t0) * r;
    Vec3<T> org(c.origin);
    return Ray<T>{org + offset, Vec3<T>(c.llc) + scale(s, Vec3<T>(c.horiz)) + scale(t, Vec3<T>(c.vert)) - org - offset, tm};
  }
};

// =====================================================================================================
// Render drivers
// =====================================================================================================
struct RenderOut { std::vector<double> sum; Trace stats; };  // sum: [H][W][3], row 0 = top (y = H-1), radiance SUMS in double

// Reference mode: runRender (Lib.hs:1491-1523) + uniformRandomUVs (1352-1371) + renderPos/sampleColor (1335-1350).
// gens = one generator per column, carried row to row; column 0 = scene.g1 (Main.hs:48-49), the others stand in
// for Main's entropy-seeded newRandGen and are derived from `entropy`.
template <class T, class S>
inline RenderOut renderReference(const Scene& sc, const Camera& cam, int W, int H, int ns, int maxDepth, uint64_t entropy, int nThreads, bool stats) {
  RenderOut out; out.sum.assign((size_t)W * H * 3, 0.0);
  std::vector<SMGen> gens(W);
  gens[0] = sc.g1;
  for (int x = 1; x < W; ++x) gens[x] = mkSMGen(entropy + (uint64_t)x);
  std::vector<Trace> traces(std::max(1, nThreads));
  std::atomic<int> nextCol{0};
  auto work = [&](int tid) {
    Trace* tr = stats ? &traces[tid] : nullptr;
    std::vector<T> us(ns), vs(ns);
    for (;;) {
      int x = nextCol.fetch_add(1); if (x >= W) break;
      SMGen g = gens[x];
      for (int row = 0; row < H; ++row) {
        int y = H - 1 - row;
        // uniformRandomUVs: generated up front; foldr conses, so the list head is the LAST pair generated
        for (int k = 0; k < ns; ++k) { T ru = nextReal<T>(g); T rv = nextReal<T>(g); us[k] = ((T)x + ru) / (T)W; vs[k] = ((T)y + rv) / (T)H; }
        Tracer<T, S> t(sc, g, tr);
        Vec3<T> acc(0, 0, 0);
        for (int k = ns - 1; k >= 0; --k) {  // foldM sampleColor over the list, head first
          Ray<T> r = t.getRay(cam, us[k], vs[k]);
          if (tr) ++tr->paths;
          acc = acc + t.rayColor(r, maxDepth);
        }
        double* px = &out.sum[((size_t)row * W + x) * 3];
        px[0] = (double)acc.x; px[1] = (double)acc.y; px[2] = (double)acc.z;
      }
    }
  };
  std::vector<std::thread> th;
  for (int i = 1; i < nThreads; ++i) th.emplace_back(work, i);
  work(0);
  for (auto& t : th) t.join();
  for (auto& t : traces) out.stats.add(t);
  return out;
}

// Device-stream mode: same per-sample algorithm, but sample s of pixel (x,y) owns streamGen(seed, pixel, s)
// and draws (ru, rv) first; per-path radiance is accumulated in double (DESIGN.md "accumulation").
// region: rows [row0,row1) x cols [col0,col1) of the full WxH image; samples [s0, s0+ns)
template <class T, class S>
inline RenderOut renderStreams(const Scene& sc, const Camera& cam, int W, int H, int s0, int ns, int maxDepth, uint64_t seed,
                               int row0, int row1, int col0, int col1, int nThreads, bool stats, bool strict = false, bool betaOrder = false,
                               uint32_t* pathPrim = nullptr, double* pathT = nullptr, uint32_t pathStride = 0) {
  RenderOut out; out.sum.assign((size_t)W * H * 3, 0.0);
  std::vector<Trace> traces(std::max(1, nThreads));
  std::atomic<int> nextRow{row0};
  auto work = [&](int tid) {
    Trace* tr = (stats || pathPrim) ? &traces[tid] : nullptr;
    for (;;) {
      int row = nextRow.fetch_add(1); if (row >= row1) break;
      int y = H - 1 - row;
      for (int x = col0; x < col1; ++x) {
        uint32_t pixel = (uint32_t)(row * W + x);
        double acc[3] = {0, 0, 0};
        for (int s = s0; s < s0 + ns; ++s) {
          SMGen g = streamGen(seed, pixel, (uint32_t)s);
          Tracer<T, S> t(sc, g, tr, strict, betaOrder);
          if (pathPrim) {  // path index: region pixels row-major, samples innermost (the device's rt_render_pathlog order)
            const size_t path = ((size_t)(row - row0) * (col1 - col0) + (x - col0)) * ns + (s - s0);
            tr->pathPrim = pathPrim + path * pathStride; tr->pathT = pathT + path * pathStride; tr->pathStride = pathStride; tr->pathBounce = 0;
          }
          T ru = nextReal<T>(g); T rv = nextReal<T>(g);
          T u = ((T)x + ru) / (T)W, v = ((T)y + rv) / (T)H;
          Ray<T> r = t.getRay(cam, u, v);
          if (tr) ++tr->paths;
          Vec3<T> c = t.rayColor(r, maxDepth);
          acc[0] += (double)c.x; acc[1] += (double)c.y; acc[2] += (double)c.z;
        }
        double* px = &out.sum[((size_t)row * W + x) * 3];
        px[0] = acc[0]; px[1] = acc[1]; px[2] = acc[2];
      }
    }
  };
  std::vector<std::thread> th;
  for (int i = 1; i < nThreads; ++i) th.emplace_back(work, i);
  work(0);
  for (auto& t : th) t.join();
  for (auto& t : traces) out.stats.add(t);
  return out;
}

// albedoToColor / scaleColor (Lib.hs:287-292)
inline uint8_t scaleColor(double x) {
  double c = clampH<double>(0.0, 0.999, std::sqrt(x));
  double f = std::floor(256 * c);
  if (!(f == f)) return 0;  // floor NaN -> Word8 is GHC/arch-defined; 0 chosen and documented
  return (uint8_t)(long long)f;
}

}  // namespace orc
