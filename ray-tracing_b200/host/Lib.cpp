// host/Lib.cpp — see Lib.hpp.  Reference lines cited are src/Lib.hs unless noted.
#include "Lib.hpp"
#include <string>

#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <stdexcept>
#include <unordered_map>

namespace Lib {

Vec3 operator+(const Vec3& a, const Vec3& b) { return Vec3{a.x + b.x, a.y + b.y, a.z + b.z}; }
Vec3 operator-(const Vec3& a, const Vec3& b) { return Vec3{a.x - b.x, a.y - b.y, a.z - b.z}; }
Vec3 operator*(const Vec3& a, const Vec3& b) { return Vec3{a.x * b.x, a.y * b.y, a.z * b.z}; }
double length(const Vec3& v) { return std::sqrt((v.x * v.x) + (v.y * v.y) + (v.z * v.z)); }
static Vec3 scaleV(double k, const Vec3& v) { return Vec3{v.x * k, v.y * k, v.z * k}; }
static Vec3 unitV(const Vec3& v) { double l = length(v); return Vec3{v.x / l, v.y / l, v.z / l}; }
static Vec3 crossV(const Vec3& a, const Vec3& b) { return Vec3{a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
// Prelude min/max
static double hmin(double x, double y) { return x <= y ? x : y; }
static double hmax(double x, double y) { return x <= y ? y : x; }

// ---- splitmix-0.1 / random-1.2.0 (pinned by the reference's stack.yaml:42-44)
static uint64_t mixA(uint64_t z) { z = (z ^ (z >> 33)) * 0xff51afd7ed558ccdULL; z = (z ^ (z >> 33)) * 0xc4ceb9fe1a85ec53ULL; return z ^ (z >> 33); }
static uint64_t mixB(uint64_t z) { z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ULL; z = (z ^ (z >> 27)) * 0x94d049bb133111ebULL; return z ^ (z >> 31); }
RandGen randGen(int64_t s) {
  uint64_t u = (uint64_t)s;
  uint64_t g = mixB(u + 0x9e3779b97f4a7c15ULL) | 1ULL;
  if (__builtin_popcountll(g ^ (g >> 1)) < 24) g ^= 0xaaaaaaaaaaaaaaaaULL;
  return RandGen{mixA(u), g};
}
uint64_t genWord64(RandGen& g) { g.seed += g.gamma; return mixA(g.seed); }
double randomDouble(RandGen& g) {
  uint64_t bits = (genWord64(g) >> 12) | 0x3ff0000000000000ULL;
  double d; std::memcpy(&d, &bits, 8);
  return d - 1.0;
}
double randomDoubleM(RandGen& g) { return randomDouble(g); }
double randomDoubleRM(RandGen& g, double mn, double mx) { double rd = randomDoubleM(g); return mn + (mx - mn) * rd; }
Vec3 randomVec3DoubleM(RandGen& g) { double x = randomDouble(g), y = randomDouble(g), z = randomDouble(g); return Vec3{x, y, z}; }
Vec3 randomVec3DoubleRM(RandGen& g, double mn, double mx) {
  double x = randomDoubleRM(g, mn, mx), y = randomDoubleRM(g, mn, mx), z = randomDoubleRM(g, mn, mx);
  return Vec3{x, y, z};
}

Camera newCamera(Point3 lookfrom, Point3 lookat, Vec3 vup, double vfov, double aspect, double aperture, double focusDist, Time t0, Time t1) {
  Camera c;
  c.lensRadius = aperture / 2.0;
  const double theta = vfov * 3.14159265358979323846 / 180.0;
  const double halfHeight = std::tan(theta / 2.0), halfWidth = aspect * halfHeight;
  c.origin = lookfrom;
  c.w = unitV(lookfrom - lookat);
  c.u = unitV(crossV(vup, c.w));
  c.v = crossV(c.w, c.u);
  c.llc = c.origin - scaleV(halfWidth * focusDist, c.u) - scaleV(halfHeight * focusDist, c.v) - scaleV(focusDist, c.w);
  c.horiz = scaleV(2 * halfWidth * focusDist, c.u);
  c.vert = scaleV(2 * halfHeight * focusDist, c.v);
  c.t0 = t0; c.t1 = t1;
  return c;
}

// ---- Arena: textures / materials
Texture Arena::ConstantColor(Albedo a) { texs_.push_back(Tex{RT_TEX_CONSTANT, {a.v.x, a.v.y, a.v.z}, RT_NONE, RT_NONE, 1.0}); return Texture{(uint32_t)texs_.size() - 1}; }
Texture Arena::CheckerTexture(Texture odd, Texture even) { texs_.push_back(Tex{RT_TEX_CHECKER, {0, 0, 0}, odd.id, even.id, 1.0}); return Texture{(uint32_t)texs_.size() - 1}; }
Texture Arena::ImageTexture(const Image* im) {
  uint32_t id = RT_NONE;
  if (im && im->w && im->h) {
    auto it = std::find(images_.begin(), images_.end(), im);
    if (it == images_.end()) { images_.push_back(im); id = (uint32_t)images_.size() - 1; } else id = (uint32_t)(it - images_.begin());
  }
  texs_.push_back(Tex{RT_TEX_IMAGE, {0, 0, 0}, id, RT_NONE, 1.0});
  return Texture{(uint32_t)texs_.size() - 1};
}
Texture Arena::makePerlin(RandGen& g, double sc) {
  Perlin p; p.ranvec.resize(768); p.perm.resize(768);
  for (int i = 0; i < 256; ++i) { Vec3 v = randomVec3DoubleRM(g, -1.0, 1.0); p.ranvec[i * 3] = v.x; p.ranvec[i * 3 + 1] = v.y; p.ranvec[i * 3 + 2] = v.z; }
  for (int a = 0; a < 3; ++a) {  // perlinGeneratePerm (430-439): swap p[i] with p[target], target = floor(U[0,i))
    int32_t* q = &p.perm[a * 256];
    for (int i = 0; i < 256; ++i) q[i] = i;
    for (int i = 255; i >= 1; --i) { int target = (int)std::floor(randomDoubleRM(g, 0.0, (double)i)); std::swap(q[i], q[target]); }
  }
  perlins_.push_back(std::move(p));
  texs_.push_back(Tex{RT_TEX_PERLIN, {0, 0, 0}, (uint32_t)perlins_.size() - 1, RT_NONE, sc});
  return Texture{(uint32_t)texs_.size() - 1};
}
Material Arena::addMat(uint32_t kind, uint32_t tex, double param) { mats_.push_back(Mat{kind, tex, param}); return Material{(uint32_t)mats_.size() - 1}; }
Material Arena::Lambertian(Texture t) { return addMat(RT_LAMBERTIAN, t.id, 0); }
Material Arena::Metal(Texture t, Fuzz f) { return addMat(RT_METAL, t.id, f.v); }
Material Arena::Dielectric(RefractiveIdx r) { return addMat(RT_DIELECTRIC, RT_NONE, r.v); }
Material Arena::DiffuseLight(Texture t) { return addMat(RT_DIFFUSE_LIGHT, t.id, 0); }
Material Arena::Isotropic(Texture t) { return addMat(RT_ISOTROPIC, t.id, 0); }

// ---- Arena: hittables
Hittable Arena::sphere(Point3 c, double r, Material m) { spheres_.push_back(SphereO{c, r, m.id}); return Hittable{RT_REF(RT_SPHERE, spheres_.size() - 1)}; }
Hittable Arena::movingSphere(Point3 c0, Point3 c1, Time t0, Time t1, double r, Material m) {
  moving_.push_back(MovingO{c0, c1, t0, t1, t1 - t0, r, m.id});
  return Hittable{RT_REF(RT_MOVING_SPHERE, moving_.size() - 1)};
}
Hittable Arena::rect(Plane p, double a0, double a1, double b0, double b1, double k, Material m) {
  rects_.push_back(RectO{a0, a1, b0, b1, k, m.id});
  return Hittable{RT_REF(RT_RECT_XY + (int)p, rects_.size() - 1)};
}
Hittable Arena::cuboid(Point3 mn, Point3 mx, Material m) { cuboids_.push_back(CuboidO{mn, mx, m.id}); return Hittable{RT_REF(RT_CUBOID, cuboids_.size() - 1)}; }
Hittable Arena::translate(Point3 off, Hittable h) { translates_.push_back(TranslateO{off, h.ref}); return Hittable{RT_REF(RT_TRANSLATE, translates_.size() - 1)}; }

static Vec3 rotatePointD(int axis, double s, double c, const Vec3& p) {  // 763-774
  if (axis == XAxis) return Vec3{p.x, c * p.y - s * p.z, s * p.y + c * p.z};
  if (axis == YAxis) return Vec3{c * p.x + s * p.z, p.y, -s * p.x + c * p.z};
  return Vec3{c * p.x - s * p.y, s * p.x + c * p.y, p.z};
}
Hittable Arena::rotate(Axis axis, double angleDeg, Hittable h) {  // 732-761 incl. the i,j,k in {0,1,2} corner sweep
  RotateO r; r.axis = axis; r.child = h.ref;
  const double rad = angleDeg * 3.14159265358979323846 / 180.0;
  r.s = std::sin(rad); r.c = std::cos(rad);
  const Box bb = boundingBox(h);
  const double inf = std::numeric_limits<double>::infinity();
  Vec3 mn{inf, inf, inf}, mx{-inf, -inf, -inf};
  for (int i = 2; i >= 0; --i) for (int j = 2; j >= 0; --j) for (int k = 2; k >= 0; --k) {
    const Vec3 corner{i * bb.box_max.x + (1 - i) * bb.box_min.x, j * bb.box_max.y + (1 - j) * bb.box_min.y, k * bb.box_max.z + (1 - k) * bb.box_min.z};
    const Vec3 n = rotatePointD(axis, r.s, r.c, corner);
    mn = Vec3{hmin(n.x, mn.x), hmin(n.y, mn.y), hmin(n.z, mn.z)};
    mx = Vec3{hmax(n.x, mx.x), hmax(n.y, mx.y), hmax(n.z, mx.z)};
  }
  r.box = Box{mn, mx};
  rotates_.push_back(r);
  return Hittable{RT_REF(RT_ROTATE, rotates_.size() - 1)};
}
Hittable Arena::constantMedium(double density, Texture tex, Hittable boundary) {  // 789-791
  media_.push_back(MediumO{-1 / density, Isotropic(tex).id, boundary.ref});
  return Hittable{RT_REF(RT_CONSTANT_MEDIUM, media_.size() - 1)};
}

static Box surroundingBox(const Box& a, const Box& b) {  // 929-939
  return Box{Vec3{hmin(a.box_min.x, b.box_min.x), hmin(a.box_min.y, b.box_min.y), hmin(a.box_min.z, b.box_min.z)},
             Vec3{hmax(a.box_max.x, b.box_max.x), hmax(a.box_max.y, b.box_max.y), hmax(a.box_max.z, b.box_max.z)}};
}
Box Arena::boundingBox(Hittable h) const {
  const uint32_t k = RT_REF_KIND(h.ref), i = RT_REF_INDEX(h.ref);
  const double eps = 0.0001;
  switch (k) {
    case RT_SPHERE: { const SphereO& s = spheres_[i]; Vec3 r{s.r, s.r, s.r}; return Box{s.c - r, s.c + r}; }
    case RT_MOVING_SPHERE: { const MovingO& s = moving_[i]; Vec3 r{s.r, s.r, s.r}; return surroundingBox(Box{s.c0 - r, s.c0 + r}, Box{s.c1 - r, s.c1 + r}); }
    case RT_RECT_XY: { const RectO& r = rects_[i]; return Box{Vec3{r.a0, r.b0, r.k - eps}, Vec3{r.a1, r.b1, r.k + eps}}; }
    case RT_RECT_XZ: { const RectO& r = rects_[i]; return Box{Vec3{r.a0, r.k - eps, r.b0}, Vec3{r.a1, r.k + eps, r.b1}}; }
    case RT_RECT_YZ: { const RectO& r = rects_[i]; return Box{Vec3{r.k - eps, r.a0, r.b0}, Vec3{r.k + eps, r.a1, r.b1}}; }
    case RT_CUBOID: return Box{cuboids_[i].mn, cuboids_[i].mx};
    case RT_BVH_NODE: return nodes_[i].box;
    case RT_TRANSLATE: { Box b = boundingBox(Hittable{translates_[i].child}); return Box{b.box_min + translates_[i].off, b.box_max + translates_[i].off}; }
    case RT_ROTATE: return rotates_[i].box;
    case RT_CONSTANT_MEDIUM: return boundingBox(Hittable{media_[i].boundary});
    default: throw std::runtime_error("Should not be trying to bound an Unhittable");  // 927
  }
}

size_t Arena::deviceBVHMin = 0;
double Arena::lastDeviceBVHms = 0.0;

Hittable Arena::makeBVH(RandGen& g, const std::vector<Hittable>& htbls) {
  if (deviceBVHMin && htbls.size() >= deviceBVHMin) {
    // The device build needs the leaf boxes and every node's axis draw.  The recursion draws one axis per node, a node before
    // its left subtree before its right subtree, so the pre-order axis array is the next N(n) draws of the generator.
    const size_t n = htbls.size();
    const uint32_t N = rt_bvh_node_count((uint32_t)n);
    std::vector<double> boxes(n * 6);
    for (size_t i = 0; i < n; ++i) {
      const Box b = boundingBox(htbls[i]);
      double* o = &boxes[i * 6];
      o[0] = b.box_min.x; o[1] = b.box_min.y; o[2] = b.box_min.z; o[3] = b.box_max.x; o[4] = b.box_max.y; o[5] = b.box_max.z;
    }
    std::vector<uint8_t> axes(N);
    for (uint32_t k = 0; k < N; ++k) axes[k] = (uint8_t)(int)std::floor(randomDoubleRM(g, 0, 3));
    std::vector<double> nbox((size_t)N * 6); std::vector<uint32_t> nl(N), nr(N), ns(N);
    if (rt_bvh_build((uint32_t)n, boxes.data(), axes.data(), N, nbox.data(), nl.data(), nr.data(), ns.data(), -1, &lastDeviceBVHms) != RT_OK)
      throw std::runtime_error(std::string("rt_bvh_build: ") + rt_last_error());
    const uint32_t base = (uint32_t)nodes_.size();
    nodes_.resize((size_t)base + N);
    auto child = [&](uint32_t c) { return (c & 0x80000000u) ? (uint32_t)RT_REF(RT_BVH_NODE, base + (c & 0x7fffffffu)) : htbls[c].ref; };
    for (uint32_t k = 0; k < N; ++k) {
      const double* b = &nbox[(size_t)k * 6];
      nodes_[base + k] = NodeO{child(nl[k]), child(nr[k]), Box{Vec3{b[0], b[1], b[2]}, Vec3{b[3], b[4], b[5]}}, ns[k]};
    }
    return Hittable{RT_REF(RT_BVH_NODE, base)};
  }
  const int axis = (int)std::floor(randomDoubleRM(g, 0, 3));
  const size_t n = htbls.size();
  if (n == 0) throw std::runtime_error("makeBVH: empty sequence");
  auto keyOf = [&](Hittable h) { Box b = boundingBox(h); return axis == 0 ? b.box_min.x : (axis == 1 ? b.box_min.y : b.box_min.z); };
  uint32_t lt, rt;
  if (n == 1) { lt = rt = htbls[0].ref; }
  else if (n == 2) {
    if (keyOf(htbls[0]) < keyOf(htbls[1])) { lt = htbls[0].ref; rt = htbls[1].ref; } else { lt = htbls[1].ref; rt = htbls[0].ref; }
  } else {
    std::vector<double> keys(n); std::vector<uint32_t> order(n);
    for (size_t i = 0; i < n; ++i) { keys[i] = keyOf(htbls[i]); order[i] = (uint32_t)i; }
    std::stable_sort(order.begin(), order.end(), [&](uint32_t a, uint32_t b) { return keys[a] < keys[b]; });
    const size_t half = n / 2;
    std::vector<Hittable> l(half), r(n - half);
    for (size_t i = 0; i < half; ++i) l[i] = htbls[order[i]];
    for (size_t i = half; i < n; ++i) r[i - half] = htbls[order[i]];
    lt = makeBVH(g, l).ref;
    rt = makeBVH(g, r).ref;
  }
  nodes_.push_back(NodeO{lt, rt, surroundingBox(boundingBox(Hittable{lt}), boundingBox(Hittable{rt})), (uint32_t)n});
  return Hittable{RT_REF(RT_BVH_NODE, nodes_.size() - 1)};
}

// ---- flatten: depth-first from world then lights, left before right, each object numbered at first encounter
struct FlattenCtx {
  const Arena& A; Arena::Flat& F;
  std::unordered_map<uint32_t, uint32_t> refMap, matMap, texMap, perlinMap, imageMap;
  uint32_t nRect = 0;
  FlattenCtx(const Arena& a, Arena::Flat& f) : A(a), F(f) {}
  static void push3(std::vector<double>& v, const Vec3& p) { v.push_back(p.x); v.push_back(p.y); v.push_back(p.z); }
  uint32_t tex(uint32_t id) {
    if (id == RT_NONE) return RT_NONE;
    auto it = texMap.find(id);
    if (it != texMap.end()) return it->second;
    const auto& t = A.texs_[id];
    const uint32_t n = (uint32_t)F.tex_kind.size();
    texMap[id] = n;
    F.tex_kind.push_back(t.kind); F.tex_rgb.push_back(t.rgb[0]); F.tex_rgb.push_back(t.rgb[1]); F.tex_rgb.push_back(t.rgb[2]);
    F.tex_scale.push_back(t.scale); F.tex_a.push_back(RT_NONE); F.tex_b.push_back(RT_NONE);
    if (t.kind == RT_TEX_CHECKER) { uint32_t o = tex(t.a); uint32_t e = tex(t.b); F.tex_a[n] = o; F.tex_b[n] = e; }
    else if (t.kind == RT_TEX_PERLIN) {
      auto pt = perlinMap.find(t.a);
      if (pt == perlinMap.end()) {
        const uint32_t pn = (uint32_t)perlinMap.size(); perlinMap[t.a] = pn;
        F.perlin_ranvec.insert(F.perlin_ranvec.end(), A.perlins_[t.a].ranvec.begin(), A.perlins_[t.a].ranvec.end());
        F.perlin_perm.insert(F.perlin_perm.end(), A.perlins_[t.a].perm.begin(), A.perlins_[t.a].perm.end());
        F.tex_a[n] = pn;
      } else F.tex_a[n] = pt->second;
    } else if (t.kind == RT_TEX_IMAGE && t.a != RT_NONE) {
      auto im = imageMap.find(t.a);
      if (im == imageMap.end()) {
        const uint32_t in = (uint32_t)imageMap.size(); imageMap[t.a] = in;
        F.image_rgb.push_back(A.images_[t.a]->rgb.data()); F.image_w.push_back(A.images_[t.a]->w); F.image_h.push_back(A.images_[t.a]->h);
        F.tex_a[n] = in;
      } else F.tex_a[n] = im->second;
    }
    return n;
  }
  uint32_t mat(uint32_t id) {
    auto it = matMap.find(id);
    if (it != matMap.end()) return it->second;
    const auto& m = A.mats_[id];
    const uint32_t n = (uint32_t)F.mat_kind.size();
    matMap[id] = n;
    F.mat_kind.push_back(m.kind); F.mat_param.push_back(m.param); F.mat_tex.push_back(RT_NONE);
    F.mat_tex[n] = tex(m.tex);
    return n;
  }
  uint32_t go(uint32_t ref) {
    const uint32_t k = RT_REF_KIND(ref), i = RT_REF_INDEX(ref);
    if (k == RT_UNHITTABLE) return RT_REF(RT_UNHITTABLE, 0);
    auto it = refMap.find(ref);
    if (it != refMap.end()) return it->second;
    uint32_t out = 0;
    switch (k) {
      case RT_SPHERE: {
        const auto& s = A.spheres_[i]; out = RT_REF(k, F.sphere_mat.size()); refMap[ref] = out;
        push3(F.sphere_cr, s.c); F.sphere_cr.push_back(s.r); F.sphere_mat.push_back(mat(s.mat)); break;
      }
      case RT_MOVING_SPHERE: {
        const auto& s = A.moving_[i]; out = RT_REF(k, F.moving_mat.size()); refMap[ref] = out;
        push3(F.moving, s.c0); push3(F.moving, s.c1); F.moving.push_back(s.t0); F.moving.push_back(s.dur); F.moving.push_back(s.r);
        F.moving_mat.push_back(mat(s.mat)); break;
      }
      case RT_RECT_XY: case RT_RECT_XZ: case RT_RECT_YZ: {
        const auto& r = A.rects_[i]; out = RT_REF(k, F.rect_mat.size()); refMap[ref] = out;
        F.rect.push_back(r.a0); F.rect.push_back(r.a1); F.rect.push_back(r.b0); F.rect.push_back(r.b1); F.rect.push_back(r.k);
        F.rect_mat.push_back(mat(r.mat)); break;
      }
      case RT_CUBOID: {
        const auto& c = A.cuboids_[i]; out = RT_REF(k, F.cuboid_mat.size()); refMap[ref] = out;
        push3(F.cuboid, c.mn); push3(F.cuboid, c.mx); F.cuboid_mat.push_back(mat(c.mat)); break;
      }
      case RT_BVH_NODE: {
        const auto& n = A.nodes_[i]; const size_t me = F.node_size.size(); out = RT_REF(k, me); refMap[ref] = out;
        push3(F.node_box, n.box.box_min); push3(F.node_box, n.box.box_max); F.node_size.push_back(n.size); F.node_left.push_back(0); F.node_right.push_back(0);
        const uint32_t l = go(n.left); const uint32_t r = go(n.right);
        F.node_left[me] = l; F.node_right[me] = r; break;
      }
      case RT_TRANSLATE: {
        const auto& t = A.translates_[i]; const size_t me = F.translate_child.size(); out = RT_REF(k, me); refMap[ref] = out;
        push3(F.translate_off, t.off); F.translate_child.push_back(0);
        const uint32_t c = go(t.child); F.translate_child[me] = c; break;
      }
      case RT_ROTATE: {
        const auto& r = A.rotates_[i]; const size_t me = F.rotate_child.size(); out = RT_REF(k, me); refMap[ref] = out;
        F.rotate_sc.push_back(r.s); F.rotate_sc.push_back(r.c); F.rotate_axis.push_back(r.axis); F.rotate_child.push_back(0);
        const uint32_t c = go(r.child); F.rotate_child[me] = c; break;
      }
      case RT_CONSTANT_MEDIUM: {
        const auto& m = A.media_[i]; const size_t me = F.medium_mat.size(); out = RT_REF(k, me); refMap[ref] = out;
        F.medium_nid.push_back(m.nInvD); F.medium_mat.push_back(0); F.medium_boundary.push_back(0);
        const uint32_t mm = mat(m.mat); F.medium_mat[me] = mm;
        const uint32_t b = go(m.boundary); F.medium_boundary[me] = b; break;
      }
      default: throw std::runtime_error("flatten: bad hittable kind");
    }
    return out;
  }
};

void Arena::flatten(Hittable world, Hittable lights, Albedo background, const Camera& cam, Flat& F) const {
  F = Flat{};
  FlattenCtx cx(*this, F);
  const uint32_t w = cx.go(world.ref);
  const uint32_t l = cx.go(lights.ref);
  RtSceneDesc& d = F.desc;
  d.abi_version = RT_ABI_VERSION;
  d.n_nodes = (uint32_t)F.node_size.size(); d.node_box = F.node_box.data(); d.node_left = F.node_left.data(); d.node_right = F.node_right.data(); d.node_size = F.node_size.data();
  d.n_spheres = (uint32_t)F.sphere_mat.size(); d.sphere_cr = F.sphere_cr.data(); d.sphere_mat = F.sphere_mat.data();
  d.n_moving = (uint32_t)F.moving_mat.size(); d.moving = F.moving.data(); d.moving_mat = F.moving_mat.data();
  d.n_rects = (uint32_t)F.rect_mat.size(); d.rect = F.rect.data(); d.rect_mat = F.rect_mat.data();
  d.n_cuboids = (uint32_t)F.cuboid_mat.size(); d.cuboid = F.cuboid.data(); d.cuboid_mat = F.cuboid_mat.data();
  d.n_translates = (uint32_t)F.translate_child.size(); d.translate_off = F.translate_off.data(); d.translate_child = F.translate_child.data();
  d.n_rotates = (uint32_t)F.rotate_child.size(); d.rotate_sc = F.rotate_sc.data(); d.rotate_axis = F.rotate_axis.data(); d.rotate_child = F.rotate_child.data();
  d.n_media = (uint32_t)F.medium_mat.size(); d.medium_neg_inv_density = F.medium_nid.data(); d.medium_mat = F.medium_mat.data(); d.medium_boundary = F.medium_boundary.data();
  d.n_materials = (uint32_t)F.mat_kind.size(); d.mat_kind = F.mat_kind.data(); d.mat_tex = F.mat_tex.data(); d.mat_param = F.mat_param.data();
  d.n_textures = (uint32_t)F.tex_kind.size(); d.tex_kind = F.tex_kind.data(); d.tex_rgb = F.tex_rgb.data(); d.tex_a = F.tex_a.data(); d.tex_b = F.tex_b.data(); d.tex_scale = F.tex_scale.data();
  d.n_perlin = (uint32_t)(F.perlin_perm.size() / 768); d.perlin_ranvec = F.perlin_ranvec.data(); d.perlin_perm = F.perlin_perm.data();
  d.n_images = (uint32_t)F.image_rgb.size(); d.image_rgb = F.image_rgb.data(); d.image_w = F.image_w.data(); d.image_h = F.image_h.data();
  d.world_root = w; d.lights_root = l;
  d.background[0] = background.v.x; d.background[1] = background.v.y; d.background[2] = background.v.z;
  const Vec3* vs[7] = {&cam.origin, &cam.llc, &cam.horiz, &cam.vert, &cam.u, &cam.v, &cam.w};
  for (int i = 0; i < 7; ++i) { d.camera[i * 3] = vs[i]->x; d.camera[i * 3 + 1] = vs[i]->y; d.camera[i * 3 + 2] = vs[i]->z; }
  d.camera[21] = cam.lensRadius; d.camera[22] = cam.t0; d.camera[23] = cam.t1;
}

// ---- render entry (Lib.hs:92-108, 287-292, 1491-1523)
RenderStaticEnv mkRenderStaticEnv(const Arena& arena, Scene s, Camera c, std::pair<int, int> wh, int numSamples, int maxDepth, int numThreads) {
  return RenderStaticEnv{&arena, s, c, wh.first, wh.second, numSamples, maxDepth, numThreads};
}
static uint8_t scaleColor(double x) {
  const double s = std::sqrt(x);
  const double c = (s < 0.0) ? 0.0 : ((s > 0.999) ? 0.999 : s);
  const double f = std::floor(256 * c);
  return (f == f) ? (uint8_t)(int)f : (uint8_t)0;  // floor NaN :: Word8 is platform-defined in GHC; 0 here
}
RGB albedoToColor(const Albedo& a) { return RGB{scaleColor(a.v.x), scaleColor(a.v.y), scaleColor(a.v.z)}; }

std::string showRow(const std::vector<RGB>& row) {
  std::string s;
  s.reserve(row.size() * 12);
  for (size_t x = 0; x < row.size(); ++x) {
    if (x) s += ' ';
    s += std::to_string((int)row[x].r); s += ' '; s += std::to_string((int)row[x].g); s += ' '; s += std::to_string((int)row[x].b);
  }
  return s;
}
void printRow(FILE* handle, int imageHeight, int i, const std::vector<RGB>& row) {
  std::fprintf(stderr, "\rRendering row %d of %d", i, imageHeight);
  const std::string s = showRow(row);
  std::fwrite(s.data(), 1, s.size(), handle);
  std::fputc('\n', handle);
}
void writeImage(FILE* handle, const std::vector<std::vector<RGB>>& rows, bool binary, bool progress) {
  const int H = (int)rows.size(), W = H ? (int)rows[0].size() : 0;
  std::fprintf(handle, "%s\n%d %d\n255\n", binary ? "P6" : "P3", W, H);
  for (int i = 0; i < H; ++i) {
    if (binary) {
      static_assert(sizeof(RGB) == 3, "RGB is three packed bytes");
      std::fwrite(rows[i].data(), 3, rows[i].size(), handle);
    } else if (progress) {
      printRow(handle, H, i + 1, rows[i]);
    } else {
      const std::string s = showRow(rows[i]);
      std::fwrite(s.data(), 1, s.size(), handle);
      std::fputc('\n', handle);
    }
  }
}

std::vector<std::vector<RGB>> runRender(const RenderStaticEnv& env, std::vector<RandGen> gens) {
  if (gens.empty()) throw std::runtime_error("runRender: no generators");
  Arena::Flat flat;
  env.arena->flatten(env.scene.world, env.scene.lights, env.scene.background, env.camera, flat);
  // The reference parallelises inside runRender (sparks over the pixels of a row, Lib.hs:1519-1520); the drop-in shards inside
  // rt_render over every visible GPU.  RTB_GPUS=n pins the count (1 = single device); decided once per process.
  static const int gpus = [] {
    const char* v = std::getenv("RTB_GPUS");
    const int want = v ? std::atoi(v) : 0;
    return (rt_device_count() > 1 && want != 1) ? rt_init(want, 0) : 1;
  }();
  (void)gpus;
  rt_scene_t sc = 0;
  if (rt_scene_create(&flat.desc, -1, &sc) != RT_OK) throw std::runtime_error(std::string("rt_scene_create: ") + rt_last_error());
  RtRenderParams p{};
  p.width = (uint32_t)env.width; p.height = (uint32_t)env.height; p.sample_offset = 0; p.sample_count = (uint32_t)env.numSamples;
  p.total_samples = (uint32_t)env.numSamples; p.max_depth = (uint32_t)env.maxDepth; p.seed = genWord64(gens[0]); p.precision = RT_PREC_DEFAULT; p.device = -1;
  p.n_gpus = 0;  // the library default: one device, or what rt_init(n) set (all the box's GPUs for rt_init(0))
  std::vector<double> sums((size_t)env.width * env.height * 3, 0.0);
  const int rc = rt_render(sc, &p, sums.data(), nullptr);
  const std::string err = rc == RT_OK ? "" : rt_last_error();
  rt_scene_destroy(sc);
  if (rc != RT_OK) throw std::runtime_error("rt_render: " + err);
  std::vector<std::vector<RGB>> rows((size_t)env.height, std::vector<RGB>((size_t)env.width));
  const double ns = (double)env.numSamples;
  for (int r = 0; r < env.height; ++r)
    for (int x = 0; x < env.width; ++x) {
      const double* s = &sums[((size_t)r * env.width + x) * 3];
      rows[r][x] = albedoToColor(albedo(s[0] / ns, s[1] / ns, s[2] / ns));  // divide summedColor ns (Lib.hs:1349)
    }
  return rows;
}

}  // namespace Lib
