// host/Lib.hpp — C++ mirror of the reference's host-side scene API (src/Lib.hs export list, lines 4-51), for hosts
// that cannot run the Haskell original: same names, argument order and meaning.  Scene values live in an Arena
// (struct-of-arrays, index handles) instead of a Haskell heap; `flatten` walks them into the RtSceneDesc the C ABI
// takes — the job INTEGRATION.md gives to the Haskell-side flattener.  `runRender` keeps the reference's shape
// (env + generators in, rows of RGB out) and does all per-sample work through librt_b200.so.
// Nothing here traces rays on the CPU.
#pragma once
#include <cstdint>
#include <string>
#include <utility>
#include <cstdio>
#include <vector>

#include "../../include/rt_abi.h"

namespace Lib {

// ---- Vec3 / Albedo (Lib.hs:200-261, 353-357)
struct Vec3 { double x = 0, y = 0, z = 0; };
using Point3 = Vec3;
inline Vec3 point3(double x, double y, double z) { return Vec3{x, y, z}; }
inline double vecX(const Vec3& v) { return v.x; }
inline double vecY(const Vec3& v) { return v.y; }
inline double vecZ(const Vec3& v) { return v.z; }
Vec3 operator+(const Vec3& a, const Vec3& b);  // |+|
Vec3 operator-(const Vec3& a, const Vec3& b);  // |-|
Vec3 operator*(const Vec3& a, const Vec3& b);  // |*|
double length(const Vec3& v);
struct Albedo { Vec3 v; };
inline Albedo albedo(double r, double g, double b) { return Albedo{Vec3{r, g, b}}; }
using Time = double;

// ---- RandGen (Random.hs:11-29): System.Random.StdGen of random-1.2.0 = splitmix-0.1 SMGen
struct RandGen { uint64_t seed = 0, gamma = 1; };
RandGen randGen(int64_t s);                       // Random.hs:20-21
double randomDouble(RandGen& g);                  // Random.hs:23-25
uint64_t genWord64(RandGen& g);
// the RandomState monad threads one generator; here it is passed by reference (Lib.hs:1119-1150)
double randomDoubleM(RandGen& g);
double randomDoubleRM(RandGen& g, double mn, double mx);
Vec3 randomVec3DoubleM(RandGen& g);
Vec3 randomVec3DoubleRM(RandGen& g, double mn, double mx);

// ---- handles into the Arena
struct Texture { uint32_t id = RT_NONE; };
struct Material { uint32_t id = RT_NONE; };
struct Hittable { uint32_t ref = 0; };  // RT_REF(kind, index); default = Unhittable
struct Box { Vec3 box_min, box_max; };  // Lib.hs:793-796
enum Axis { XAxis = 0, YAxis = 1, ZAxis = 2 };
enum Plane { XYPlane = 0, XZPlane = 1, YZPlane = 2 };
struct Fuzz { double v; };
struct RefractiveIdx { double v; };
struct Image { uint32_t w = 0, h = 0; std::vector<uint8_t> rgb; };  // JP.Image PixelRGB8

struct Camera { Vec3 origin, llc, horiz, vert, u, v, w; double lensRadius = 0, t0 = 0, t1 = 0; };  // Lib.hs:1230-1251
Camera newCamera(Point3 lookfrom, Point3 lookat, Vec3 vup, double vfov, double aspect, double aperture, double focusDist, Time t0, Time t1);  // 1269-1295

// The heap the scene values live in.
class Arena {
 public:
  // textures (Lib.hs:394-419) and materials (339-345)
  Texture ConstantColor(Albedo a);
  Texture CheckerTexture(Texture odd, Texture even);
  Texture ImageTexture(const Image* im);  // nullptr = ImageTexture Nothing
  Texture makePerlin(RandGen& g, double sc);  // Lib.hs:424-439
  Material Lambertian(Texture t);
  Material Metal(Texture t, Fuzz f);
  Material Dielectric(RefractiveIdx r);
  Material DiffuseLight(Texture t);
  Material Isotropic(Texture t);
  // smart constructors (Lib.hs:587-660, 726-761, 789-791)
  Hittable sphere(Point3 c, double r, Material m);
  Hittable movingSphere(Point3 c0, Point3 c1, Time t0, Time t1, double r, Material m);
  Hittable rect(Plane p, double a0, double a1, double b0, double b1, double k, Material m);
  Hittable cuboid(Point3 mn, Point3 mx, Material m);
  Hittable translate(Point3 off, Hittable h);
  Hittable rotate(Axis axis, double angleDeg, Hittable h);
  Hittable constantMedium(double density, Texture tex, Hittable boundary);
  Hittable makeBVH(RandGen& g, const std::vector<Hittable>& htbls);  // Lib.hs:941-961 (Maybe (Time,Time) is never read)
  // Sequences of at least this many objects are built by rt_bvh_build on the device (same tree, same generator state
  // afterwards); 0 = always the host recursion.  Process-wide default, set by rth_set_device_bvh / RTB_DEVICE_BVH_MIN.
  static size_t deviceBVHMin;
  static double lastDeviceBVHms;  // device time of the most recent rt_bvh_build issued from makeBVH
  Box boundingBox(Hittable h) const;                                 // Lib.hs:905-927

  // flattened, canonically numbered copy of everything reachable from (world, lights): the RtSceneDesc arrays
  struct Flat {
    std::vector<double> node_box, sphere_cr, moving, rect, cuboid, translate_off, rotate_sc, medium_nid, mat_param, tex_rgb, tex_scale, perlin_ranvec;
    std::vector<uint32_t> node_left, node_right, node_size, sphere_mat, moving_mat, rect_mat, cuboid_mat, translate_child, rotate_axis, rotate_child,
        medium_mat, medium_boundary, mat_kind, mat_tex, tex_kind, tex_a, tex_b, image_w, image_h;
    std::vector<int32_t> perlin_perm;
    std::vector<const uint8_t*> image_rgb;
    RtSceneDesc desc{};
  };
  // Scene = (world, lights, background) (Lib.hs:84)
  void flatten(Hittable world, Hittable lights, Albedo background, const Camera& cam, Flat& out) const;

 private:
  friend struct FlattenCtx;
  struct Tex { uint32_t kind; double rgb[3]; uint32_t a, b; double scale; };
  struct Mat { uint32_t kind; uint32_t tex; double param; };
  struct Perlin { std::vector<double> ranvec; std::vector<int32_t> perm; };
  std::vector<Tex> texs_;
  std::vector<Mat> mats_;
  std::vector<Perlin> perlins_;
  std::vector<const Image*> images_;
  // objects, one array per kind
  struct SphereO { Vec3 c; double r; uint32_t mat; };
  struct MovingO { Vec3 c0, c1; double t0, t1, dur, r; uint32_t mat; };
  struct RectO { double a0, a1, b0, b1, k; uint32_t mat; };
  struct CuboidO { Vec3 mn, mx; uint32_t mat; };
  struct NodeO { uint32_t left, right; Box box; uint32_t size; };
  struct TranslateO { Vec3 off; uint32_t child; };
  struct RotateO { uint32_t axis; double s, c; Box box; uint32_t child; };
  struct MediumO { double nInvD; uint32_t mat; uint32_t boundary; };
  std::vector<SphereO> spheres_;
  std::vector<MovingO> moving_;
  std::vector<RectO> rects_;
  std::vector<CuboidO> cuboids_;
  std::vector<NodeO> nodes_;
  std::vector<TranslateO> translates_;
  std::vector<RotateO> rotates_;
  std::vector<MediumO> media_;
  Material addMat(uint32_t kind, uint32_t tex, double param);
};

using Scene = struct SceneT { Hittable world, lights; Albedo background; };

// ---- render env + entry point (Lib.hs:89-108, 1491-1523)
struct RGB { uint8_t r, g, b; };
struct RenderStaticEnv { const Arena* arena; Scene scene; Camera camera; int width, height, numSamples, maxDepth, numThreads; };
RenderStaticEnv mkRenderStaticEnv(const Arena& arena, Scene s, Camera c, std::pair<int, int> wh, int numSamples, int maxDepth, int numThreads);
// Returns rows top first, like the reference.  gens: the reference passes one generator per column; the device RNG is
// counter-based, so only the head generator is consumed (one genWord64 = the stream seed).  Throws std::runtime_error
// with rt_last_error() if the device path fails — there is no CPU fallback.
std::vector<std::vector<RGB>> runRender(const RenderStaticEnv& env, std::vector<RandGen> gens);
RGB albedoToColor(const Albedo& a);  // Lib.hs:287-292 (host-side, as in the reference)
// Lib.hs:299-305: one text row "r g b r g b ..." + newline on `handle`, progress "\rRendering row i of H" on stderr
std::string showRow(const std::vector<RGB>& row);
void printRow(FILE* handle, int imageHeight, int i, const std::vector<RGB>& row);
// The image file app/Main.hs:58-62 prints: "P3\nW H\n255\n" + printRow per row (binary = false), or the same header
// with P6 and raw bytes (binary = true; SURVEY 8(f) rank 2: at 4K the text path dominates once rendering takes seconds)
void writeImage(FILE* handle, const std::vector<std::vector<RGB>>& rows, bool binary, bool progress);

}  // namespace Lib
