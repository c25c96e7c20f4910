// host/Scenes.cpp — C++ mirror of the reference's src/Scenes.hs (the ten scene builders and four cameras) on the
// Arena API of Lib.hpp, plus the config-5 synthetic scene of BASELINE.md.  Draw order follows the Haskell monadic
// order line by line so that `randGen 1024` produces the same scene as the reference's constructors would.
#include "Scenes.hpp"

#include <stdexcept>

namespace Scenes {
using namespace Lib;

// `lit` (not in the reference): replace the HEAD light list, which makes the mixture pdf degenerate to 0/0 (Lib.hs:823-836 with an
// Unhittable or XY-rect light), by a usable one; used only for NaN-free radiance parity tests.  The phantom light is not in the world.
static Hittable phantomLight(Arena& A) { return A.sphere(point3(0, 60, 0), 10, A.DiffuseLight(A.ConstantColor(albedo(1, 1, 1)))); }

static SceneAndGen done(Hittable world, Hittable lights, Albedo bg, const RandGen& g) { return SceneAndGen{Scene{world, lights, bg}, g}; }

// Scenes.hs:32-73
SceneAndGen makeCornellBoxScene(Arena& A, Time, Time, RandGen gen) {
  Material red = A.Lambertian(A.ConstantColor(albedo(0.65, 0.05, 0.05)));
  Material white = A.Lambertian(A.ConstantColor(albedo(0.73, 0.73, 0.73)));
  Material green = A.Lambertian(A.ConstantColor(albedo(0.12, 0.45, 0.15)));
  Material light = A.DiffuseLight(A.ConstantColor(albedo(15, 15, 15)));
  Hittable lightHittable = A.rect(XZPlane, 213, 343, 227, 332, 554, light);
  Hittable box1 = A.translate(point3(265, 0, 295), A.rotate(YAxis, 15, A.cuboid(point3(0, 0, 0), point3(165, 330, 165), white)));
  Hittable glassSphere = A.sphere(point3(190, 90, 190), 90, A.Dielectric(RefractiveIdx{1.5}));
  std::vector<Hittable> objs = {A.rect(YZPlane, 0, 555, 0, 555, 555, green), A.rect(YZPlane, 0, 555, 0, 555, 0, red), lightHittable,
                                A.rect(XZPlane, 0, 555, 0, 555, 0, white), A.rect(XZPlane, 0, 555, 0, 555, 555, white),
                                A.rect(XYPlane, 0, 555, 0, 555, 555, white), box1, glassSphere};
  Hittable world = A.makeBVH(gen, objs);
  Hittable lights = A.makeBVH(gen, {lightHittable, glassSphere});
  return done(world, lights, albedo(0, 0, 0), gen);
}

// Scenes.hs:75-118
SceneAndGen makeCornellSmokeBoxScene(Arena& A, Time, Time, RandGen gen) {
  Material light = A.DiffuseLight(A.ConstantColor(albedo(7, 7, 7)));
  Hittable lightHittable = A.rect(XZPlane, 113, 443, 127, 432, 554, light);
  Material red = A.Lambertian(A.ConstantColor(albedo(0.65, 0.05, 0.05)));
  Material white = A.Lambertian(A.ConstantColor(albedo(0.73, 0.73, 0.73)));
  Material green = A.Lambertian(A.ConstantColor(albedo(0.12, 0.45, 0.15)));
  std::vector<Hittable> objs = {
      A.rect(YZPlane, 0, 555, 0, 555, 555, green), A.rect(YZPlane, 0, 555, 0, 555, 0, red), lightHittable,
      A.rect(XZPlane, 0, 555, 0, 555, 0, white), A.rect(XZPlane, 0, 555, 0, 555, 555, white), A.rect(XYPlane, 0, 555, 0, 555, 555, white),
      A.constantMedium(0.01, A.ConstantColor(albedo(0, 0, 0)),
                       A.translate(point3(265, 0, 295), A.rotate(YAxis, 15, A.cuboid(point3(0, 0, 0), point3(165, 330, 165), white)))),
      A.constantMedium(0.01, A.ConstantColor(albedo(1, 1, 1)),
                       A.translate(point3(130, 0, 65), A.rotate(YAxis, -18, A.cuboid(point3(0, 0, 0), point3(165, 165, 165), white))))};
  Hittable world = A.makeBVH(gen, objs);
  return done(world, lightHittable, albedo(0, 0, 0), gen);
}

// Scenes.hs:120-131
Camera cornellCamera(std::pair<int, int> wh) {
  return newCamera(point3(278, 278, -800), point3(278, 278, 0.0), point3(0, 1, 0), 40.0, (double)wh.first / (double)wh.second, 0.0, 10.0, 0.0, 1.0);
}

// Scenes.hs:133-155
SceneAndGen makeSimpleLightScene(Arena& A, Time, Time, RandGen gen, bool lit) {
  Material difflight = A.DiffuseLight(A.ConstantColor(albedo(4, 4, 4)));
  Hittable sphereLight = A.sphere(point3(0, 7, 0), 2, difflight);
  Hittable rectLight = A.rect(XYPlane, 3, 5, 1, 3, -2, difflight);
  Texture perText = A.makePerlin(gen, 1.0);
  Hittable world = A.makeBVH(gen, {A.sphere(point3(0, -1000, 0), 1000, A.Lambertian(perText)), A.sphere(point3(0, 2, 0), 2, A.Lambertian(perText)), sphereLight, rectLight});
  Hittable lights = A.makeBVH(gen, {sphereLight, rectLight});
  if (lit) lights = sphereLight;
  return done(world, lights, albedo(0, 0, 0), gen);
}

// Scenes.hs:167-179
SceneAndGen makeEarthScene(Arena& A, Texture earthTex, Time, Time, RandGen gen, bool lit) {
  Hittable world = A.makeBVH(gen, {A.sphere(point3(0, 0, 0), 2, A.Lambertian(earthTex))});
  return done(world, lit ? phantomLight(A) : Hittable{}, albedo(1.0, 1.0, 1.0), gen);
}

// Scenes.hs:181-192
Camera twoSpheresSceneCamera(std::pair<int, int> wh) {
  return newCamera(point3(26, 4, 6), point3(0, 2, 0), point3(0, 1, 0), 20.0, (double)wh.first / (double)wh.second, 0.1, 20.0, 0.0, 1.0);
}

// Scenes.hs:194-211
SceneAndGen makeTwoPerlinSpheresScene(Arena& A, Time, Time, RandGen gen, bool lit) {
  Texture perText = A.makePerlin(gen, 1.5);
  Hittable world = A.makeBVH(gen, {A.sphere(point3(0, -1000, 0), 1000, A.Lambertian(perText)), A.sphere(point3(0, 2, 0), 2, A.Lambertian(perText))});
  return done(world, lit ? phantomLight(A) : Hittable{}, albedo(0, 0, 0), gen);
}

// Scenes.hs:213-237
SceneAndGen makeTwoSpheresScene(Arena& A, Time, Time, RandGen gen, bool lit) {
  Material checkerMaterial = A.Metal(A.CheckerTexture(A.ConstantColor(albedo(0.2, 0.3, 0.1)), A.ConstantColor(albedo(0.9, 0.9, 0.9))), Fuzz{0.0});
  Material flatMaterial = A.Lambertian(A.ConstantColor(albedo(0.6, 0.2, 0.1)));
  Hittable world = A.makeBVH(gen, {A.sphere(point3(0, -10, 0), 10, checkerMaterial), A.sphere(point3(0, 10, 0), 10, flatMaterial)});
  return done(world, lit ? phantomLight(A) : Hittable{}, albedo(0.8, 0.8, 0.9), gen);
}

// Scenes.hs:239-250
Camera randomSceneCamera(std::pair<int, int> wh) {
  return newCamera(point3(13, 2, 3), point3(0, 0, 0), point3(0, 1, 0), 20.0, (double)wh.first / (double)wh.second, 0.1, 10.0, 0.0, 1.0);
}

// makeRandomSphereM of Scenes.hs:284-317 (moving = false) and 364-399 (moving = true)
static bool randomSphere(Arena& A, RandGen& gen, int a, int b, bool moving, Hittable& out) {
  const double mat = randomDoubleM(gen);
  const double px = randomDoubleM(gen);
  const double py = randomDoubleM(gen);
  const Point3 center = point3((double)a + 0.9 * px, 0.2, (double)b + 0.9 * py);
  if (Lib::length(center - point3(4.0, 0.2, 0)) <= 0.9) return false;
  if (mat < 0.8) {
    const Vec3 a1 = randomVec3DoubleM(gen);
    const Vec3 a2 = randomVec3DoubleM(gen);
    if (moving) {
      const double mx = randomDoubleRM(gen, -0.25, 0.25);
      const double mz = randomDoubleRM(gen, -0.25, 0.25);
      out = A.movingSphere(center, center + point3(mx, 0, mz), 0.0, 1.0, 0.2, A.Lambertian(A.ConstantColor(Albedo{a1 * a2})));
    } else {
      out = A.sphere(center, 0.2, A.Lambertian(A.ConstantColor(Albedo{a1 * a2})));
    }
  } else if (mat < 0.95) {
    const Vec3 alb = randomVec3DoubleRM(gen, 0.5, 1.0);
    const double fuzz = randomDoubleRM(gen, 0.0, 0.5);
    out = A.sphere(center, 0.2, A.Metal(A.ConstantColor(Albedo{alb}), Fuzz{fuzz}));
  } else {
    out = A.sphere(center, 0.2, A.Dielectric(RefractiveIdx{1.5}));
  }
  return true;
}

// Scenes.hs:253-283
SceneAndGen makeRandomSceneBookOne(Arena& A, RandGen gen, bool lit) {
  std::vector<Hittable> objs = {A.sphere(point3(0, -1000, 0), 1000, A.Lambertian(A.ConstantColor(albedo(0.5, 0.5, 0.5)))),
                                A.sphere(point3(0, 1, 0), 1.0, A.Dielectric(RefractiveIdx{1.5})),
                                A.sphere(point3(-4, 1, 0), 1.0, A.Lambertian(A.ConstantColor(albedo(0.4, 0.2, 0.1)))),
                                A.sphere(point3(4, 1, 0), 1.0, A.Metal(A.ConstantColor(albedo(0.7, 0.6, 0.5)), Fuzz{0.0}))};
  for (int a = -11; a <= 10; ++a) for (int b = -11; b <= 10; ++b) { Hittable h; if (randomSphere(A, gen, a, b, false, h)) objs.push_back(h); }
  Hittable world = A.makeBVH(gen, objs);
  return done(world, lit ? phantomLight(A) : Hittable{}, albedo(0.7, 0.8, 0.9), gen);
}

// Scenes.hs:321-363
SceneAndGen makeRandomScene(Arena& A, Texture earthtex, Time, Time, RandGen gen, bool lit) {
  std::vector<Hittable> objs = {
      A.sphere(point3(0, -1000, 0), 1000, A.Lambertian(A.CheckerTexture(A.ConstantColor(albedo(0.2, 0.3, 0.1)), A.ConstantColor(albedo(0.9, 0.9, 0.9))))),
      A.cuboid(point3(-0.75, 0.0, -0.75), point3(0.75, 1.5, 0.75), A.Dielectric(RefractiveIdx{1.5})),
      A.sphere(point3(-4, 1, 0), 1.0, A.Lambertian(earthtex)),
      A.sphere(point3(4, 1, 0), 1.0, A.Metal(A.ConstantColor(albedo(0.7, 0.6, 0.5)), Fuzz{0.0}))};
  for (int a = -11; a <= 10; ++a) for (int b = -11; b <= 10; ++b) { Hittable h; if (randomSphere(A, gen, a, b, true, h)) objs.push_back(h); }
  Hittable world = A.makeBVH(gen, objs);
  return done(world, lit ? phantomLight(A) : Hittable{}, albedo(0.7, 0.8, 0.9), gen);
}

// Scenes.hs:401-412
Camera nextWeekFinalSceneCamera(std::pair<int, int> wh) {
  return newCamera(point3(575, 278, -525), point3(320, 278, 0.0), point3(0, 1, 0), 40.0, (double)wh.first / (double)wh.second, 0.1, 580.0, 0.0, 1.0);
}

// Scenes.hs:414-466
SceneAndGen makeNextWeekFinalScene(Arena& A, Texture earthtex, Time t0, Time t1, RandGen gen, bool lit) {
  Material ground = A.Lambertian(A.ConstantColor(albedo(0.48, 0.83, 0.53)));
  Material white = A.Lambertian(A.ConstantColor(albedo(0.73, 0.73, 0.73)));
  const double w = 100, y0 = 0;
  std::vector<Hittable> b1;
  for (int ii = 0; ii <= 19; ++ii) for (int jj = 0; jj <= 19; ++jj) {
    const double x0 = ii * w - 1000, z0 = jj * w - 1000, x1 = x0 + w;
    const double y1 = randomDoubleRM(gen, 1, 101);
    const double z1 = z0 + w;
    b1.push_back(A.cuboid(point3(x0, y0, z0), point3(x1, y1, z1), ground));
  }
  Hittable boxes1 = A.makeBVH(gen, b1);
  Material light = A.DiffuseLight(A.ConstantColor(albedo(7, 7, 7)));
  Hittable boundary1 = A.sphere(point3(360, 150, 145), 70, A.Dielectric(RefractiveIdx{1.5}));
  Hittable boundary2 = A.sphere(point3(0, 0, 0), 5000, A.Dielectric(RefractiveIdx{1.5}));
  Texture pertext = A.makePerlin(gen, 0.1);
  std::vector<Hittable> b2;
  for (int i = 0; i < 1000; ++i) { const Vec3 p = randomVec3DoubleRM(gen, 0, 165); b2.push_back(A.sphere(p, 10, white)); }
  Hittable boxes2 = A.makeBVH(gen, b2);
  Hittable lightRect = A.rect(XZPlane, 113, 443, 127, 432, 554, light);
  std::vector<Hittable> objs = {
      boxes1, lightRect,
      A.movingSphere(point3(400, 400, 200), point3(430, 400, 200), t0, t1, 50, A.Lambertian(A.ConstantColor(albedo(0.7, 0.3, 0.1)))),
      A.sphere(point3(260, 150, 45), 50, A.Dielectric(RefractiveIdx{1.5})),
      A.sphere(point3(0, 150, 145), 50, A.Metal(A.ConstantColor(albedo(0.8, 0.8, 0.9)), Fuzz{10.0})),
      boundary1,
      A.constantMedium(0.2, A.ConstantColor(albedo(0.2, 0.4, 0.9)), boundary1),
      A.constantMedium(0.0001, A.ConstantColor(albedo(1, 1, 1)), boundary2),
      A.sphere(point3(400, 200, 400), 100, A.Lambertian(earthtex)),
      A.sphere(point3(220, 280, 300), 80, A.Lambertian(pertext)),
      A.translate(point3(-100, 270, 395), A.rotate(YAxis, 15, boxes2))};
  Hittable world = A.makeBVH(gen, objs);
  return done(world, lit ? lightRect : Hittable{}, albedo(0, 0, 0), gen);
}

// Config 5: n x n jittered spheres, Book-1 material mix, XZ-rect ground, one XZ-rect light
SceneAndGen makeSyntheticScene(Arena& A, int n, RandGen gen) {
  const int lo = -(n / 2);
  const double ext = (double)n / 2 + 1;
  std::vector<Hittable> objs;
  objs.reserve((size_t)n * n + 2);
  objs.push_back(A.rect(XZPlane, -ext, ext, -ext, ext, 0.0, A.Lambertian(A.ConstantColor(albedo(0.5, 0.5, 0.5)))));
  Hittable lightRect = A.rect(XZPlane, -ext * 0.5, ext * 0.5, -ext * 0.5, ext * 0.5, ext, A.DiffuseLight(A.ConstantColor(albedo(4, 4, 4))));
  objs.push_back(lightRect);
  for (int a = lo; a < lo + n; ++a) for (int b = lo; b < lo + n; ++b) {
    const double mat = randomDoubleM(gen);
    const double px = randomDoubleM(gen);
    const double py = randomDoubleM(gen);
    const Point3 center = point3((double)a + 0.9 * px, 0.2, (double)b + 0.9 * py);
    if (mat < 0.8) {
      const Vec3 a1 = randomVec3DoubleM(gen); const Vec3 a2 = randomVec3DoubleM(gen);
      objs.push_back(A.sphere(center, 0.2, A.Lambertian(A.ConstantColor(Albedo{a1 * a2}))));
    } else if (mat < 0.95) {
      const Vec3 alb = randomVec3DoubleRM(gen, 0.5, 1.0); const double fuzz = randomDoubleRM(gen, 0.0, 0.5);
      objs.push_back(A.sphere(center, 0.2, A.Metal(A.ConstantColor(Albedo{alb}), Fuzz{fuzz})));
    } else {
      objs.push_back(A.sphere(center, 0.2, A.Dielectric(RefractiveIdx{1.5})));
    }
  }
  Hittable world = A.makeBVH(gen, objs);
  return done(world, lightRect, albedo(0.7, 0.8, 0.9), gen);
}
Camera syntheticCamera(std::pair<int, int> wh, int n) {
  const double e = (double)n / 2;
  return newCamera(point3(0.9 * e, 0.45 * e, 0.9 * e), point3(0, 0, 0), point3(0, 1, 0), 40.0, (double)wh.first / (double)wh.second, 0.0, 10.0, 0.0, 1.0);
}

// "quirks": an API-coverage scene that is NOT in Scenes.hs, built only with the reference's constructors (Lib.hs:587-660, 726-761,
// 789-791, 941-961) in a Cornell-sized room: Rotate about X and Z (Lib.hs:763-787), three nested Translate / Rotate frames, Rotate
// above a BVH, one-object makeBVH (BVHNode x x) in the world and in the light list, an XY rect among the lights (htblRandom /
// htblPdfValue fallbacks, Lib.hs:684, 724), Metal fuzz > 1, a Perlin-textured medium in a rotated cuboid, a checker of a checker.
SceneAndGen makeQuirksScene(Arena& A, Time t0, Time t1, RandGen gen, bool lit) {
  Texture pertext = A.makePerlin(gen, 0.05);
  Texture chk = A.CheckerTexture(A.CheckerTexture(A.ConstantColor(albedo(0.2, 0.3, 0.1)), A.ConstantColor(albedo(0.9, 0.9, 0.9))), pertext);
  Material white = A.Lambertian(A.ConstantColor(albedo(0.73, 0.73, 0.73)));
  Material red = A.Lambertian(A.ConstantColor(albedo(0.65, 0.05, 0.05)));
  Material mirror = A.Metal(A.ConstantColor(albedo(0.8, 0.85, 0.9)), Fuzz{0.3});
  Material fuzzy = A.Metal(A.ConstantColor(albedo(0.9, 0.6, 0.2)), Fuzz{2.0});
  Material glass = A.Dielectric(RefractiveIdx{1.5});
  Material light = A.DiffuseLight(A.ConstantColor(albedo(15, 15, 15)));
  Material dim = A.DiffuseLight(A.ConstantColor(albedo(4, 4, 4)));
  Hittable xzLight = A.rect(XZPlane, 163, 393, 177, 382, 554, light);
  Hittable sphereLight = A.sphere(point3(460, 420, 120), 40, dim);
  Hittable xyLight = A.rect(XYPlane, 100, 250, 350, 500, 554, dim);
  std::vector<Hittable> cluster;
  for (int i = 0; i < 12; ++i) {
    const Vec3 p = randomVec3DoubleRM(gen, 0, 120);
    const double r = randomDoubleRM(gen, 8, 20);
    cluster.push_back(A.sphere(p, r, i % 3 == 0 ? glass : (i % 3 == 1 ? white : mirror)));
  }
  Hittable clusterBVH = A.makeBVH(gen, cluster);
  Hittable lone = A.makeBVH(gen, {A.sphere(point3(90, 60, 380), 60, fuzzy)});
  Material floorMat = A.Lambertian(chk);
  std::vector<Hittable> objs = {
      A.rect(YZPlane, 0, 555, 0, 555, 555, mirror), A.rect(YZPlane, 0, 555, 0, 555, 0, red), xzLight, sphereLight, xyLight,
      A.rect(XZPlane, 0, 555, 0, 555, 0, floorMat), A.rect(XZPlane, 0, 555, 0, 555, 555, white), A.rect(XYPlane, 0, 555, 0, 555, 555, white),
      A.translate(point3(360, 100, 260), A.rotate(XAxis, 25, A.cuboid(point3(-60, -60, -60), point3(60, 60, 60), white))),
      A.translate(point3(200, 250, 250), A.rotate(ZAxis, -40, clusterBVH)),
      A.rotate(XAxis, 10, A.translate(point3(420, 300, 300), A.rotate(ZAxis, 20, A.cuboid(point3(-40, -40, -40), point3(40, 40, 40), glass)))),
      lone,
      A.constantMedium(0.02, A.ConstantColor(albedo(0.9, 0.9, 0.9)), A.sphere(point3(140, 90, 150), 70, white)),
      A.constantMedium(0.03, pertext, A.translate(point3(420, 80, 120), A.rotate(ZAxis, 30, A.cuboid(point3(-50, -50, -50), point3(50, 50, 50), white)))),
      A.movingSphere(point3(280, 400, 300), point3(310, 430, 300), t0, t1, 35, A.Lambertian(A.ConstantColor(albedo(0.7, 0.3, 0.1))))};
  Hittable world = A.makeBVH(gen, objs);
  // lit: without the XY rect, whose (1,0,0) fallback direction makes the mixture pdf 0 on surfaces facing y or z (0/0)
  Hittable lights = lit ? A.makeBVH(gen, {xzLight, sphereLight}) : A.makeBVH(gen, {xzLight, sphereLight, xyLight});
  return done(world, lights, albedo(0, 0, 0), gen);
}

}  // namespace Scenes
