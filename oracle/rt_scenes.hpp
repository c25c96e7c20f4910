// oracle/rt_scenes.hpp — restatement of /root/reference/src/Scenes.hs (scene constructors + cameras) and of the
// config-5 synthetic scene described in BASELINE.md, on top of rt_oracle.hpp.  TEST INFRASTRUCTURE ONLY.
// RNG draw order follows the monadic order of the Haskell source line by line.
#pragma once
#include "rt_oracle.hpp"

namespace orc {

// lightsMode: 0 = literal HEAD light list; 1 = "lit" variant (a usable light list so that the HEAD mixture pdf
// never degenerates to 0/0) — used for NaN-free radiance parity, never for the literal configs.
struct SceneArgs { double t0 = 0.0, t1 = 1.0; std::shared_ptr<ImageData> earth; int earthW = 0, earthH = 0; int lightsMode = 0; int synthN = 1000; };

inline HitP phantomLight() {  // not in the world; only a sampling target for lit variants of sky-lit scenes
  return sphere(V3d(0, 60, 0), 10, diffuseLight(constantColor(1, 1, 1)));
}

// Scenes.hs:32-73
inline Scene makeCornellBoxScene(const SceneArgs& a, SMGen gen) {
  MatP red = lambertian(constantColor(0.65, 0.05, 0.05));
  MatP white = lambertian(constantColor(0.73, 0.73, 0.73));
  MatP green = lambertian(constantColor(0.12, 0.45, 0.15));
  MatP light = diffuseLight(constantColor(15, 15, 15));
  HitP lightHittable = rect(XZPlane, 213, 343, 227, 332, 554, light);
  HitP box1 = translate(V3d(265, 0, 295), rotate(YAxis, 15, cuboid(V3d(0, 0, 0), V3d(165, 330, 165), white)));
  HitP glassSphere = sphere(V3d(190, 90, 190), 90, dielectric(1.5));
  Scene s;
  s.world = makeBVH(gen, {rect(YZPlane, 0, 555, 0, 555, 555, green), rect(YZPlane, 0, 555, 0, 555, 0, red), lightHittable,
                          rect(XZPlane, 0, 555, 0, 555, 0, white), rect(XZPlane, 0, 555, 0, 555, 555, white),
                          rect(XYPlane, 0, 555, 0, 555, 555, white), box1, glassSphere});
  s.lights = makeBVH(gen, {lightHittable, glassSphere});
  s.g1 = gen;
  (void)a;
  return s;
}

// Scenes.hs:75-118
inline Scene makeCornellSmokeBoxScene(const SceneArgs& a, SMGen gen) {
  MatP light = diffuseLight(constantColor(7, 7, 7));
  HitP lightHittable = rect(XZPlane, 113, 443, 127, 432, 554, light);
  MatP red = lambertian(constantColor(0.65, 0.05, 0.05));
  MatP white = lambertian(constantColor(0.73, 0.73, 0.73));
  MatP green = lambertian(constantColor(0.12, 0.45, 0.15));
  Scene s;
  s.world = makeBVH(gen, {rect(YZPlane, 0, 555, 0, 555, 555, green), rect(YZPlane, 0, 555, 0, 555, 0, red), lightHittable,
                          rect(XZPlane, 0, 555, 0, 555, 0, white), rect(XZPlane, 0, 555, 0, 555, 555, white),
                          rect(XYPlane, 0, 555, 0, 555, 555, white),
                          constantMedium(0.01, constantColor(0, 0, 0),
                                         translate(V3d(265, 0, 295), rotate(YAxis, 15, cuboid(V3d(0, 0, 0), V3d(165, 330, 165), white)))),
                          constantMedium(0.01, constantColor(1, 1, 1),
                                         translate(V3d(130, 0, 65), rotate(YAxis, -18, cuboid(V3d(0, 0, 0), V3d(165, 165, 165), white))))});
  s.lights = lightHittable;
  s.g1 = gen;
  (void)a;
  return s;
}

// Scenes.hs:120-131
inline Camera cornellCamera(int W, int H) {
  return newCamera(V3d(278, 278, -800), V3d(278, 278, 0.0), V3d(0, 1, 0), 40.0, (double)W / (double)H, 0.0, 10.0, 0.0, 1.0);
}

// Scenes.hs:133-155
inline Scene makeSimpleLightScene(const SceneArgs& a, SMGen gen) {
  MatP difflight = diffuseLight(constantColor(4, 4, 4));
  HitP sphereLight = sphere(V3d(0, 7, 0), 2, difflight);
  HitP rectLight = rect(XYPlane, 3, 5, 1, 3, -2, difflight);
  TexP perText = makePerlin(gen, 1.0);
  Scene s;
  s.world = makeBVH(gen, {sphere(V3d(0, -1000, 0), 1000, lambertian(perText)), sphere(V3d(0, 2, 0), 2, lambertian(perText)), sphereLight, rectLight});
  s.lights = makeBVH(gen, {sphereLight, rectLight});
  if (a.lightsMode == 1) s.lights = sphereLight;
  s.g1 = gen;
  return s;
}

// Scenes.hs:167-179
inline Scene makeEarthScene(const SceneArgs& a, SMGen gen) {
  TexP earthTex = imageTexture(a.earth, a.earthW, a.earthH);
  Scene s;
  s.world = makeBVH(gen, {sphere(V3d(0, 0, 0), 2, lambertian(earthTex))});
  s.lights = a.lightsMode == 1 ? phantomLight() : unhittable();
  s.background[0] = s.background[1] = s.background[2] = 1.0;
  s.g1 = gen;
  return s;
}

// Scenes.hs:181-192
inline Camera twoSpheresSceneCamera(int W, int H) {
  return newCamera(V3d(26, 4, 6), V3d(0, 2, 0), V3d(0, 1, 0), 20.0, (double)W / (double)H, 0.1, 20.0, 0.0, 1.0);
}

// Scenes.hs:194-211
inline Scene makeTwoPerlinSpheresScene(const SceneArgs& a, SMGen gen) {
  TexP perText = makePerlin(gen, 1.5);
  Scene s;
  s.world = makeBVH(gen, {sphere(V3d(0, -1000, 0), 1000, lambertian(perText)), sphere(V3d(0, 2, 0), 2, lambertian(perText))});
  s.lights = a.lightsMode == 1 ? phantomLight() : unhittable();
  s.g1 = gen;
  return s;
}

// Scenes.hs:213-237
inline Scene makeTwoSpheresScene(const SceneArgs& a, SMGen gen) {
  MatP checkerMaterial = metal(checkerTexture(constantColor(0.2, 0.3, 0.1), constantColor(0.9, 0.9, 0.9)), 0.0);
  MatP flatMaterial = lambertian(constantColor(0.6, 0.2, 0.1));
  Scene s;
  s.world = makeBVH(gen, {sphere(V3d(0, -10, 0), 10, checkerMaterial), sphere(V3d(0, 10, 0), 10, flatMaterial)});
  s.lights = a.lightsMode == 1 ? phantomLight() : unhittable();
  s.background[0] = 0.8; s.background[1] = 0.8; s.background[2] = 0.9;
  s.g1 = gen;
  return s;
}

// Scenes.hs:239-250
inline Camera randomSceneCamera(int W, int H) {
  return newCamera(V3d(13, 2, 3), V3d(0, 0, 0), V3d(0, 1, 0), 20.0, (double)W / (double)H, 0.1, 10.0, 0.0, 1.0);
}

// Scenes.hs:253-317 (moving=false) and 321-399 (moving=true; the "tweaked" Next Week variant)
inline Scene makeRandomSceneImpl(const SceneArgs& a, SMGen gen, bool nextWeek) {
  HitP ground, s1, s2;
  if (!nextWeek) {
    ground = sphere(V3d(0, -1000, 0), 1000, lambertian(constantColor(0.5, 0.5, 0.5)));
    s1 = sphere(V3d(0, 1, 0), 1.0, dielectric(1.5));
    s2 = sphere(V3d(-4, 1, 0), 1.0, lambertian(constantColor(0.4, 0.2, 0.1)));
  } else {
    ground = sphere(V3d(0, -1000, 0), 1000, lambertian(checkerTexture(constantColor(0.2, 0.3, 0.1), constantColor(0.9, 0.9, 0.9))));
    s1 = cuboid(V3d(-0.75, 0.0, -0.75), V3d(0.75, 1.5, 0.75), dielectric(1.5));
    s2 = sphere(V3d(-4, 1, 0), 1.0, lambertian(imageTexture(a.earth, a.earthW, a.earthH)));
  }
  HitP s3 = sphere(V3d(4, 1, 0), 1.0, metal(constantColor(0.7, 0.6, 0.5), 0.0));
  std::vector<HitP> objs = {ground, s1, s2, s3};
  for (int aa = -11; aa <= 10; ++aa) for (int bb = -11; bb <= 10; ++bb) {  // makeRandomSphereM
    double mat = randomDoubleM(gen);
    double px = randomDoubleM(gen);
    double py = randomDoubleM(gen);
    V3d center((double)aa + 0.9 * px, 0.2, (double)bb + 0.9 * py);
    if (length(center - V3d(4.0, 0.2, 0)) <= 0.9) continue;
    if (mat < 0.8) {
      V3d a1 = randomVec3DoubleM(gen);
      V3d a2 = randomVec3DoubleM(gen);
      if (nextWeek) {
        double mx = randomDoubleRM(gen, -0.25, 0.25);
        double mz = randomDoubleRM(gen, -0.25, 0.25);
        V3d alb = a1 * a2;
        objs.push_back(movingSphere(center, center + V3d(mx, 0, mz), 0.0, 1.0, 0.2, lambertian(constantColor(alb.x, alb.y, alb.z))));
      } else {
        V3d alb = a1 * a2;
        objs.push_back(sphere(center, 0.2, lambertian(constantColor(alb.x, alb.y, alb.z))));
      }
    } else if (mat < 0.95) {
      V3d alb = randomVec3DoubleRM(gen, 0.5, 1.0);
      double fuzz = randomDoubleRM(gen, 0.0, 0.5);
      objs.push_back(sphere(center, 0.2, metal(constantColor(alb.x, alb.y, alb.z), fuzz)));
    } else {
      objs.push_back(sphere(center, 0.2, dielectric(1.5)));
    }
  }
  Scene s;
  s.world = makeBVH(gen, objs);
  s.lights = a.lightsMode == 1 ? phantomLight() : unhittable();
  s.background[0] = 0.7; s.background[1] = 0.8; s.background[2] = 0.9;
  s.g1 = gen;
  return s;
}
inline Scene makeRandomSceneBookOne(const SceneArgs& a, SMGen gen) { return makeRandomSceneImpl(a, gen, false); }
inline Scene makeRandomScene(const SceneArgs& a, SMGen gen) { return makeRandomSceneImpl(a, gen, true); }

// Scenes.hs:401-412
inline Camera nextWeekFinalSceneCamera(int W, int H) {
  return newCamera(V3d(575, 278, -525), V3d(320, 278, 0.0), V3d(0, 1, 0), 40.0, (double)W / (double)H, 0.1, 580.0, 0.0, 1.0);
}

// Scenes.hs:414-466
inline Scene makeNextWeekFinalScene(const SceneArgs& a, SMGen gen) {
  MatP ground = lambertian(constantColor(0.48, 0.83, 0.53));
  MatP white = lambertian(constantColor(0.73, 0.73, 0.73));
  const double w = 100, y0 = 0;
  std::vector<HitP> b1;
  for (int ii = 0; ii <= 19; ++ii) for (int jj = 0; jj <= 19; ++jj) {
    double i = ii, j = jj;
    double x0 = i * w - 1000, z0 = j * w - 1000;
    double x1 = x0 + w;
    double y1 = randomDoubleRM(gen, 1, 101);
    double z1 = z0 + w;
    b1.push_back(cuboid(V3d(x0, y0, z0), V3d(x1, y1, z1), ground));
  }
  HitP boxes1 = makeBVH(gen, b1);
  MatP light = diffuseLight(constantColor(7, 7, 7));
  HitP boundary1 = sphere(V3d(360, 150, 145), 70, dielectric(1.5));
  HitP boundary2 = sphere(V3d(0, 0, 0), 5000, dielectric(1.5));
  TexP pertext = makePerlin(gen, 0.1);
  std::vector<HitP> b2;
  for (int i = 0; i < 1000; ++i) { V3d p = randomVec3DoubleRM(gen, 0, 165); b2.push_back(sphere(p, 10, white)); }
  HitP boxes2 = makeBVH(gen, b2);
  HitP lightRect = rect(XZPlane, 113, 443, 127, 432, 554, light);
  Scene s;
  s.world = makeBVH(gen, {boxes1, lightRect,
                          movingSphere(V3d(400, 400, 200), V3d(430, 400, 200), a.t0, a.t1, 50, lambertian(constantColor(0.7, 0.3, 0.1))),
                          sphere(V3d(260, 150, 45), 50, dielectric(1.5)),
                          sphere(V3d(0, 150, 145), 50, metal(constantColor(0.8, 0.8, 0.9), 10.0)),
                          boundary1,
                          constantMedium(0.2, constantColor(0.2, 0.4, 0.9), boundary1),
                          constantMedium(0.0001, constantColor(1, 1, 1), boundary2),
                          sphere(V3d(400, 200, 400), 100, lambertian(imageTexture(a.earth, a.earthW, a.earthH))),
                          sphere(V3d(220, 280, 300), 80, lambertian(pertext)),
                          translate(V3d(-100, 270, 395), rotate(YAxis, 15, boxes2))});
  s.lights = a.lightsMode == 1 ? lightRect : unhittable();
  s.g1 = gen;
  return s;
}

// Config 5 (not in the reference, built with the reference's constructors):
// n x n spheres r=0.2 on a jittered grid, Book-1 material mix in makeRandomSphereM's draw order, XZ-rect ground,
// one overhead XZ-rect light which is also the light list, makeBVH over everything.
inline Scene makeSyntheticScene(const SceneArgs& a, SMGen gen) {
  int n = a.synthN; int lo = -(n / 2);
  double ext = (double)n / 2 + 1;
  std::vector<HitP> objs;
  objs.reserve((size_t)n * n + 2);
  objs.push_back(rect(XZPlane, -ext, ext, -ext, ext, 0.0, lambertian(constantColor(0.5, 0.5, 0.5))));
  HitP lightRect = rect(XZPlane, -ext * 0.5, ext * 0.5, -ext * 0.5, ext * 0.5, ext, diffuseLight(constantColor(4, 4, 4)));
  objs.push_back(lightRect);
  for (int aa = lo; aa < lo + n; ++aa) for (int bb = lo; bb < lo + n; ++bb) {
    double mat = randomDoubleM(gen);
    double px = randomDoubleM(gen);
    double py = randomDoubleM(gen);
    V3d center((double)aa + 0.9 * px, 0.2, (double)bb + 0.9 * py);
    if (mat < 0.8) {
      V3d a1 = randomVec3DoubleM(gen); V3d a2 = randomVec3DoubleM(gen); V3d alb = a1 * a2;
      objs.push_back(sphere(center, 0.2, lambertian(constantColor(alb.x, alb.y, alb.z))));
    } else if (mat < 0.95) {
      V3d alb = randomVec3DoubleRM(gen, 0.5, 1.0); double fuzz = randomDoubleRM(gen, 0.0, 0.5);
      objs.push_back(sphere(center, 0.2, metal(constantColor(alb.x, alb.y, alb.z), fuzz)));
    } else {
      objs.push_back(sphere(center, 0.2, dielectric(1.5)));
    }
  }
  Scene s;
  s.world = makeBVH(gen, std::move(objs));
  s.lights = lightRect;
  s.background[0] = 0.7; s.background[1] = 0.8; s.background[2] = 0.9;
  s.g1 = gen;
  return s;
}
// elevated look-down camera for config 5, aperture 0; scaled with the grid so the whole field is in view
inline Camera syntheticCamera(int W, int H, int n) {
  double e = (double)n / 2;
  return newCamera(V3d(0.9 * e, 0.45 * e, 0.9 * e), V3d(0, 0, 0), V3d(0, 1, 0), 40.0, (double)W / (double)H, 0.0, 10.0, 0.0, 1.0);
}

// API-coverage scene "quirks" (NOT in the reference; built only with the reference's constructors, Lib.hs:587-660, 726-761,
// 789-791, 941-961): the combinations Scenes.hs never uses but Lib.hs supports, inside a Cornell-sized room —
// Rotate about the X and Z axes (Lib.hs:763-787), three nested Translate/Rotate frames, a Rotate above a BVH, a one-object
// makeBVH (Q3: BVHNode x x) in the world and in the light list, an XY rect in the light list (Q10: htblRandom falls back to
// (1,0,0), htblPdfValue to 0), Metal with fuzz > 1 (Q12), a ConstantMedium with a Perlin phase texture around a rotated
// cuboid, a checker of a checker and a Perlin texture (Lib.hs:496-503), a moving sphere next to all of them.
inline Scene makeQuirksScene(const SceneArgs& a, SMGen gen) {
  TexP pertext = makePerlin(gen, 0.05);
  TexP chk = checkerTexture(checkerTexture(constantColor(0.2, 0.3, 0.1), constantColor(0.9, 0.9, 0.9)), pertext);
  MatP white = lambertian(constantColor(0.73, 0.73, 0.73));
  MatP red = lambertian(constantColor(0.65, 0.05, 0.05));
  MatP mirror = metal(constantColor(0.8, 0.85, 0.9), 0.3);
  MatP fuzzy = metal(constantColor(0.9, 0.6, 0.2), 2.0);
  MatP glass = dielectric(1.5);
  MatP light = diffuseLight(constantColor(15, 15, 15));
  MatP dim = diffuseLight(constantColor(4, 4, 4));
  HitP xzLight = rect(XZPlane, 163, 393, 177, 382, 554, light);
  HitP sphereLight = sphere(V3d(460, 420, 120), 40, dim);
  HitP xyLight = rect(XYPlane, 100, 250, 350, 500, 554, dim);
  std::vector<HitP> cluster;
  for (int i = 0; i < 12; ++i) {
    V3d p = randomVec3DoubleRM(gen, 0, 120);
    double r = randomDoubleRM(gen, 8, 20);
    cluster.push_back(sphere(p, r, i % 3 == 0 ? glass : (i % 3 == 1 ? white : mirror)));
  }
  HitP clusterBVH = makeBVH(gen, cluster);
  HitP lone = makeBVH(gen, {sphere(V3d(90, 60, 380), 60, fuzzy)});
  Scene s;
  s.world = makeBVH(gen, {rect(YZPlane, 0, 555, 0, 555, 555, mirror), rect(YZPlane, 0, 555, 0, 555, 0, red), xzLight, sphereLight, xyLight,
                          rect(XZPlane, 0, 555, 0, 555, 0, lambertian(chk)), rect(XZPlane, 0, 555, 0, 555, 555, white),
                          rect(XYPlane, 0, 555, 0, 555, 555, white),
                          translate(V3d(360, 100, 260), rotate(XAxis, 25, cuboid(V3d(-60, -60, -60), V3d(60, 60, 60), white))),
                          translate(V3d(200, 250, 250), rotate(ZAxis, -40, clusterBVH)),
                          rotate(XAxis, 10, translate(V3d(420, 300, 300), rotate(ZAxis, 20, cuboid(V3d(-40, -40, -40), V3d(40, 40, 40), glass)))),
                          lone,
                          constantMedium(0.02, constantColor(0.9, 0.9, 0.9), sphere(V3d(140, 90, 150), 70, white)),
                          constantMedium(0.03, pertext, translate(V3d(420, 80, 120), rotate(ZAxis, 30, cuboid(V3d(-50, -50, -50), V3d(50, 50, 50), white)))),
                          movingSphere(V3d(280, 400, 300), V3d(310, 430, 300), a.t0, a.t1, 35, lambertian(constantColor(0.7, 0.3, 0.1)))});
  // lightsMode 1: without the XY rect, whose (1,0,0) fallback direction makes the mixture pdf 0 on surfaces facing y or z (0/0, Q11)
  s.lights = a.lightsMode == 1 ? makeBVH(gen, {xzLight, sphereLight}) : makeBVH(gen, {xzLight, sphereLight, xyLight});
  s.g1 = gen;
  return s;
}

inline Scene makeSceneByName(const std::string& name, const SceneArgs& a, int64_t seed) {
  SMGen gen = randGen(seed);
  Scene s;
  if (name == "cornell") s = makeCornellBoxScene(a, gen);
  else if (name == "cornell_smoke") s = makeCornellSmokeBoxScene(a, gen);
  else if (name == "simple_light") s = makeSimpleLightScene(a, gen);
  else if (name == "earth") s = makeEarthScene(a, gen);
  else if (name == "two_perlin") s = makeTwoPerlinSpheresScene(a, gen);
  else if (name == "two_spheres") s = makeTwoSpheresScene(a, gen);
  else if (name == "random_book1") s = makeRandomSceneBookOne(a, gen);
  else if (name == "random") s = makeRandomScene(a, gen);
  else if (name == "next_week_final") s = makeNextWeekFinalScene(a, gen);
  else if (name == "synthetic") s = makeSyntheticScene(a, gen);
  else if (name == "quirks") s = makeQuirksScene(a, gen);
  else throw std::runtime_error("unknown scene: " + name);
  s.number();
  return s;
}
inline Camera cameraByName(const std::string& name, int W, int H, int synthN) {
  if (name == "cornell" || name == "cornell_smoke" || name == "quirks") return cornellCamera(W, H);
  if (name == "simple_light" || name == "earth" || name == "two_perlin" || name == "two_spheres") return twoSpheresSceneCamera(W, H);
  if (name == "random_book1" || name == "random") return randomSceneCamera(W, H);
  if (name == "next_week_final") return nextWeekFinalSceneCamera(W, H);
  if (name == "synthetic") return syntheticCamera(W, H, synthN);
  throw std::runtime_error("unknown camera: " + name);
}

}  // namespace orc
