// host/capi.cpp — C entry points over the host-side mirror (Lib.hpp / Scenes.hpp) for ctypes callers (tests, bench.py).
// Builds a reference scene by name with the mirrored constructors, flattens it, and hands out the RtSceneDesc that
// goes through the C ABI of librt_b200.so.
#include <cstdlib>
#include <algorithm>
#include <cstring>
#include <memory>
#include <stdexcept>
#include <string>

#include "Scenes.hpp"

using namespace Lib;

namespace {
struct HostHandle {
  Arena arena; Image earth; Scene scene; RandGen g1; std::string name; int synthN = 0;
  Arena::Flat flat; Camera cam; int camW = 0, camH = 0;
};
thread_local std::string g_err;

Camera cameraFor(const std::string& n, int W, int H, int synthN) {
  std::pair<int, int> wh{W, H};
  if (n == "cornell" || n == "cornell_smoke" || n == "quirks") return Scenes::cornellCamera(wh);
  if (n == "simple_light" || n == "earth" || n == "two_perlin" || n == "two_spheres") return Scenes::twoSpheresSceneCamera(wh);
  if (n == "random_book1" || n == "random") return Scenes::randomSceneCamera(wh);
  if (n == "next_week_final") return Scenes::nextWeekFinalSceneCamera(wh);
  if (n == "synthetic") return Scenes::syntheticCamera(wh, synthN);
  throw std::runtime_error("unknown camera: " + n);
}
}  // namespace

extern "C" {

const char* rth_last_error() { return g_err.c_str(); }

// makeBVH of sequences with at least min_objects objects goes to rt_bvh_build (0 = host recursion only); returns the old value
int rth_set_device_bvh(int min_objects) { const int old = (int)Arena::deviceBVHMin; Arena::deviceBVHMin = min_objects > 0 ? (size_t)min_objects : 0; return old; }
double rth_last_device_bvh_ms() { return Arena::lastDeviceBVHms; }

// names: cornell cornell_smoke simple_light earth two_perlin two_spheres random_book1 random next_week_final synthetic quirks
void* rth_scene_create(const char* name, int64_t seed, double t0, double t1, const uint8_t* earth, int ew, int eh, int lit, int synthN) {
  try {
    auto h = std::make_unique<HostHandle>();
    h->name = name; h->synthN = synthN;
    if (earth && ew > 0 && eh > 0) { h->earth.w = (uint32_t)ew; h->earth.h = (uint32_t)eh; h->earth.rgb.assign(earth, earth + (size_t)ew * eh * 3); }
    if (const char* v = std::getenv("RTB_DEVICE_BVH_MIN")) Arena::deviceBVHMin = (size_t)std::max(0, std::atoi(v));
    Arena& A = h->arena;
    RandGen gen = randGen(seed);
    const std::string n = name;
    const bool L = lit != 0;
    Scenes::SceneAndGen sg;
    // earthTexture (Scenes.hs:157-165): decoding earthmap.jpg is the caller's job here; a missing image gives ImageTexture Nothing
    auto earthTex = [&]() { return A.ImageTexture(h->earth.w ? &h->earth : nullptr); };
    if (n == "cornell") sg = Scenes::makeCornellBoxScene(A, t0, t1, gen);
    else if (n == "cornell_smoke") sg = Scenes::makeCornellSmokeBoxScene(A, t0, t1, gen);
    else if (n == "simple_light") sg = Scenes::makeSimpleLightScene(A, t0, t1, gen, L);
    else if (n == "earth") sg = Scenes::makeEarthScene(A, earthTex(), t0, t1, gen, L);
    else if (n == "two_perlin") sg = Scenes::makeTwoPerlinSpheresScene(A, t0, t1, gen, L);
    else if (n == "two_spheres") sg = Scenes::makeTwoSpheresScene(A, t0, t1, gen, L);
    else if (n == "random_book1") sg = Scenes::makeRandomSceneBookOne(A, gen, L);
    else if (n == "random") sg = Scenes::makeRandomScene(A, earthTex(), t0, t1, gen, L);
    else if (n == "next_week_final") sg = Scenes::makeNextWeekFinalScene(A, earthTex(), t0, t1, gen, L);
    else if (n == "synthetic") sg = Scenes::makeSyntheticScene(A, synthN, gen);
    else if (n == "quirks") sg = Scenes::makeQuirksScene(A, t0, t1, gen, L);
    else throw std::runtime_error("unknown scene: " + n);
    h->scene = sg.scene; h->g1 = sg.gen;
    return h.release();
  } catch (const std::exception& e) { g_err = e.what(); return nullptr; }
}
void rth_scene_destroy(void* p) { delete static_cast<HostHandle*>(p); }

// flattens with the scene's camera for a W x H image; the returned desc stays valid until the next call on this handle
const RtSceneDesc* rth_scene_desc(void* p, int W, int H) {
  try {
    auto* h = static_cast<HostHandle*>(p);
    h->cam = cameraFor(h->name, W, H, h->synthN); h->camW = W; h->camH = H;
    h->arena.flatten(h->scene.world, h->scene.lights, h->scene.background, h->cam, h->flat);
    return &h->flat.desc;
  } catch (const std::exception& e) { g_err = e.what(); return nullptr; }
}
void rth_scene_g1(void* p, uint64_t* out2) { auto* h = static_cast<HostHandle*>(p); out2[0] = h->g1.seed; out2[1] = h->g1.gamma; }

// runRender through the mirrored API: mkRenderStaticEnv + runRender (gens = [g1]) -> RGB8 rows, top first
int rth_run_render(void* p, int W, int H, int numSamples, int maxDepth, uint8_t* out_rgb8) {
  try {
    auto* h = static_cast<HostHandle*>(p);
    Camera cam = cameraFor(h->name, W, H, h->synthN);
    RenderStaticEnv env = mkRenderStaticEnv(h->arena, h->scene, cam, {W, H}, numSamples, maxDepth, 2);
    auto rows = runRender(env, {h->g1});
    for (int r = 0; r < H; ++r) for (int x = 0; x < W; ++x) { const RGB& c = rows[r][x]; uint8_t* o = out_rgb8 + ((size_t)r * W + x) * 3; o[0] = c.r; o[1] = c.g; o[2] = c.b; }
    return 0;
  } catch (const std::exception& e) { g_err = e.what(); return -1; }
}

// app/Main.hs:58-62 for an already rendered frame: P3 text (the reference's format) or P6 binary
int rth_write_image(const char* path, const uint8_t* rgb8, int W, int H, int binary) {
  try {
    FILE* f = std::fopen(path, "wb");
    if (!f) throw std::runtime_error(std::string("cannot open ") + path);
    std::vector<std::vector<RGB>> rows((size_t)H, std::vector<RGB>((size_t)W));
    for (int r = 0; r < H; ++r) for (int x = 0; x < W; ++x) { const uint8_t* o = rgb8 + ((size_t)r * W + x) * 3; rows[r][x] = RGB{o[0], o[1], o[2]}; }
    writeImage(f, rows, binary != 0, false);
    std::fclose(f);
    return 0;
  } catch (const std::exception& e) { g_err = e.what(); return -1; }
}

}  // extern "C"
