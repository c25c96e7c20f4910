// examples/Main.cpp — app/Main.hs (Main.hs:9-63) written against the C++ mirror of the reference's host API: the same constants, the same
// sequence of calls (randGen 1024 -> cornellCamera -> makeCornellBoxScene -> mkRenderStaticEnv -> runRender -> P3 header + printRow per
// row), with runRender going through the C ABI (include/rt_abi.h) to the B200 kernels instead of the CPU.  What a Haskell maintainer gets
// by importing runRenderB200 as runRender is this program.
//
//   build + run: see tests/test_host_abi.py::test_main_example_builds (g++ ... -lrt_host -lrt_b200), then `examples/Main > cornell.ppm`
#include <cstdio>
#include <exception>
#include <vector>

#include "Lib.hpp"
#include "Scenes.hpp"

// X and Y dimensions of output image, threads, samples, maximum number of reflections (Main.hs:9-24)
static const int defaultImageWidth = 500, defaultImageHeight = 500, defaultnThreads = 2, defaultNs = 1000, defaultMaxDepth = 50;

int main() {
  try {
    const int imageWidth = defaultImageWidth, imageHeight = defaultImageHeight;
    Lib::Arena arena;
    const Lib::RandGen gen = Lib::randGen(1024);  // Fix a seed for comparable performance tests (Main.hs:30)
    const Lib::Camera camera = Scenes::cornellCamera({imageWidth, imageHeight});
    Scenes::SceneAndGen sg = Scenes::makeCornellBoxScene(arena, 0.0, 1.0, gen);
    // Main.hs:49-50 draws one fresh generator per image column; the device RNG is counter-based and consumes only the head of the list
    const std::vector<Lib::RandGen> gens{sg.gen};
    const Lib::RenderStaticEnv staticEnv =
        Lib::mkRenderStaticEnv(arena, sg.scene, camera, {imageWidth, imageHeight}, defaultNs, defaultMaxDepth, defaultnThreads);
    const std::vector<std::vector<Lib::RGB>> vals = Lib::runRender(staticEnv, gens);
    std::printf("P3\n%d %d\n255\n", imageWidth, imageHeight);
    for (int i = 0; i < imageHeight; ++i) Lib::printRow(stdout, imageHeight, i + 1, vals[(size_t)i]);
    std::fputs("\nDone.\n", stderr);
    return 0;
  } catch (const std::exception& e) {
    std::fprintf(stderr, "rt_b200: %s\n", e.what());  // e.g. no CUDA device: the library has no CPU path
    return 1;
  }
}
