// host/Scenes.hpp — C++ mirror of the reference's src/Scenes.hs export list (Scenes.hs:2-18); see Scenes.cpp.
#pragma once
#include <utility>

#include "Lib.hpp"

namespace Scenes {

// The trailing `lit` flag is an extension (see Scenes.cpp); the default is the reference's HEAD behaviour.
struct SceneAndGen { Lib::Scene scene; Lib::RandGen gen; };  // (Scene, RandGen)

Lib::Camera cornellCamera(std::pair<int, int> wh);
Lib::Camera twoSpheresSceneCamera(std::pair<int, int> wh);
Lib::Camera randomSceneCamera(std::pair<int, int> wh);
Lib::Camera nextWeekFinalSceneCamera(std::pair<int, int> wh);
Lib::Camera syntheticCamera(std::pair<int, int> wh, int n);

SceneAndGen makeCornellBoxScene(Lib::Arena& A, Lib::Time t0, Lib::Time t1, Lib::RandGen gen);
SceneAndGen makeCornellSmokeBoxScene(Lib::Arena& A, Lib::Time t0, Lib::Time t1, Lib::RandGen gen);
SceneAndGen makeSimpleLightScene(Lib::Arena& A, Lib::Time t0, Lib::Time t1, Lib::RandGen gen, bool lit = false);
SceneAndGen makeEarthScene(Lib::Arena& A, Lib::Texture earthTex, Lib::Time t0, Lib::Time t1, Lib::RandGen gen, bool lit = false);
SceneAndGen makeTwoPerlinSpheresScene(Lib::Arena& A, Lib::Time t0, Lib::Time t1, Lib::RandGen gen, bool lit = false);
SceneAndGen makeTwoSpheresScene(Lib::Arena& A, Lib::Time t0, Lib::Time t1, Lib::RandGen gen, bool lit = false);
SceneAndGen makeRandomSceneBookOne(Lib::Arena& A, Lib::RandGen gen, bool lit = false);
SceneAndGen makeRandomScene(Lib::Arena& A, Lib::Texture earthtex, Lib::Time t0, Lib::Time t1, Lib::RandGen gen, bool lit = false);
SceneAndGen makeNextWeekFinalScene(Lib::Arena& A, Lib::Texture earthtex, Lib::Time t0, Lib::Time t1, Lib::RandGen gen, bool lit = false);
SceneAndGen makeSyntheticScene(Lib::Arena& A, int n, Lib::RandGen gen);
SceneAndGen makeQuirksScene(Lib::Arena& A, Lib::Time t0, Lib::Time t1, Lib::RandGen gen, bool lit = false);  // API-coverage scene, not in Scenes.hs

}  // namespace Scenes
