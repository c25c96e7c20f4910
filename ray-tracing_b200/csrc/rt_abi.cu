// rt_abi.cu — host side of librt_b200.so: the C ABI declared in include/rt_abi.h.
// Validates the flattened scene, packs it into the device layout of rt_device.cuh (fp32 or fp64), launches the
// kernels and moves framebuffers.  There is no CPU rendering path in this library: without a CUDA device every
// compute entry point returns RT_ERR_CUDA.
#include "../../include/rt_abi.h"
#include "rt_device.cuh"
#include "rt_traverse.cuh"  // kBlock, column layout
#include "rt_launch.h"

#include <dlfcn.h>

#include <algorithm>
#include <chrono>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <functional>
#include <map>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

using namespace rtb;

namespace {

thread_local std::string g_err;
int fail(int code, const std::string& msg) { g_err = msg; return code; }
#define CU(call)                                                                                             \
  do {                                                                                                       \
    cudaError_t e_ = (call);                                                                                 \
    if (e_ != cudaSuccess) return fail(RT_ERR_CUDA, std::string(#call) + ": " + cudaGetErrorString(e_));     \
  } while (0)

// deep copy of the caller's RtSceneDesc (pointers are only borrowed during rt_scene_create)
struct HostScene {
  std::vector<double> node_box, sphere_cr, moving, rect, cuboid, translate_off, rotate_sc, medium_nid, mat_param, tex_rgb, tex_scale, perlin_ranvec;
  std::vector<uint32_t> node_left, node_right, node_size, sphere_mat, moving_mat, rect_mat, cuboid_mat, translate_child, rotate_axis, rotate_child,
      medium_mat, medium_boundary, mat_kind, mat_tex, tex_kind, tex_a, tex_b, image_w, image_h;
  std::vector<int32_t> perlin_perm;
  std::vector<std::vector<uint8_t>> images;
  uint32_t n_nodes = 0, n_spheres = 0, n_moving = 0, n_rects = 0, n_cuboids = 0, n_translates = 0, n_rotates = 0, n_media = 0, n_materials = 0,
           n_textures = 0, n_perlin = 0, n_images = 0;
  uint32_t world_root = 0, lights_root = 0;
  double background[3] = {0, 0, 0}, camera[24] = {0};
  std::vector<uint32_t> node_lsize, node_rsize;  // htblSize of the children (Lib.hs:662-671)
  int stackDepth = 1;
  uint32_t lightCap = 0;  // entries of the light-pdf walk stack (rt_traverse.cuh)
  // scan stream of the warp-cooperative traversal (rt_coop.cuh): eligible when the world is a BVH whose leaves are spheres / rects /
  // Unhittable only and the light list is one primitive or Unhittable; coopRecords = records of the pre-order expansion
  bool coopEligible = false;
  uint64_t coopRecords = 0;
};

template <class T> struct DevBuffers {
  void* blob = nullptr;  // ONE device allocation holds every array of the scene (256-byte aligned sub-ranges)
  DScene<T> sc{};
  bool ready = false;
  ~DevBuffers() { if (blob) cudaFree(blob); }
};

// The wavefront's path-state slots (~360 MB on a B200) outlive their scene in a small per-device pool: a host that creates a scene per
// frame (bench.py's e2e loop, a progressive viewer) otherwise pays a cudaMalloc + cudaFree of the blob per frame.  Blobs are handed out
// whole and exclusively (a DeviceCtx owns what it took until it is destroyed), at most kKeep per device are kept, rt_shutdown frees them.
struct ScratchPool {
  struct Item { int device; void* p; size_t bytes; };
  static constexpr size_t kKeep = 2;
  std::mutex mu;
  std::vector<Item> items;
  void* take(int device, size_t need, size_t* got) {
    std::lock_guard<std::mutex> lk(mu);
    int best = -1;
    for (size_t i = 0; i < items.size(); ++i)
      if (items[i].device == device && items[i].bytes >= need && (best < 0 || items[i].bytes < items[(size_t)best].bytes)) best = (int)i;
    if (best < 0) return nullptr;
    void* p = items[(size_t)best].p; *got = items[(size_t)best].bytes;
    items.erase(items.begin() + best);
    return p;
  }
  void give(int device, void* p, size_t bytes) {
    std::lock_guard<std::mutex> lk(mu);
    size_t have = 0;
    for (const Item& it : items) have += it.device == device;
    if (have >= kKeep) { cudaFree(p); return; }
    items.push_back(Item{device, p, bytes});
  }
  void clear() {
    std::lock_guard<std::mutex> lk(mu);
    for (const Item& it : items) cudaFree(it.p);
    items.clear();
  }
};
ScratchPool g_scratch;  // defined before the scene table: destroyed after it at exit (nothing is freed then, the context is going away)

// everything a scene owns on ONE device: the packed scene (fp32 eagerly, fp64 on first use), the work counters, a framebuffer
struct DeviceCtx {
  int device = 0, numSMs = 148;
  DevBuffers<float> f32;
  DevBuffers<double> f64;
  unsigned long long* d_counter = nullptr;  // kCtrTotal words: [0] path counter, work counters, scheduler statistics, tail stamps (rt_device.cuh kCtr*)
  double* d_fb = nullptr; size_t fbElems = 0;
  void* wfBlob = nullptr; size_t wfBytes = 0; WfState wf{};   // wavefront path-state slots (rt_wavefront.cuh), grown on demand
  unsigned long long* h_wfCtl = nullptr;    // pinned: {frame finished, next free path index} read back between batches of rounds
  int wfRounds = 0;                         // rounds launched for the call in flight
  RenderArgs wfArgs{}; const void* wfScene = nullptr; int wfPrec = 0; bool wfPending = false;  // the wavefront frame in flight
  cudaStream_t wfStream = nullptr; cudaEvent_t wfProbe = nullptr;
  cudaStream_t stream = nullptr;            // multi-GPU shares only (the single-device entry points use the caller's / default stream)
  cudaEvent_t evStart = nullptr, evDone = nullptr;
  ~DeviceCtx() {
    if (d_counter) cudaFree(d_counter);
    if (d_fb) cudaFree(d_fb);
    if (wfBlob) {  // cudaFree would have waited for the device; a pooled blob may be handed to another scene at once, so wait here
      int cur = 0; cudaGetDevice(&cur); cudaSetDevice(device); cudaDeviceSynchronize(); cudaSetDevice(cur);
      g_scratch.give(device, wfBlob, wfBytes);
    }
    if (h_wfCtl) cudaFreeHost(h_wfCtl);
    if (wfProbe) cudaEventDestroy(wfProbe);
    if (evStart) cudaEventDestroy(evStart);
    if (evDone) cudaEventDestroy(evDone);
    if (stream) cudaStreamDestroy(stream);
  }
};

struct SceneEntry {
  std::mutex mu;  // calls on one handle are serialised: they share the device scene, the work counter and the staging buffers
  HostScene hs;
  DeviceCtx main;                                         // the device given to rt_scene_create
  std::map<int, std::unique_ptr<DeviceCtx>> replicas;     // other devices, filled by multi-GPU rt_render calls
  double* h_fb = nullptr; size_t hfbElems = 0;  // pinned staging
  ~SceneEntry() { if (h_fb) cudaFreeHost(h_fb); }
};

// library-wide multi-GPU state (rt_init)
struct MultiGpu {
  std::mutex mu;
  int defaultGpus = 1;
  bool wantNccl = false;
  // NCCL, loaded at run time so that the library has no link-time dependency on it
  void* lib = nullptr;
  std::vector<void*> comms;  // ncclComm_t per device ordinal 0..n-1
  int (*commInitAll)(void**, int, const int*) = nullptr;
  int (*commDestroy)(void*) = nullptr;
  int (*reduce)(const void*, void*, size_t, int, int, int, void*, cudaStream_t) = nullptr;
  int (*groupStart)() = nullptr;
  int (*groupEnd)() = nullptr;
  const char* (*errorString)(int) = nullptr;
} g_multi;

std::mutex g_mu;
std::map<uint64_t, std::shared_ptr<SceneEntry>> g_scenes;
uint64_t g_nextId = 1;

template <class V, class P> void copyIn(V& dst, const P* src, size_t n) {
  dst.assign(n, typename V::value_type{});
  if (n) std::memcpy(dst.data(), src, n * sizeof(typename V::value_type));
}

bool refOk(const HostScene& h, uint32_t ref) {
  uint32_t k = ref >> 28, i = ref & 0x0fffffffu;  // full 28-bit field: counts are limited to 27 bits, so a set bit 27 is out of range
  switch (k) {
    case K_UNHITTABLE: return true;
    case K_SPHERE: return i < h.n_spheres;
    case K_MOVING: return i < h.n_moving;
    case K_RECT_XY: case K_RECT_XZ: case K_RECT_YZ: return i < h.n_rects;
    case K_CUBOID: return i < h.n_cuboids;
    case K_BVH: return i < h.n_nodes;
    case K_TRANSLATE: return i < h.n_translates;
    case K_ROTATE: return i < h.n_rotates;
    case K_MEDIUM: return i < h.n_media;
    default: return false;
  }
}

// Structural analysis of the hittable DAG: traversal stack need (push/pop discipline of rtb::traverse), open
// frames, medium nesting, htblSize; detects cycles.  Memoised per ref.
struct Analysis {
  // per hittable ref, memoised over the DAG: traversal-stack entries needed, open frames, nested media, htblSize.  Flat tables
  // indexed by a dense id (a 1 M-sphere scene is 3 M refs: std::map tables made rt_scene_create the slowest part of an e2e step)
  const HostScene& h;
  size_t base[16];
  std::vector<int> stackNeedV, sizeV;
  std::vector<uint8_t> framesV, mediaV, state;  // state: 0 new, 1 in progress, 2 done
  bool cycle = false;
  int maxStack = 1, maxFrames = 0, maxMedia = 0;
  explicit Analysis(const HostScene& hs) : h(hs) {
    const size_t cnt[16] = {1, hs.n_spheres, hs.n_moving, hs.n_rects, hs.n_rects, hs.n_rects, hs.n_cuboids, hs.n_nodes, hs.n_translates, hs.n_rotates, hs.n_media, 0, 0, 0, 0, 0};
    size_t tot = 0;
    for (int k = 0; k < 16; ++k) { base[k] = tot; tot += cnt[k]; }
    stackNeedV.assign(tot, 0); sizeV.assign(tot, 0); framesV.assign(tot, 0); mediaV.assign(tot, 0); state.assign(tot, 0);
  }
  size_t id(uint32_t ref) const { return base[ref >> 28] + ((ref >> 28) == K_UNHITTABLE ? 0 : (ref & kIdxMask)); }  // refs were range-checked (refOk)
  int size(uint32_t ref) const { return sizeV[id(ref)]; }
  void go(uint32_t ref) {
    if (cycle) return;
    const size_t me = id(ref);
    if (state[me] == 2) return;
    if (state[me] == 1) { cycle = true; return; }
    state[me] = 1;
    uint32_t k = ref >> 28, i = ref & kIdxMask;
    int sn = 0, fr = 0, md = 0, sz = 1;
    if (k == K_BVH) {
      uint32_t l = h.node_left[i], r = h.node_right[i];
      go(l); go(r);
      if (cycle) return;
      const size_t il = id(l), ir = id(r);
      sn = std::max(2, std::max(1 + stackNeedV[il], stackNeedV[ir]));
      fr = std::max(framesV[il], framesV[ir]); md = std::max(mediaV[il], mediaV[ir]); sz = (int)h.node_size[i];
    } else if (k == K_TRANSLATE || k == K_ROTATE) {
      uint32_t c = k == K_TRANSLATE ? h.translate_child[i] : h.rotate_child[i];
      go(c);
      if (cycle) return;
      const size_t ic = id(c);
      sn = std::max(2, 1 + stackNeedV[ic]); fr = 1 + framesV[ic]; md = mediaV[ic]; sz = sizeV[ic];
    } else if (k == K_MEDIUM) {
      uint32_t c = h.medium_boundary[i];
      go(c);
      if (cycle) return;
      const size_t ic = id(c);
      sn = std::max(2, 1 + stackNeedV[ic]); fr = framesV[ic]; md = 1 + mediaV[ic]; sz = 1;
    } else if (k == K_UNHITTABLE) {
      sz = 0;
    }
    stackNeedV[me] = sn; framesV[me] = (uint8_t)std::min(fr, 255); mediaV[me] = (uint8_t)std::min(md, 255); sizeV[me] = sz;
    maxStack = std::max(maxStack, sn); maxFrames = std::max(maxFrames, fr); maxMedia = std::max(maxMedia, md);
    state[me] = 2;
  }
};
int lightDepth(const HostScene& h, uint32_t ref, int d = 0) {
  if (d > 64) return d;
  if ((ref >> 28) != K_BVH) return 1;
  uint32_t i = ref & kIdxMask;
  return 1 + std::max(lightDepth(h, h.node_left[i], d + 1), lightDepth(h, h.node_right[i], d + 1));
}

template <class T> T fromBits(uint32_t b);
template <> float fromBits<float>(uint32_t b) { float f; std::memcpy(&f, &b, 4); return f; }
template <> double fromBits<double>(uint32_t b) { uint64_t w = b; double d; std::memcpy(&d, &w, 8); return d; }

// Scene upload: every array is laid out in ONE pinned staging buffer (kept for the life of the library, grown on demand) and goes
// to the device as ONE allocation and ONE cudaMemcpyAsync (round 1 did a cudaMalloc + a pageable cudaMemcpy per array, ~16 per scene).
struct PinnedStage {
  std::mutex mu; char* p = nullptr; size_t cap = 0;
  ~PinnedStage() { if (p) cudaFreeHost(p); }
} g_stage;
// host-side packing of large scenes (10^6 spheres: ~130 MB of records) over a few threads; small scenes stay on the calling thread
template <class F> void parallelFor(size_t n, F f, size_t kMin = 1u << 16) {
  const unsigned hw = std::max(1u, std::min(8u, std::thread::hardware_concurrency()));
  const size_t parts = std::min<size_t>(hw, n / kMin);
  if (parts <= 1) { f((size_t)0, n); return; }
  std::vector<std::thread> th;
  for (size_t p = 1; p < parts; ++p) th.emplace_back([=] { f(n * p / parts, n * (p + 1) / parts); });
  f((size_t)0, n / parts);
  for (auto& t : th) t.join();
}

struct BlobLayout {
  struct Piece { size_t off; const void* src; size_t bytes; };
  std::vector<Piece> pieces; size_t total = 0;
  size_t add(const void* src, size_t bytes) {
    const size_t off = total;
    if (bytes) pieces.push_back({off, src, bytes});
    total += (bytes + 255) & ~(size_t)255;
    return off;
  }
  template <class V> size_t add(const std::vector<V>& v) { return add(v.data(), v.size() * sizeof(V)); }
};

// which leaf classes are tested as pairs (bit 0 sphere-like, bit 1 rect / cuboid): sphere pairs always pay; rect / cuboid pairs pay
// in cuboid-heavy scenes (cfg3 +3 %) and cost in rect scenes (cfg4 -9 %).  RTB_PAIR overrides for experiments.
static int pairMask(const HostScene& h) {
  const char* v = std::getenv("RTB_PAIR");
  return v ? std::atoi(v) : (h.n_cuboids > h.n_rects ? 3 : 1);
}

// pack the flattened scene into the device layout (rt_device.cuh DScene<T>); doubles are rounded to T here
template <class T> int buildDevice(const HostScene& h, DevBuffers<T>& db) {
  if (db.ready) return RT_OK;
  using Q = Q4<T>;
  auto B = [](uint32_t b) { return fromBits<T>(b); };
  std::vector<Q> nodes(2 * (size_t)h.n_nodes), spheres(h.n_spheres), moving(3 * (size_t)h.n_moving), rects(2 * (size_t)h.n_rects), cuboids(2 * (size_t)h.n_cuboids),
      translates(h.n_translates), rotates(h.n_rotates), media(h.n_media), mats(h.n_materials), texs(2 * (size_t)h.n_textures), pvec(256 * (size_t)h.n_perlin);
  std::vector<uint4> nsz(h.n_nodes);
  std::vector<uint8_t> pperm(768 * (size_t)h.n_perlin);
  const int pmask = pairMask(h);
  parallelFor(h.n_nodes, [&](size_t i0, size_t i1) {
  for (size_t i = i0; i < i1; ++i) {
    const double* b = &h.node_box[i * 6];
    // both children leaves of one step class (sphere-like, or rect / cuboid): mark the pair (rt_device.cuh kPairBit)
    auto leafClass = [](uint32_t r) { uint32_t k = r >> 28; return (k == K_SPHERE || k == K_MOVING) ? 1 : ((k >= K_RECT_XY && k <= K_CUBOID) ? 2 : 0); };
    const int lcl = leafClass(h.node_left[i]);
    const uint32_t pair = (lcl != 0 && ((pmask >> (lcl - 1)) & 1) && lcl == leafClass(h.node_right[i])) ? kPairBit : 0u;
    nodes[2 * i] = Q{(T)b[0], (T)b[1], (T)b[2], B(h.node_left[i] | pair)};
    nodes[2 * i + 1] = Q{(T)b[3], (T)b[4], (T)b[5], B(h.node_right[i])};
    nsz[i] = make_uint4(h.node_size[i], h.node_lsize[i], h.node_rsize[i], 0);
  }
  });
  parallelFor(h.n_spheres, [&](size_t i0, size_t i1) {
    for (size_t i = i0; i < i1; ++i) { const double* sp = &h.sphere_cr[i * 4]; spheres[i] = Q{(T)sp[0], (T)sp[1], (T)sp[2], (T)sp[3]}; }
  });
  for (size_t i = 0; i < h.n_moving; ++i) {
    const double* m = &h.moving[i * 9];
    moving[3 * i] = Q{(T)m[0], (T)m[1], (T)m[2], (T)m[8]};
    moving[3 * i + 1] = Q{(T)m[3], (T)m[4], (T)m[5], (T)m[6]};
    // exact reciprocal of a power-of-two duration (movingFrac, rt_device.cuh): mantissa 0.5 in T's own rounding, and 1 / d representable
    const T dur = (T)m[7];
    int e2 = 0;
    const bool pow2 = std::isfinite((double)dur) && dur != (T)0 && std::fabs(std::frexp((double)dur, &e2)) == 0.5 && e2 > -100 && e2 < 100;
    moving[3 * i + 2] = Q{dur, B(h.moving_mat[i]), pow2 ? (T)1 / dur : (T)0, (T)0};
  }
  for (size_t i = 0; i < h.n_rects; ++i) {
    const double* r = &h.rect[i * 5];
    rects[2 * i] = Q{(T)r[0], (T)r[1], (T)r[2], (T)r[3]};
    rects[2 * i + 1] = Q{(T)r[4], B(h.rect_mat[i]), (T)0, (T)0};
  }
  for (size_t i = 0; i < h.n_cuboids; ++i) {
    const double* c = &h.cuboid[i * 6];
    cuboids[2 * i] = Q{(T)c[0], (T)c[1], (T)c[2], B(h.cuboid_mat[i])};
    cuboids[2 * i + 1] = Q{(T)c[3], (T)c[4], (T)c[5], (T)0};
  }
  for (size_t i = 0; i < h.n_translates; ++i) { const double* t = &h.translate_off[i * 3]; translates[i] = Q{(T)t[0], (T)t[1], (T)t[2], B(h.translate_child[i])}; }
  for (size_t i = 0; i < h.n_rotates; ++i) rotates[i] = Q{(T)h.rotate_sc[i * 2], (T)h.rotate_sc[i * 2 + 1], B(h.rotate_axis[i]), B(h.rotate_child[i])};
  for (size_t i = 0; i < h.n_media; ++i) media[i] = Q{(T)h.medium_nid[i], B(h.medium_mat[i]), B(h.medium_boundary[i]), (T)0};
  for (size_t i = 0; i < h.n_materials; ++i) mats[i] = Q{B(h.mat_kind[i]), B(h.mat_tex[i]), (T)h.mat_param[i], (T)0};
  for (size_t i = 0; i < h.n_textures; ++i) {
    texs[2 * i] = Q{(T)h.tex_rgb[i * 3], (T)h.tex_rgb[i * 3 + 1], (T)h.tex_rgb[i * 3 + 2], (T)h.tex_scale[i]};
    texs[2 * i + 1] = Q{B(h.tex_kind[i]), B(h.tex_a[i]), B(h.tex_b[i]), (T)0};
  }
  for (size_t i = 0; i < (size_t)h.n_perlin * 256; ++i) pvec[i] = Q{(T)h.perlin_ranvec[i * 3], (T)h.perlin_ranvec[i * 3 + 1], (T)h.perlin_ranvec[i * 3 + 2], (T)0};
  for (size_t i = 0; i < (size_t)h.n_perlin * 768; ++i) pperm[i] = (uint8_t)(h.perlin_perm[i] & 255);

  // scan stream (rt_coop.cuh): the world BVH in reference visit order — a node, its whole left subtree, its right subtree
  // (Lib.hs:971-988) — every node carrying the index of the first record behind its subtree
  std::vector<Q> coop;
  if (h.coopEligible) {
    coop.resize(2 * (size_t)h.coopRecords);
    size_t at = 0;  // records written
    // explicit stack instead of recursion: an entry is a ref to emit, or (high bit set) a node record waiting for the end of its subtree
    std::vector<uint64_t> todo;
    todo.push_back(h.world_root);
    const uint64_t kClose = 1ull << 63;
    while (!todo.empty()) {
      const uint64_t e = todo.back(); todo.pop_back();
      if (e & kClose) { coop[2 * (size_t)(e & ~kClose)].w = B((uint32_t)at); continue; }  // first record behind the subtree
      const uint32_t ref = (uint32_t)e, k = ref >> 28, i = ref & kIdxMask;
      Q* r = &coop[2 * at];
      if (k == K_BVH) {
        const double* b = &h.node_box[(size_t)i * 6];
        r[0] = Q{(T)b[0], (T)b[1], (T)b[2], B(0u)};
        r[1] = Q{(T)b[3], (T)b[4], (T)b[5], B(ref)};
        todo.push_back(kClose | (uint64_t)at);
        todo.push_back(h.node_right[i]); todo.push_back(h.node_left[i]);  // left subtree first (Lib.hs:971-988)
      } else if (k == K_SPHERE) {
        const double* sp = &h.sphere_cr[(size_t)i * 4];
        r[0] = Q{(T)sp[0], (T)sp[1], (T)sp[2], (T)sp[3]};
        r[1] = Q{(T)0, (T)0, (T)0, B(ref)};
      } else if (k >= K_RECT_XY && k <= K_RECT_YZ) {
        const double* rc = &h.rect[(size_t)i * 5];
        r[0] = Q{(T)rc[0], (T)rc[1], (T)rc[2], (T)rc[3]};
        r[1] = Q{(T)rc[4], (T)0, (T)0, B(ref)};
      } else {  // Unhittable
        r[0] = Q{(T)0, (T)0, (T)0, (T)0};
        r[1] = Q{(T)0, (T)0, (T)0, B(ref)};
      }
      ++at;
    }
  }
  DScene<T>& sc = db.sc;
  BlobLayout L;
  const size_t oCoop = L.add(coop);
  const size_t oNodes = L.add(nodes), oNsz = L.add(nsz), oSph = L.add(spheres), oSmat = L.add(h.sphere_mat), oMov = L.add(moving), oRect = L.add(rects),
               oCub = L.add(cuboids), oTr = L.add(translates), oRot = L.add(rotates), oMed = L.add(media), oMat = L.add(mats), oTex = L.add(texs),
               oPv = L.add(pvec), oPp = L.add(pperm);
  std::vector<DImage> imgs(h.n_images);
  std::vector<size_t> oImg(h.n_images);
  for (size_t i = 0; i < h.n_images; ++i) oImg[i] = L.add(nullptr, (size_t)h.image_w[i] * h.image_h[i] * sizeof(uchar4));
  const size_t oImgTab = L.add(nullptr, imgs.size() * sizeof(DImage));
  const size_t total = std::max<size_t>(L.total, 256);
  CU(cudaMalloc(&db.blob, total));
  char* base = static_cast<char*>(db.blob);
  {
    std::lock_guard<std::mutex> lk(g_stage.mu);
    if (g_stage.cap < total) {
      if (g_stage.p) cudaFreeHost(g_stage.p);
      g_stage.p = nullptr; g_stage.cap = 0;
      CU(cudaHostAlloc(reinterpret_cast<void**>(&g_stage.p), total, cudaHostAllocDefault));
      g_stage.cap = total;
    }
    for (const auto& pc : L.pieces) if (pc.src)
      parallelFor(pc.bytes, [&](size_t b0, size_t b1) { std::memcpy(g_stage.p + pc.off + b0, static_cast<const char*>(pc.src) + b0, b1 - b0); }, (size_t)4 << 20);
    for (size_t i = 0; i < h.n_images; ++i) {  // RGB8 -> uchar4 directly in the staging buffer
      uchar4* px = reinterpret_cast<uchar4*>(g_stage.p + oImg[i]);
      const size_t npx = (size_t)h.image_w[i] * h.image_h[i];
      for (size_t q = 0; q < npx; ++q) px[q] = make_uchar4(h.images[i][q * 3], h.images[i][q * 3 + 1], h.images[i][q * 3 + 2], 255);
      imgs[i] = DImage{reinterpret_cast<const uchar4*>(base + oImg[i]), h.image_w[i], h.image_h[i], 0};
    }
    if (!imgs.empty()) std::memcpy(g_stage.p + oImgTab, imgs.data(), imgs.size() * sizeof(DImage));
    CU(cudaMemcpyAsync(base, g_stage.p, L.total, cudaMemcpyHostToDevice, 0));
    CU(cudaStreamSynchronize(0));
  }
  auto at = [&](size_t off, size_t count) -> const void* { return count ? base + off : nullptr; };
  sc.nodes = static_cast<const Q*>(at(oNodes, nodes.size())); sc.node_sz = static_cast<const uint4*>(at(oNsz, nsz.size()));
  sc.spheres = static_cast<const Q*>(at(oSph, spheres.size())); sc.sphere_mat = static_cast<const uint32_t*>(at(oSmat, h.sphere_mat.size()));
  sc.moving = static_cast<const Q*>(at(oMov, moving.size())); sc.rects = static_cast<const Q*>(at(oRect, rects.size()));
  sc.cuboids = static_cast<const Q*>(at(oCub, cuboids.size())); sc.translates = static_cast<const Q*>(at(oTr, translates.size()));
  sc.rotates = static_cast<const Q*>(at(oRot, rotates.size())); sc.media = static_cast<const Q*>(at(oMed, media.size()));
  sc.mats = static_cast<const Q*>(at(oMat, mats.size())); sc.texs = static_cast<const Q*>(at(oTex, texs.size()));
  sc.perlin_vec = static_cast<const Q*>(at(oPv, pvec.size())); sc.perlin_perm = static_cast<const uint8_t*>(at(oPp, pperm.size()));
  sc.images = static_cast<const DImage*>(at(oImgTab, imgs.size()));
  sc.coop = static_cast<const Q*>(at(oCoop, coop.size())); sc.coop_n = (uint32_t)(coop.size() / 2); sc.pad2_ = 0;
  sc.world_root = h.world_root; sc.lights_root = h.lights_root;
  sc.bigCoords = 0; sc.pad_ = 0;
  for (double v : h.node_box) if (!(std::fabs(v) <= 1099511627776.0)) sc.bigCoords = 1;  // beyond 2^40 or non-finite (rt_device.cuh kDivHi)
  for (double v : h.rect) if (!(std::fabs(v) <= 1099511627776.0)) sc.bigCoords = 1;      // plane quotients use the same hoisted reciprocal (rectTHoisted)
  for (double v : h.cuboid) if (!(std::fabs(v) <= 1099511627776.0)) sc.bigCoords = 1;
  for (int i = 0; i < 3; ++i) sc.bg[i] = (T)h.background[i];
  for (int i = 0; i < 24; ++i) sc.cam[i] = (T)h.camera[i];
  db.ready = true;
  return RT_OK;
}

// the entry stays alive while a call holds the pointer, even if another thread destroys the handle meanwhile
std::shared_ptr<SceneEntry> lookup(rt_scene_t id) {
  std::lock_guard<std::mutex> lk(g_mu);
  auto it = g_scenes.find(id);
  return it == g_scenes.end() ? nullptr : it->second;
}

// CUDA events that do not outlive an early error return
struct Events {
  cudaEvent_t ev[3] = {nullptr, nullptr, nullptr};
  int n = 0;
  cudaError_t create(int count) { for (; n < count; ++n) { cudaError_t e = cudaEventCreate(&ev[n]); if (e != cudaSuccess) return e; } return cudaSuccess; }
  ~Events() { for (int i = 0; i < n; ++i) cudaEventDestroy(ev[i]); }
};

int deviceScene(const HostScene& hs, DeviceCtx* c, uint32_t precision, const void** out) {
  int rc;
  if (precision == RT_PREC_F64) { if ((rc = buildDevice<double>(hs, c->f64))) return rc; *out = &c->f64.sc; }
  else if (precision == RT_PREC_F32 || precision == RT_PREC_MIXED) { if ((rc = buildDevice<float>(hs, c->f32))) return rc; *out = &c->f32.sc; }
  else return fail(RT_ERR_ARG, "bad precision");
  return RT_OK;
}

int normaliseParams(const RtRenderParams* p, RtRenderParams& q) {
  if (!p) return fail(RT_ERR_ARG, "params is null");
  q = *p;
  if (q.width == 0 || q.height == 0) return fail(RT_ERR_ARG, "width/height must be > 0");
  if ((uint64_t)q.width * q.height > 0x0fffffffu) return fail(RT_ERR_ARG, "image too large");
  if (q.row0 == 0 && q.row1 == 0) q.row1 = q.height;
  if (q.col0 == 0 && q.col1 == 0) q.col1 = q.width;
  if (q.row1 > q.height || q.col1 > q.width || q.row0 >= q.row1 || q.col0 >= q.col1) return fail(RT_ERR_ARG, "bad region");
  if (q.precision > RT_PREC_F64) return fail(RT_ERR_ARG, "bad precision");
  if ((q.flags & RT_FLAG_FAST_MATH) && (q.flags & RT_FLAG_STRICT_MATH)) return fail(RT_ERR_ARG, "RT_FLAG_FAST_MATH and RT_FLAG_STRICT_MATH exclude each other");
  return RT_OK;
}

// scheduler knobs of renderKernel; environment overrides are for tuning experiments only
uint32_t tuneParam(const char* name, uint32_t dflt) {
  const char* v = std::getenv(name);
  return v ? (uint32_t)std::strtoul(v, nullptr, 10) : dflt;
}

struct PathLog { uint32_t* prim = nullptr; double* t = nullptr; uint32_t stride = 0; };

// launches the render kernel of one device into d_out (accumulating).  instr: 0 production, 1 counters, 3 path log
// scan-stream traversal: forced by RT_FLAG_COOP, forbidden by RT_FLAG_NO_COOP / fast math, otherwise chosen for eligible scenes whose
// BVH is large enough that one path is a long dependent chain (RTB_COOP_MIN_NODES, default 32768 nodes)
bool useCoop(const HostScene& hs, uint32_t flags) {
  if (!hs.coopEligible || (flags & (RT_FLAG_NO_COOP | RT_FLAG_FAST_MATH))) return false;
  if (flags & RT_FLAG_COOP) return true;
  return hs.n_nodes >= tuneParam("RTB_COOP_MIN_NODES", 32768);
}

// split traversal / shade kernels: forced by RT_FLAG_WAVEFRONT; otherwise (RTB_WAVEFRONT=auto|0|1, default auto) for scenes whose
// shading dominates — a light list to sample (every Lambertian bounce runs the mixture pdf) and a small BVH
bool wavefrontCapable(const RtRenderParams& q, int instr) {
  return instr == 0 && (q.precision == RT_PREC_F32 || q.precision == RT_PREC_MIXED) && !(q.flags & (RT_FLAG_FAST_MATH | RT_FLAG_STRICT_MATH));
}
bool useWavefront(const HostScene& hs, const RtRenderParams& q, int instr) {
  if (!wavefrontCapable(q, instr) || (q.flags & RT_FLAG_NO_WAVEFRONT)) return false;
  if (q.flags & RT_FLAG_WAVEFRONT) return true;
  const char* v = std::getenv("RTB_WAVEFRONT");
  if (v && v[0] == '0') return false;
  if (v && v[0] == '1') return true;
  return (hs.lights_root >> 28) != K_UNHITTABLE && hs.n_nodes <= 64 && hs.n_media == 0;
}

// more rounds of the wavefront until the frame-finished flag is set (blocks the host; the first batch was queued by renderLaunch)
// A wavefront frame in flight on one device: batches of rounds are queued with a PROBE in their middle (a copy of {done, next free
// path index} to pinned memory + an event), so that when the probe fires the second half of the batch is still queued and the device
// never waits for the host.  wavefrontPoll never blocks; wavefrontFinishAll polls all the devices of a call round-robin.
int wavefrontQueue(const HostScene& hs, DeviceCtx* c, int rounds, bool init) {
  CU(cudaSetDevice(c->device));  // polled round-robin over the devices of a multi-GPU call
  const int first = std::max(1, rounds / 2);
  cudaError_t ce = launchWavefront(c->wfPrec, c->wfScene, c->wfArgs, c->wf, hs.stackDepth, c->numSMs, first, init, c->wfStream, nullptr);
  if (ce == cudaSuccess) ce = cudaMemcpyAsync(c->h_wfCtl, c->wf.ctl + 2, 2 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, c->wfStream);
  if (ce == cudaSuccess) ce = cudaEventRecord(c->wfProbe, c->wfStream);
  if (ce == cudaSuccess && rounds - first > 0) ce = launchWavefront(c->wfPrec, c->wfScene, c->wfArgs, c->wf, hs.stackDepth, c->numSMs, rounds - first, false, c->wfStream, nullptr);
  if (ce != cudaSuccess) return fail(RT_ERR_CUDA, std::string("wavefront launch: ") + cudaGetErrorString(ce));
  c->wfRounds += rounds;
  return RT_OK;
}
// -> 1 finished, 0 still running (more rounds were queued if needed), < 0 error
int wavefrontPoll(const HostScene& hs, DeviceCtx* c) {
  const cudaError_t qe = cudaEventQuery(c->wfProbe);
  if (qe == cudaErrorNotReady) return 0;
  if (qe != cudaSuccess) return fail(RT_ERR_CUDA, std::string("wavefront probe: ") + cudaGetErrorString(qe));
  if (c->h_wfCtl[0]) return 1;
  const RenderArgs& a = c->wfArgs;
  // every round finishes at least one bounce of every live path: a frame needs at most paths * (maxDepth + 1) / slots rounds plus the tail of one chunk
  const unsigned long long bound = a.total_paths / std::max(1u, c->wf.P) * (a.maxDepth + 1ull) * 2ull + 2ull * a.wfChunk * (a.maxDepth + 2ull) + 64ull;
  if ((unsigned long long)c->wfRounds > bound) return fail(RT_ERR_CUDA, "wavefront: the frame did not finish within its round bound (internal error)");
  // size the next batch from the progress at the probe: rounds per handed-out path so far, then the drain of the last paths
  const double handed = (double)std::min<unsigned long long>(c->h_wfCtl[1], a.total_paths);
  const double frac = handed / (double)a.total_paths;
  int more = 32;
  if (frac > 0.02 && frac < 1.0) more = (int)std::min(2048.0, (double)c->wfRounds * (1.0 - frac) / frac * 0.6 + 16.0);
  const int rc = wavefrontQueue(hs, c, more, false);
  return rc ? rc : 0;
}
int wavefrontFinishAll(const HostScene& hs, DeviceCtx* const* ctx, int n) {
  for (;;) {
    bool pending = false, progressed = false;
    for (int k = 0; k < n; ++k) {
      if (!ctx[k]->wfPending) continue;
      const int r = wavefrontPoll(hs, ctx[k]);
      if (r < 0) return r;
      if (r == 1) { ctx[k]->wfPending = false; progressed = true; } else pending = true;
    }
    if (!pending) return RT_OK;
    if (!progressed) std::this_thread::yield();
  }
}

int renderLaunch(const HostScene& hs, DeviceCtx* c, const RtRenderParams& q, double* d_out, cudaStream_t st, int instr, int* gridOut, const PathLog* plog = nullptr,
                 uint32_t* kindOut = nullptr, bool wfDefer = false, uint32_t rowStep = 1) {
  if ((q.flags & RT_FLAG_WAVEFRONT) && !wavefrontCapable(q, instr))
    return fail(RT_ERR_UNSUPPORTED, "RT_FLAG_WAVEFRONT: fp32 / mixed arithmetic with literal math and no instrumentation only");
  if ((q.flags & RT_FLAG_COOP) && !hs.coopEligible)
    return fail(RT_ERR_UNSUPPORTED, "RT_FLAG_COOP: the scene is not eligible for the scan-stream traversal (world must be a BVH of spheres / rects, light list one primitive)");
  const void* dsc = nullptr;
  int rc = deviceScene(hs, c, q.precision, &dsc);
  if (rc) return rc;
  if (!c->d_counter) { CU(cudaMalloc(&c->d_counter, kCtrTotal * sizeof(unsigned long long))); }
  CU(cudaMemsetAsync(c->d_counter, 0, kCtrTotal * sizeof(unsigned long long), st));
  if (q.sample_count == 0) return RT_OK;
  RenderArgs a{};
  a.W = q.width; a.H = q.height; a.s0 = q.sample_offset; a.ns = q.sample_count; a.maxDepth = q.max_depth;
  a.row0 = q.row0; a.row1 = q.row1; a.col0 = q.col0; a.col1 = q.col1; a.rowStep = std::max(1u, rowStep);
  a.seed = q.seed;
  a.total_paths = (unsigned long long)((q.row1 - q.row0 + a.rowStep - 1) / a.rowStep) * (q.col1 - q.col0) * q.sample_count;
  // chunk: ~1024 paths, halved until there are at least 8 chunks per warp slot for balance, down to 4 paths: launches with few, very
  // long paths (the cfg5 sample: seconds per path) are tail-bound, and a warp that refills 4 lanes at a time keeps its lanes busier than
  // one that waits for a 32-path chunk to drain (tools/chunk_probe.py: 90 k spheres, 480x270x1: 1424 -> 1242 ms; results are identical)
  unsigned long long warps = (unsigned long long)c->numSMs * 16ull;
  unsigned long long chunk = std::max(1u, tuneParam("RTB_CHUNK_MAX", 1024));
  const unsigned long long chunkMin = std::max(1u, tuneParam("RTB_CHUNK_MIN", 4));
  while (chunk > chunkMin && a.total_paths / chunk < warps * 8ull) chunk /= 2;
  a.chunk = (uint32_t)chunk;
  a.chunkMin = (uint32_t)std::min<unsigned long long>(chunk, std::max(chunkMin, (unsigned long long)tuneParam("RTB_CHUNK_TAIL", 32)));
  a.chunkDiv = (uint32_t)std::max<unsigned long long>(1, warps * tuneParam("RTB_CHUNK_DIV", 8));  // 8: cfg3 +15 % over 2 (its 1024-path chunks are 170 ms of warp time), others -0.2 % (profiles/r2_chunk_div_probe.log)
  // scheduler thresholds (rt_render_kernel.inl): single rects are cheap steps and worth running for 3 lanes, cuboid-heavy scenes batch
  // 6; media make the rare class frequent (3 rare steps per medium visit) and worth running for 4 lanes.  Measured on the
  // reference scenes (tools/knob_sweep.py); results never depend on these.
  a.shadeMin = tuneParam("RTB_SHADE_MIN", 16);
  a.rectMin = std::max(1u, tuneParam("RTB_RECT_MIN", hs.n_cuboids > hs.n_rects ? 6u : 3u));
  a.rareMin = std::max(1u, tuneParam("RTB_RARE_MIN", hs.n_media > 0 ? 4u : 8u));
  // fast path of the vote: with a handful of BVH nodes (the Cornell boxes: 8) the node class rarely collects 12 waiting lanes and every
  // iteration would pay the full vote; 8 / 6 there is +27 % on cfg4's renderKernel, 12 / 8 stays the optimum of cfg1-3 (profiles/r2_keep_sweep.log)
  const bool fewNodes = hs.n_nodes <= 64;
  a.keepNode = std::max(1u, tuneParam("RTB_KEEP_NODE", fewNodes ? 8 : 12)); a.keepLeaf = std::max(1u, tuneParam("RTB_KEEP_LEAF", fewNodes ? 6 : 8));
  a.lightCap = hs.lightCap;
  a.counter = c->d_counter; a.ctr = instr == 1 ? c->d_counter + kCtrWorkBase : nullptr; a.out = d_out;
  if (plog) { a.plogPrim = plog->prim; a.plogT = plog->t; a.plogStride = plog->stride; }
  const bool coop = useCoop(hs, q.flags);
  if (kindOut) *kindOut = coop ? 1u : 0u;
  cudaError_t ce;
  if (!coop && useWavefront(hs, q, instr)) {
    // slots: enough to keep every resident traversal lane fed several times per round, never more than chunks of paths
    const unsigned long long maxP = (unsigned long long)c->numSMs * 8ull * kBlock * tuneParam("RTB_WF_SLOTS_PER_LANE", 16);
    a.wfChunk = (uint32_t)std::max<unsigned long long>(1, std::min<unsigned long long>(32, a.total_paths / maxP));
    unsigned long long P = std::min(maxP, (a.total_paths + a.wfChunk - 1) / a.wfChunk);
    P = (P + kBlock - 1) / kBlock * kBlock;
    const size_t bytes = wfLayout((uint32_t)P, nullptr, nullptr);
    if (c->wfBytes < bytes) {
      if (c->wfBlob) cudaFree(c->wfBlob);
      c->wfBlob = nullptr; c->wfBytes = 0;
      size_t got = 0;
      if (void* pooled = g_scratch.take(c->device, bytes, &got)) { c->wfBlob = pooled; c->wfBytes = got; }
      else { CU(cudaMalloc(&c->wfBlob, bytes)); c->wfBytes = bytes; }
    }
    if (!c->h_wfCtl) CU(cudaHostAlloc(reinterpret_cast<void**>(&c->h_wfCtl), 2 * sizeof(unsigned long long), cudaHostAllocDefault));
    wfLayout((uint32_t)P, static_cast<char*>(c->wfBlob), &c->wf);
    a.wfMedia = hs.n_media > 0 ? 1u : 0u;
    a.wfGuideInv = 1.0f / ((float)P * (float)std::max(1u, tuneParam("RTB_WF_GUIDE", 4)));  // a slot asks for 1 / (4 P) of what is left (tools/wf_guide_sweep.py)
    a.wfSort = tuneParam("RTB_WF_SORT", 1);  // block-level material sort of the shade kernel: +4 % once the drain is short (profiles/r2_wf_sort_probe.log; it measured neutral while the drain dominated)
    a.shadeMin = tuneParam("RTB_WF_ITEM_MIN", 8);
    // the scenes the wavefront is chosen for (<= 64 nodes): 8 / 4 is +4 % over 12 / 8 (profiles/r2_keep_sweep_wavefront.log); a forced wavefront on a
    // large BVH keeps renderKernel's 12 / 8 (8 / 4 costs cfg1 / cfg2 9 % there)
    a.keepNode = std::max(1u, tuneParam("RTB_WF_KEEP_NODE", fewNodes ? 8 : 12)); a.keepLeaf = std::max(1u, tuneParam("RTB_WF_KEEP_LEAF", fewNodes ? 4 : 8));
    const unsigned long long lanes = (unsigned long long)c->numSMs * 8ull * kBlock;
    a.chunk = (uint32_t)std::max<unsigned long long>(32, std::min<unsigned long long>(256, P / std::max<unsigned long long>(1, lanes / 32 * 4)));
    const int rounds0 = (int)std::min<unsigned long long>(4096, a.total_paths * tuneParam("RTB_WF_EST_BOUNCES", 4) / P + 24);
    if (!c->wfProbe) CU(cudaEventCreateWithFlags(&c->wfProbe, cudaEventDisableTiming));
    c->wfArgs = a; c->wfScene = dsc; c->wfPrec = (int)q.precision; c->wfStream = st; c->wfRounds = 0; c->wfPending = true;
    if (gridOut) *gridOut = 0;
    if ((rc = wavefrontQueue(hs, c, rounds0, true))) return rc;
    if (kindOut) *kindOut = 2u;
    if (wfDefer) return RT_OK;  // multi-GPU shares: every device gets its first batch before any is topped up (wavefrontFinishAll)
    DeviceCtx* one[1] = {c};
    return wavefrontFinishAll(hs, one, 1);
  }
  if (coop && (q.flags & RT_FLAG_STRICT_MATH)) ce = launchRenderCoop<MATH_STRICT>((int)q.precision, instr, dsc, a, c->numSMs, st, gridOut);
  else if (coop) ce = launchRenderCoop<MATH_LITERAL>((int)q.precision, instr, dsc, a, c->numSMs, st, gridOut);
  else if (q.flags & RT_FLAG_FAST_MATH) ce = launchRender<MATH_FAST>((int)q.precision, instr == 3 ? 0 : instr, dsc, a, hs.stackDepth, c->numSMs, st, gridOut);
  else if (q.flags & RT_FLAG_STRICT_MATH) ce = launchRender<MATH_STRICT>((int)q.precision, instr, dsc, a, hs.stackDepth, c->numSMs, st, gridOut);
  else ce = launchRender<MATH_LITERAL>((int)q.precision, instr, dsc, a, hs.stackDepth, c->numSMs, st, gridOut);
  if (ce != cudaSuccess) return fail(RT_ERR_CUDA, std::string("render launch: ") + cudaGetErrorString(ce));
  return RT_OK;
}

int ensureFramebuffer(DeviceCtx* c, size_t elems) {
  if (c->fbElems < elems) {
    if (c->d_fb) cudaFree(c->d_fb);
    c->d_fb = nullptr; c->fbElems = 0;
    CU(cudaMalloc(&c->d_fb, elems * sizeof(double)));
    c->fbElems = elems;
  }
  return RT_OK;
}
int ensureHostStaging(SceneEntry* e, size_t elems) {
  if (e->hfbElems < elems) {
    if (e->h_fb) cudaFreeHost(e->h_fb);
    e->h_fb = nullptr; e->hfbElems = 0;
    CU(cudaHostAlloc(&e->h_fb, elems * sizeof(double), cudaHostAllocDefault));
    e->hfbElems = elems;
  }
  return RT_OK;
}

// ---- multi-GPU inside rt_render: the frame's samples are split into contiguous slices over the devices (interleaved rows
// when there are fewer samples than devices), every device renders its share into its own double framebuffer on its own stream, and
// the partial sums are added onto the scene's device: by ONE kernel there that reads the peers' buffers over NVLink (peer access),
// or by ncclReduce when rt_init asked for NCCL, or through staged peer copies when the devices cannot address each other.
int ensureReplica(SceneEntry* e, int device, DeviceCtx** out) {
  if (device == e->main.device) { *out = &e->main; }
  else {
    auto& slot = e->replicas[device];
    if (!slot) {
      slot.reset(new DeviceCtx());
      slot->device = device;
      CU(cudaDeviceGetAttribute(&slot->numSMs, cudaDevAttrMultiProcessorCount, device));  // one attribute: cudaGetDeviceProperties costs milliseconds per call
    }
    *out = slot.get();
  }
  DeviceCtx* c = *out;
  CU(cudaSetDevice(device));
  if (!c->stream) {
    CU(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
    CU(cudaEventCreate(&c->evStart)); CU(cudaEventCreate(&c->evDone));
  }
  return RT_OK;
}

int renderMulti(SceneEntry* e, const RtRenderParams& q, int nGpus, double* out, RtStats* stats) {
  const size_t elems = (size_t)q.width * q.height * 3;
  std::vector<int> devs{e->main.device};
  for (int d = 0; d < nGpus && (int)devs.size() < nGpus; ++d) if (d != e->main.device) devs.push_back(d);
  const int n = (int)devs.size();
  std::vector<DeviceCtx*> ctx(n);
  const auto t0 = std::chrono::steady_clock::now();
  bool replicated = false;
  for (int k = 0; k < n; ++k) {
    int rc = ensureReplica(e, devs[k], &ctx[k]);
    if (rc) return rc;
    const bool had = (q.precision == RT_PREC_F64) ? ctx[k]->f64.ready : ctx[k]->f32.ready;
    const void* dsc = nullptr;
    if ((rc = deviceScene(e->hs, ctx[k], q.precision, &dsc))) return rc;  // replication happens here, once per device and layout
    replicated = replicated || !had;
    if ((rc = ensureFramebuffer(ctx[k], elems))) return rc;
  }
  const double replicateMs = replicated ? std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count() : 0.0;
  int rc = ensureHostStaging(e, elems);
  if (rc) return rc;
  // peer access root <- k
  DeviceCtx* root = ctx[0];
  bool allPeer = true;
  CU(cudaSetDevice(root->device));
  for (int k = 1; k < n; ++k) {
    int can = 0;
    CU(cudaDeviceCanAccessPeer(&can, root->device, devs[k]));
    if (can) {
      cudaError_t pe = cudaDeviceEnablePeerAccess(devs[k], 0);
      if (pe == cudaErrorPeerAccessAlreadyEnabled) cudaGetLastError();
      else if (pe != cudaSuccess) { cudaGetLastError(); can = 0; }
    }
    allPeer = allPeer && can;
  }
  // shares
  uint32_t kernelKind = 0;
  const bool bySamples = q.sample_count >= (uint32_t)n;
  const uint32_t rows = q.row1 - q.row0;
  for (int k = 0; k < n; ++k) {
    RtRenderParams qk = q;
    uint32_t rowStep = 1;
    if (bySamples) {
      const uint32_t base = q.sample_count / n, rem = q.sample_count % n;
      qk.sample_offset = q.sample_offset + k * base + std::min<uint32_t>(k, rem);
      qk.sample_count = base + ((uint32_t)k < rem ? 1u : 0u);
    } else {
      // fewer samples than devices: device k takes rows row0 + k, row0 + k + n, ... (interleaved: neighbouring rows cost alike, bands do not --
      // the cfg5 sample's upper band was 30 % cheaper than its lower one)
      qk.row0 = q.row0 + (uint32_t)k;
      rowStep = (uint32_t)n;
      if ((uint32_t)k >= rows) qk.sample_count = 0;
    }
    CU(cudaSetDevice(devs[k]));
    CU(cudaMemsetAsync(ctx[k]->d_fb, 0, elems * sizeof(double), ctx[k]->stream));
    CU(cudaEventRecord(ctx[k]->evStart, ctx[k]->stream));
    ctx[k]->wfPending = false;
    if ((rc = renderLaunch(e->hs, ctx[k], qk, ctx[k]->d_fb, ctx[k]->stream, 0, nullptr, nullptr, &kernelKind, /*wfDefer=*/true, rowStep))) return rc;
    if (!ctx[k]->wfPending) CU(cudaEventRecord(ctx[k]->evDone, ctx[k]->stream));
  }
  {  // wavefront shares: every device has its first batch of rounds queued; top them up round-robin until each frame is finished
    std::vector<int> wasPending(n);
    for (int k = 0; k < n; ++k) wasPending[k] = ctx[k]->wfPending ? 1 : 0;
    if ((rc = wavefrontFinishAll(e->hs, ctx.data(), n))) return rc;
    for (int k = 0; k < n; ++k) if (wasPending[k]) { CU(cudaSetDevice(devs[k])); CU(cudaEventRecord(ctx[k]->evDone, ctx[k]->stream)); }
  }
  // reduction onto the root
  CU(cudaSetDevice(root->device));
  Events evs;
  CU(evs.create(3));
  uint32_t kind = RT_REDUCE_NONE;
  bool useNccl = false;
  {
    std::lock_guard<std::mutex> lk(g_multi.mu);
    useNccl = g_multi.wantNccl && !g_multi.comms.empty();
    for (int k = 0; k < n && useNccl; ++k) if (devs[k] >= (int)g_multi.comms.size()) useNccl = false;
    if (useNccl && (int)g_multi.comms.size() != n) useNccl = false;  // the communicator spans exactly the rt_init devices
  }
  if (n > 1 && useNccl) {
    kind = RT_REDUCE_NCCL;
    CU(cudaEventRecord(evs.ev[0], root->stream));
    int nrc = g_multi.groupStart();
    for (int k = 0; k < n && nrc == 0; ++k)
      nrc = g_multi.reduce(ctx[k]->d_fb, ctx[k]->d_fb, elems, /*ncclFloat64*/ 8, /*ncclSum*/ 0, root->device, g_multi.comms[devs[k]], ctx[k]->stream);
    const int grc = g_multi.groupEnd();
    if (nrc == 0) nrc = grc;
    if (nrc != 0) return fail(RT_ERR_CUDA, std::string("ncclReduce: ") + (g_multi.errorString ? g_multi.errorString(nrc) : "error"));
    CU(cudaEventRecord(evs.ev[1], root->stream));
  } else {
    for (int k = 1; k < n; ++k) CU(cudaStreamWaitEvent(root->stream, ctx[k]->evDone, 0));
    CU(cudaEventRecord(evs.ev[0], root->stream));
    if (n > 1 && allPeer) {
      kind = RT_REDUCE_PEER;
      const double* src[16] = {nullptr};
      for (int k = 1; k < n; ++k) src[k - 1] = ctx[k]->d_fb;
      cudaError_t ce = launchReducePeer(root->d_fb, src, n - 1, elems, root->numSMs, root->stream);
      if (ce != cudaSuccess) return fail(RT_ERR_CUDA, std::string("reduce launch: ") + cudaGetErrorString(ce));
    } else if (n > 1) {
      kind = RT_REDUCE_STAGED;
      double* tmp = nullptr;
      CU(cudaMalloc(&tmp, elems * sizeof(double)));
      for (int k = 1; k < n; ++k) {
        CU(cudaMemcpyPeerAsync(tmp, root->device, ctx[k]->d_fb, devs[k], elems * sizeof(double), root->stream));
        const double* src[16] = {tmp};
        cudaError_t ce = launchReducePeer(root->d_fb, src, 1, elems, root->numSMs, root->stream);
        if (ce != cudaSuccess) { cudaFree(tmp); return fail(RT_ERR_CUDA, std::string("reduce launch: ") + cudaGetErrorString(ce)); }
      }
      CU(cudaStreamSynchronize(root->stream));
      cudaFree(tmp);
    }
    CU(cudaEventRecord(evs.ev[1], root->stream));
  }
  const size_t rowElems = (size_t)q.width * 3;
  CU(cudaMemcpyAsync(e->h_fb + q.row0 * rowElems, root->d_fb + q.row0 * rowElems, (size_t)(q.row1 - q.row0) * rowElems * sizeof(double), cudaMemcpyDeviceToHost, root->stream));
  CU(cudaEventRecord(evs.ev[2], root->stream));
  CU(cudaEventSynchronize(evs.ev[2]));
  for (int k = 1; k < n; ++k) { CU(cudaSetDevice(devs[k])); CU(cudaStreamSynchronize(ctx[k]->stream)); }
  CU(cudaSetDevice(root->device));
  for (uint32_t r = q.row0; r < q.row1; ++r)
    std::memcpy(out + r * rowElems + (size_t)q.col0 * 3, e->h_fb + r * rowElems + (size_t)q.col0 * 3, (size_t)(q.col1 - q.col0) * 3 * sizeof(double));
  if (stats) {
    std::memset(stats, 0, sizeof(*stats));
    float ms = 0;
    for (int k = 0; k < n; ++k) {
      cudaEventElapsedTime(&ms, ctx[k]->evStart, ctx[k]->evDone);
      if (k < 16) stats->gpu_kernel_ms[k] = ms;
      stats->kernel_ms = std::max(stats->kernel_ms, (double)ms);
    }
    for (int k = 0; k < n; ++k) {
      unsigned long long tt[2] = {0, 0};
      CU(cudaSetDevice(devs[k]));
      CU(cudaMemcpy(tt, ctx[k]->d_counter + kCtrTailBase, sizeof(tt), cudaMemcpyDeviceToHost));
      if (tt[0] && tt[1] > tt[0]) stats->tail_ms = std::max(stats->tail_ms, (double)(tt[1] - tt[0]) * 1e-6);
    }
    CU(cudaSetDevice(root->device));
    cudaEventElapsedTime(&ms, evs.ev[0], evs.ev[1]); stats->reduce_ms = ms;
    cudaEventElapsedTime(&ms, evs.ev[1], evs.ev[2]); stats->d2h_ms = ms;
    stats->samples = (uint64_t)(q.row1 - q.row0) * (q.col1 - q.col0) * q.sample_count;
    stats->launches = (uint64_t)n + (kind == RT_REDUCE_PEER ? 1 : 0);
    stats->gpus = (uint32_t)n; stats->reduce_kind = kind; stats->replicate_ms = replicateMs; stats->kernel_kind = kernelKind; stats->rounds = kernelKind == 2 ? (uint32_t)ctx[0]->wfRounds : 0u;
  }
  return RT_OK;
}

int renderHost(rt_scene_t scene, const RtRenderParams* params, double* out, RtStats* stats, int instr, const PathLog* hostLog = nullptr) {
  const std::shared_ptr<SceneEntry> ep = lookup(scene);
  if (!ep) return fail(RT_ERR_HANDLE, "unknown scene handle");
  SceneEntry* e = ep.get();
  std::lock_guard<std::mutex> lk(e->mu);
  if (!out) return fail(RT_ERR_ARG, "out_rgb_sum is null");
  RtRenderParams q;
  int rc = normaliseParams(params, q);
  if (rc) return rc;
  int nGpus = (int)q.n_gpus;
  if (nGpus == 0) { std::lock_guard<std::mutex> lk2(g_multi.mu); nGpus = g_multi.defaultGpus; }
  if (nGpus > 1 && instr == 0) {
    int ndev = 0;
    CU(cudaGetDeviceCount(&ndev));
    if (nGpus > ndev) return fail(RT_ERR_ARG, "n_gpus exceeds the number of visible devices");
    if (nGpus > 16) return fail(RT_ERR_ARG, "n_gpus > 16");
    return renderMulti(e, q, nGpus, out, stats);
  }
  DeviceCtx* c = &e->main;
  CU(cudaSetDevice(c->device));
  const size_t elems = (size_t)q.width * q.height * 3;
  if ((rc = ensureFramebuffer(c, elems))) return rc;
  if ((rc = ensureHostStaging(e, elems))) return rc;
  Events evs;
  CU(evs.create(3));
  const cudaEvent_t e0 = evs.ev[0], e1 = evs.ev[1], e2 = evs.ev[2];
  struct Buf { void* p = nullptr; ~Buf() { if (p) cudaFree(p); } } dPrim, dT;
  PathLog dlog;
  const size_t paths = (size_t)(q.row1 - q.row0) * (q.col1 - q.col0) * q.sample_count;
  if (instr == 3) {
    const size_t cells = paths * hostLog->stride;
    CU(cudaMalloc(&dPrim.p, std::max<size_t>(cells, 1) * 4)); CU(cudaMalloc(&dT.p, std::max<size_t>(cells, 1) * 8));
    CU(cudaMemcpy(dPrim.p, hostLog->prim, cells * 4, cudaMemcpyHostToDevice)); CU(cudaMemcpy(dT.p, hostLog->t, cells * 8, cudaMemcpyHostToDevice));
    dlog.prim = (uint32_t*)dPrim.p; dlog.t = (double*)dT.p; dlog.stride = hostLog->stride;
  }
  CU(cudaMemsetAsync(c->d_fb, 0, elems * sizeof(double), 0));
  CU(cudaEventRecord(e0, 0));
  int grid = 0;
  uint32_t kernelKind = 0;
  rc = renderLaunch(e->hs, c, q, c->d_fb, 0, instr, &grid, instr == 3 ? &dlog : nullptr, &kernelKind);
  if (rc) return rc;
  CU(cudaEventRecord(e1, 0));
  // copy back only the rows of the region
  const size_t rowElems = (size_t)q.width * 3;
  CU(cudaMemcpyAsync(e->h_fb + q.row0 * rowElems, c->d_fb + q.row0 * rowElems, (size_t)(q.row1 - q.row0) * rowElems * sizeof(double), cudaMemcpyDeviceToHost, 0));
  CU(cudaEventRecord(e2, 0));
  CU(cudaEventSynchronize(e2));
  for (uint32_t r = q.row0; r < q.row1; ++r)
    std::memcpy(out + r * rowElems + (size_t)q.col0 * 3, e->h_fb + r * rowElems + (size_t)q.col0 * 3, (size_t)(q.col1 - q.col0) * 3 * sizeof(double));
  if (instr == 3) {
    const size_t cells = paths * hostLog->stride;
    CU(cudaMemcpy(hostLog->prim, dPrim.p, cells * 4, cudaMemcpyDeviceToHost)); CU(cudaMemcpy(hostLog->t, dT.p, cells * 8, cudaMemcpyDeviceToHost));
  }
  if (stats) {
    std::memset(stats, 0, sizeof(*stats));
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1); stats->kernel_ms = ms; stats->gpu_kernel_ms[0] = ms;
    cudaEventElapsedTime(&ms, e1, e2); stats->d2h_ms = ms;
    stats->samples = (uint64_t)paths;
    stats->launches = q.sample_count ? 1 : 0;
    stats->gpus = 1; stats->kernel_kind = kernelKind; stats->rounds = kernelKind == 2 ? (uint32_t)c->wfRounds : 0u;
    if (kernelKind == 2) stats->launches = 1ull + 2ull * (uint64_t)c->wfRounds;
    unsigned long long cc[kCtrTotal];
    CU(cudaMemcpy(cc, c->d_counter, sizeof(cc), cudaMemcpyDeviceToHost));
    if (instr == 1) {
      for (int i = 0; i < kWorkCounters; ++i) stats->counters[i] = cc[kCtrWorkBase + i];
      for (int i = 0; i < kSchedCounters; ++i) stats->sched[i] = cc[kCtrSchedBase + i];
    }
    if (cc[kCtrTailBase] && cc[kCtrTailBase + 1] > cc[kCtrTailBase]) stats->tail_ms = (double)(cc[kCtrTailBase + 1] - cc[kCtrTailBase]) * 1e-6;
  }
  return RT_OK;
}

// ---- rt_render_device_static: renderKernel's scratch (the work-counter block and the double framebuffer the kernel accumulates into)
// in a STATIC device array instead of cudaMalloc'd buffers, for callers that may not allocate (include/rt_abi.h).  No counterpart in
// the reference.  Every GPU that loads the library holds the array until exit.
constexpr size_t kStaticCtrWords = 64;                                   // >= kCtrTotal, keeps the framebuffer 256-byte aligned
constexpr size_t kStaticFbElems = (size_t)RT_STATIC_MAX_PIXELS * 3;      // doubles
static_assert(kCtrTotal <= (int)kStaticCtrWords, "counter block does not fit the static scratch");
static __device__ __align__(256) unsigned long long g_staticScratch[kStaticCtrWords + kStaticFbElems];  // 8-byte words: counters, then the framebuffer

std::mutex g_staticMu;                           // host side of the calls is serialised; launches that overlap on the GPU still race (documented)
unsigned long long* g_staticPtr[64] = {};        // per device: the array's DEVICE address (the symbol itself is only a host-side stand-in)

// once per device, with that device current
int staticScratch(int device, unsigned long long** out) {
  if (device < 0 || device >= 64) return fail(RT_ERR_ARG, "rt_render_device_static: device index out of range");
  if (!g_staticPtr[device]) {
    CU(cudaSetDevice(device));
    void* p = nullptr;
    cudaError_t ce = cudaGetSymbolAddress(&p, g_staticScratch);
    if (ce != cudaSuccess || !p) return fail(RT_ERR_CUDA, std::string("cudaGetSymbolAddress(static scratch): ") + cudaGetErrorString(ce));
    g_staticPtr[device] = static_cast<unsigned long long*>(p);
  }
  *out = g_staticPtr[device];
  return RT_OK;
}

}  // namespace

namespace rtb { void setLastError(const std::string& s) { g_err = s; } }  // for the other translation units of the library

extern "C" {

int rt_abi_version(void) { return RT_ABI_VERSION; }
const char* rt_last_error(void) { return g_err.c_str(); }
int rt_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
  return n;
}
int rt_get_limits(RtLimits* out) {
  if (!out) return fail(RT_ERR_ARG, "out is null");
  out->max_stack = kMaxStack; out->max_frames = kMaxFrames; out->max_medium_depth = 1; out->max_light_depth = kMaxLightDepth;
  out->max_tex_depth = kMaxTexDepth;
  return RT_OK;
}

int rt_init(int n_gpus, uint32_t flags) {
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) { cudaGetLastError(); return fail(RT_ERR_CUDA, "no CUDA device available (this library has no CPU path)"); }
  if (n_gpus < 0 || n_gpus > ndev) return fail(RT_ERR_ARG, "n_gpus out of range");
  if (n_gpus == 0) n_gpus = ndev;
  if (n_gpus > 16) return fail(RT_ERR_ARG, "n_gpus > 16");
  std::lock_guard<std::mutex> lk(g_multi.mu);
  // peer access between every pair (the reduction reads the peers' framebuffers from whichever device owns the scene)
  for (int a = 0; a < n_gpus; ++a) {
    CU(cudaSetDevice(a));
    for (int b = 0; b < n_gpus; ++b) {
      if (a == b) continue;
      int can = 0;
      if (cudaDeviceCanAccessPeer(&can, a, b) == cudaSuccess && can) { if (cudaDeviceEnablePeerAccess(b, 0) != cudaSuccess) cudaGetLastError(); }
    }
  }
  CU(cudaSetDevice(0));
  for (void* c : g_multi.comms) if (c && g_multi.commDestroy) g_multi.commDestroy(c);
  g_multi.comms.clear();
  g_multi.wantNccl = false;
  if ((flags & RT_INIT_NCCL) && n_gpus > 1) {
    if (!g_multi.lib) {
      g_multi.lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
      if (!g_multi.lib) return fail(RT_ERR_CUDA, std::string("RT_INIT_NCCL: cannot load libnccl.so.2: ") + dlerror());
      g_multi.commInitAll = reinterpret_cast<int (*)(void**, int, const int*)>(dlsym(g_multi.lib, "ncclCommInitAll"));
      g_multi.commDestroy = reinterpret_cast<int (*)(void*)>(dlsym(g_multi.lib, "ncclCommDestroy"));
      g_multi.reduce = reinterpret_cast<int (*)(const void*, void*, size_t, int, int, int, void*, cudaStream_t)>(dlsym(g_multi.lib, "ncclReduce"));
      g_multi.groupStart = reinterpret_cast<int (*)()>(dlsym(g_multi.lib, "ncclGroupStart"));
      g_multi.groupEnd = reinterpret_cast<int (*)()>(dlsym(g_multi.lib, "ncclGroupEnd"));
      g_multi.errorString = reinterpret_cast<const char* (*)(int)>(dlsym(g_multi.lib, "ncclGetErrorString"));
      if (!g_multi.commInitAll || !g_multi.commDestroy || !g_multi.reduce || !g_multi.groupStart || !g_multi.groupEnd)
        return fail(RT_ERR_CUDA, "RT_INIT_NCCL: libnccl.so.2 lacks a required entry point");
    }
    std::vector<int> devs(n_gpus);
    for (int i = 0; i < n_gpus; ++i) devs[i] = i;
    g_multi.comms.assign(n_gpus, nullptr);
    const int nrc = g_multi.commInitAll(g_multi.comms.data(), n_gpus, devs.data());
    if (nrc != 0) { g_multi.comms.clear(); return fail(RT_ERR_CUDA, std::string("ncclCommInitAll: ") + (g_multi.errorString ? g_multi.errorString(nrc) : "error")); }
    g_multi.wantNccl = true;
  }
  g_multi.defaultGpus = n_gpus;
  return n_gpus;
}

int rt_scene_create(const RtSceneDesc* d, int device, rt_scene_t* out) {
  if (!d || !out) return fail(RT_ERR_ARG, "desc/out is null");
  if (d->abi_version != RT_ABI_VERSION) return fail(RT_ERR_ARG, "abi_version mismatch");
  auto need = [](uint32_t n, const void* p) { return n == 0 || p != nullptr; };
  if (!need(d->n_nodes, d->node_box) || !need(d->n_nodes, d->node_left) || !need(d->n_nodes, d->node_right) || !need(d->n_nodes, d->node_size) ||
      !need(d->n_spheres, d->sphere_cr) || !need(d->n_spheres, d->sphere_mat) || !need(d->n_moving, d->moving) || !need(d->n_moving, d->moving_mat) ||
      !need(d->n_rects, d->rect) || !need(d->n_rects, d->rect_mat) || !need(d->n_cuboids, d->cuboid) || !need(d->n_cuboids, d->cuboid_mat) ||
      !need(d->n_translates, d->translate_off) || !need(d->n_translates, d->translate_child) || !need(d->n_rotates, d->rotate_sc) ||
      !need(d->n_rotates, d->rotate_axis) || !need(d->n_rotates, d->rotate_child) || !need(d->n_media, d->medium_neg_inv_density) ||
      !need(d->n_media, d->medium_mat) || !need(d->n_media, d->medium_boundary) || !need(d->n_materials, d->mat_kind) || !need(d->n_materials, d->mat_tex) ||
      !need(d->n_materials, d->mat_param) || !need(d->n_textures, d->tex_kind) || !need(d->n_textures, d->tex_rgb) || !need(d->n_textures, d->tex_a) ||
      !need(d->n_textures, d->tex_b) || !need(d->n_textures, d->tex_scale) || !need(d->n_perlin, d->perlin_ranvec) || !need(d->n_perlin, d->perlin_perm) ||
      !need(d->n_images, d->image_rgb) || !need(d->n_images, d->image_w) || !need(d->n_images, d->image_h))
    return fail(RT_ERR_ARG, "a non-empty array is null");
  const uint32_t lim = kIdxMask;  // 27-bit indices: bit 27 of a ref is the pair mark
  if (d->n_nodes > lim || d->n_spheres > lim || d->n_moving > lim || d->n_rects > lim || d->n_cuboids > lim) return fail(RT_ERR_ARG, "too many objects");

  auto e = std::make_shared<SceneEntry>();
  HostScene& h = e->hs;
  h.n_nodes = d->n_nodes; h.n_spheres = d->n_spheres; h.n_moving = d->n_moving; h.n_rects = d->n_rects; h.n_cuboids = d->n_cuboids;
  h.n_translates = d->n_translates; h.n_rotates = d->n_rotates; h.n_media = d->n_media; h.n_materials = d->n_materials; h.n_textures = d->n_textures;
  h.n_perlin = d->n_perlin; h.n_images = d->n_images;
  copyIn(h.node_box, d->node_box, 6 * (size_t)h.n_nodes); copyIn(h.node_left, d->node_left, h.n_nodes); copyIn(h.node_right, d->node_right, h.n_nodes);
  copyIn(h.node_size, d->node_size, h.n_nodes);
  copyIn(h.sphere_cr, d->sphere_cr, 4 * (size_t)h.n_spheres); copyIn(h.sphere_mat, d->sphere_mat, h.n_spheres);
  copyIn(h.moving, d->moving, 9 * (size_t)h.n_moving); copyIn(h.moving_mat, d->moving_mat, h.n_moving);
  copyIn(h.rect, d->rect, 5 * (size_t)h.n_rects); copyIn(h.rect_mat, d->rect_mat, h.n_rects);
  copyIn(h.cuboid, d->cuboid, 6 * (size_t)h.n_cuboids); copyIn(h.cuboid_mat, d->cuboid_mat, h.n_cuboids);
  copyIn(h.translate_off, d->translate_off, 3 * (size_t)h.n_translates); copyIn(h.translate_child, d->translate_child, h.n_translates);
  copyIn(h.rotate_sc, d->rotate_sc, 2 * (size_t)h.n_rotates); copyIn(h.rotate_axis, d->rotate_axis, h.n_rotates); copyIn(h.rotate_child, d->rotate_child, h.n_rotates);
  copyIn(h.medium_nid, d->medium_neg_inv_density, h.n_media); copyIn(h.medium_mat, d->medium_mat, h.n_media); copyIn(h.medium_boundary, d->medium_boundary, h.n_media);
  copyIn(h.mat_kind, d->mat_kind, h.n_materials); copyIn(h.mat_tex, d->mat_tex, h.n_materials); copyIn(h.mat_param, d->mat_param, h.n_materials);
  copyIn(h.tex_kind, d->tex_kind, h.n_textures); copyIn(h.tex_rgb, d->tex_rgb, 3 * (size_t)h.n_textures); copyIn(h.tex_a, d->tex_a, h.n_textures);
  copyIn(h.tex_b, d->tex_b, h.n_textures); copyIn(h.tex_scale, d->tex_scale, h.n_textures);
  copyIn(h.perlin_ranvec, d->perlin_ranvec, 768 * (size_t)h.n_perlin); copyIn(h.perlin_perm, d->perlin_perm, 768 * (size_t)h.n_perlin);
  copyIn(h.image_w, d->image_w, h.n_images); copyIn(h.image_h, d->image_h, h.n_images);
  h.images.resize(h.n_images);
  for (uint32_t i = 0; i < h.n_images; ++i) {
    if (!d->image_rgb[i] || h.image_w[i] == 0 || h.image_h[i] == 0) return fail(RT_ERR_ARG, "bad image");
    h.images[i].assign(d->image_rgb[i], d->image_rgb[i] + (size_t)h.image_w[i] * h.image_h[i] * 3);
  }
  h.world_root = d->world_root; h.lights_root = d->lights_root;
  std::memcpy(h.background, d->background, sizeof(h.background)); std::memcpy(h.camera, d->camera, sizeof(h.camera));

  // ---- validation
  auto matOk = [&](uint32_t m) { return m < h.n_materials; };
  for (uint32_t i = 0; i < h.n_nodes; ++i) if (!refOk(h, h.node_left[i]) || !refOk(h, h.node_right[i])) return fail(RT_ERR_ARG, "BVH child ref out of range");
  for (uint32_t i = 0; i < h.n_translates; ++i) if (!refOk(h, h.translate_child[i])) return fail(RT_ERR_ARG, "translate child out of range");
  for (uint32_t i = 0; i < h.n_rotates; ++i) if (!refOk(h, h.rotate_child[i]) || h.rotate_axis[i] > 2) return fail(RT_ERR_ARG, "rotate child/axis out of range");
  for (uint32_t i = 0; i < h.n_media; ++i) if (!refOk(h, h.medium_boundary[i]) || !matOk(h.medium_mat[i])) return fail(RT_ERR_ARG, "medium boundary/material out of range");
  for (uint32_t i = 0; i < h.n_spheres; ++i) if (!matOk(h.sphere_mat[i])) return fail(RT_ERR_ARG, "sphere material out of range");
  for (uint32_t i = 0; i < h.n_moving; ++i) if (!matOk(h.moving_mat[i])) return fail(RT_ERR_ARG, "moving sphere material out of range");
  for (uint32_t i = 0; i < h.n_rects; ++i) if (!matOk(h.rect_mat[i])) return fail(RT_ERR_ARG, "rect material out of range");
  for (uint32_t i = 0; i < h.n_cuboids; ++i) if (!matOk(h.cuboid_mat[i])) return fail(RT_ERR_ARG, "cuboid material out of range");
  for (uint32_t i = 0; i < h.n_materials; ++i) {
    if (h.mat_kind[i] > M_ISOTROPIC) return fail(RT_ERR_ARG, "bad material kind");
    if (h.mat_kind[i] != M_DIELECTRIC && h.mat_tex[i] >= h.n_textures) return fail(RT_ERR_ARG, "material texture out of range");
  }
  for (uint32_t i = 0; i < h.n_textures; ++i) {
    uint32_t k = h.tex_kind[i];
    if (k > TX_IMAGE) return fail(RT_ERR_ARG, "bad texture kind");
    if (k == TX_CHECKER && (h.tex_a[i] >= h.n_textures || h.tex_b[i] >= h.n_textures)) return fail(RT_ERR_ARG, "checker child out of range");
    if (k == TX_PERLIN && h.tex_a[i] >= h.n_perlin) return fail(RT_ERR_ARG, "perlin table out of range");
    if (k == TX_IMAGE && h.tex_a[i] != RT_NONE && h.tex_a[i] >= h.n_images) return fail(RT_ERR_ARG, "image out of range");
  }
  {  // checker nesting: textureValue walks at most kMaxTexDepth levels; a cycle would never end in the reference either
    std::vector<int> depth(h.n_textures, 0);  // 0 unknown, -1 in progress, > 0 levels below and including this texture
    bool cyc = false;
    std::function<int(uint32_t)> levels = [&](uint32_t t) -> int {
      if (depth[t] > 0) return depth[t];
      if (depth[t] < 0) { cyc = true; return 1; }
      if (h.tex_kind[t] != TX_CHECKER) return depth[t] = 1;
      depth[t] = -1;
      const int d = 1 + std::max(levels(h.tex_a[t]), levels(h.tex_b[t]));
      return depth[t] = d;
    };
    int deepest = 0;
    for (uint32_t i = 0; i < h.n_textures && !cyc; ++i) deepest = std::max(deepest, levels(i));
    if (cyc) return fail(RT_ERR_ARG, "checker textures form a cycle");
    if (deepest > kMaxTexDepth) return fail(RT_ERR_UNSUPPORTED, "more than " + std::to_string(kMaxTexDepth) + " nested CheckerTexture levels");
  }
  if (!refOk(h, h.world_root) || !refOk(h, h.lights_root)) return fail(RT_ERR_ARG, "root ref out of range");
  Analysis an(h);
  an.go(h.world_root); an.go(h.lights_root);
  if (!an.cycle) for (uint32_t i = 0; i < h.n_nodes; ++i) { an.go(h.node_left[i]); an.go(h.node_right[i]); }
  if (an.cycle) return fail(RT_ERR_ARG, "hittable graph has a cycle");
  const int stackNeed = an.maxStack, frames = an.maxFrames, media = an.maxMedia;
  if (stackNeed > kMaxStack) return fail(RT_ERR_UNSUPPORTED, "scene needs a traversal stack of " + std::to_string(stackNeed) + " entries (limit " + std::to_string(kMaxStack) + ")");
  if (frames > kMaxFrames) return fail(RT_ERR_UNSUPPORTED, "more than " + std::to_string(kMaxFrames) + " nested Translate/Rotate frames");
  if (media > 1) return fail(RT_ERR_UNSUPPORTED, "ConstantMedium nested inside a ConstantMedium boundary");
  if (lightDepth(h, h.lights_root) + 1 > kMaxLightDepth) return fail(RT_ERR_UNSUPPORTED, "lights BVH too deep");
  {  // scan-stream eligibility and size (memoised expansion count; shared subtrees are expanded once per reference)
    std::vector<uint64_t> count(h.n_nodes, 0);  // per node, 0 = not yet counted
    bool ok = (h.world_root >> 28) == K_BVH;
    std::function<uint64_t(uint32_t)> expand = [&](uint32_t ref) -> uint64_t {
      const uint32_t k = ref >> 28;
      if (k == K_SPHERE || (k >= K_RECT_XY && k <= K_RECT_YZ) || k == K_UNHITTABLE) return 1;
      if (k != K_BVH) { ok = false; return 1; }
      const uint32_t i = ref & kIdxMask;
      if (count[i]) return count[i];
      const uint64_t c = 1 + expand(h.node_left[i]) + expand(h.node_right[i]);
      count[i] = c;
      return c;
    };
    const uint64_t total = ok ? expand(h.world_root) : 0;
    const uint32_t lk = h.lights_root >> 28;
    ok = ok && (lk == K_UNHITTABLE || lk == K_SPHERE || (lk >= K_RECT_XY && lk <= K_RECT_YZ));
    ok = ok && total <= 4ull * ((uint64_t)h.n_nodes + h.n_spheres + h.n_rects) + 16 && total < (1ull << 30);
    h.coopEligible = ok; h.coopRecords = ok ? total : 0;
  }
  h.stackDepth = stackNeed;
  h.lightCap = (h.lights_root >> 28) == K_BVH ? (uint32_t)lightDepth(h, h.lights_root) : 0u;  // DFS over L levels holds at most L entries
  h.node_lsize.resize(h.n_nodes); h.node_rsize.resize(h.n_nodes);
  for (uint32_t i = 0; i < h.n_nodes; ++i) { h.node_lsize[i] = (uint32_t)an.size(h.node_left[i]); h.node_rsize[i] = (uint32_t)an.size(h.node_right[i]); }

  // ---- device
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) { cudaGetLastError(); return fail(RT_ERR_CUDA, "no CUDA device available (this library has no CPU path)"); }
  if (device < 0) { CU(cudaGetDevice(&device)); }
  if (device >= ndev) return fail(RT_ERR_ARG, "device ordinal out of range");
  CU(cudaSetDevice(device));
  e->main.device = device;
  CU(cudaDeviceGetAttribute(&e->main.numSMs, cudaDevAttrMultiProcessorCount, device));
  int rc = buildDevice<float>(h, e->main.f32);  // the production layout is uploaded eagerly; fp64 on first use
  if (rc) return rc;
  std::lock_guard<std::mutex> lk(g_mu);
  *out = g_nextId++;
  g_scenes[*out] = std::move(e);
  return RT_OK;
}

int rt_scene_destroy(rt_scene_t scene) {
  std::shared_ptr<SceneEntry> e;
  {
    std::lock_guard<std::mutex> lk(g_mu);
    auto it = g_scenes.find(scene);
    if (it == g_scenes.end()) return fail(RT_ERR_HANDLE, "unknown scene handle");
    e = std::move(it->second);
    g_scenes.erase(it);
  }
  {
    std::lock_guard<std::mutex> lk(e->mu);  // a call in flight on another thread finishes first
    for (auto& kv : e->replicas) { cudaSetDevice(kv.first); cudaDeviceSynchronize(); }
    cudaSetDevice(e->main.device);
    cudaDeviceSynchronize();                // and so does asynchronous work of rt_render_device
  }
  e.reset();  // the buffers go with the last holder
  return RT_OK;
}

int rt_render(rt_scene_t scene, const RtRenderParams* params, double* out, RtStats* stats) { return renderHost(scene, params, out, stats, 0); }
int rt_render_counted(rt_scene_t scene, const RtRenderParams* params, double* out, RtStats* stats) { return renderHost(scene, params, out, stats, 1); }
int rt_render_pathlog(rt_scene_t scene, const RtRenderParams* params, double* out, uint32_t max_bounces, uint32_t* prim_log, double* t_log) {
  if (!prim_log || !t_log || max_bounces == 0) return fail(RT_ERR_ARG, "path log buffers are null / empty");
  if (params && (params->flags & RT_FLAG_FAST_MATH)) return fail(RT_ERR_ARG, "the path log is not built for RT_FLAG_FAST_MATH");
  PathLog log; log.prim = prim_log; log.t = t_log; log.stride = max_bounces;
  return renderHost(scene, params, out, nullptr, 3, &log);
}

int rt_render_device(rt_scene_t scene, const RtRenderParams* params, double* d_out, void* cuda_stream, RtStats* stats) {
  const std::shared_ptr<SceneEntry> ep = lookup(scene);
  if (!ep) return fail(RT_ERR_HANDLE, "unknown scene handle");
  SceneEntry* e = ep.get();
  std::lock_guard<std::mutex> lk(e->mu);
  if (!d_out) return fail(RT_ERR_ARG, "d_out_rgb_sum is null");
  RtRenderParams q;
  int rc = normaliseParams(params, q);
  if (rc) return rc;
  CU(cudaSetDevice(e->main.device));
  cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);
  Events evs;
  if (stats) { CU(evs.create(2)); CU(cudaEventRecord(evs.ev[0], st)); }
  const cudaEvent_t e0 = evs.ev[0], e1 = evs.ev[1];
  uint32_t kernelKind = 0;
  rc = renderLaunch(e->hs, &e->main, q, d_out, st, 0, nullptr, nullptr, &kernelKind);
  if (rc) return rc;
  if (stats) {
    CU(cudaEventRecord(e1, st));
    CU(cudaEventSynchronize(e1));
    std::memset(stats, 0, sizeof(*stats));
    float ms = 0; cudaEventElapsedTime(&ms, e0, e1);
    stats->kernel_ms = ms; stats->samples = (uint64_t)(q.row1 - q.row0) * (q.col1 - q.col0) * q.sample_count; stats->launches = q.sample_count ? 1 : 0;
    stats->gpus = 1; stats->gpu_kernel_ms[0] = ms; stats->kernel_kind = kernelKind;
    if (kernelKind == 2) { stats->rounds = (uint32_t)e->main.wfRounds; stats->launches = 1ull + 2ull * (uint64_t)e->main.wfRounds; }
  }
  return RT_OK;
}

// No cudaMalloc / cudaFree / event creation on this path: memsets, one renderKernel launch and one 2-D device copy, all on the
// caller's stream.  The device copy of the scene in the requested arithmetic must already exist (rt_scene_create uploads fp32 /
// mixed; fp64 is uploaded by the first fp64 call of another entry point).
int rt_render_device_static(rt_scene_t scene, const RtRenderParams* params, double* d_out, void* cuda_stream) {
  const std::shared_ptr<SceneEntry> ep = lookup(scene);
  if (!ep) return fail(RT_ERR_HANDLE, "unknown scene handle");
  SceneEntry* e = ep.get();
  std::lock_guard<std::mutex> lk(e->mu);
  if (!d_out) return fail(RT_ERR_ARG, "d_out_rgb_sum is null");
  RtRenderParams q;
  int rc = normaliseParams(params, q);
  if (rc) return rc;
  if ((uint64_t)q.width * q.height > RT_STATIC_MAX_PIXELS)
    return fail(RT_ERR_UNSUPPORTED, "rt_render_device_static: width*height = " + std::to_string((uint64_t)q.width * q.height) + " pixels exceeds RT_STATIC_MAX_PIXELS = " +
                                        std::to_string((uint64_t)RT_STATIC_MAX_PIXELS) + " (the static scratch framebuffer); use rt_render_device");
  if (q.flags & (RT_FLAG_COOP | RT_FLAG_WAVEFRONT)) return fail(RT_ERR_UNSUPPORTED, "rt_render_device_static runs renderKernel only (no RT_FLAG_COOP / RT_FLAG_WAVEFRONT)");
  if (q.n_gpus > 1) return fail(RT_ERR_UNSUPPORTED, "rt_render_device_static renders on the scene's device only");
  DeviceCtx* c = &e->main;
  const void* dsc = nullptr;
  if (q.precision == RT_PREC_F64) { if (c->f64.ready) dsc = &c->f64.sc; }
  else if (c->f32.ready) dsc = &c->f32.sc;
  if (!dsc) return fail(RT_ERR_UNSUPPORTED, "rt_render_device_static: the scene is not on the device in this arithmetic yet (uploading allocates): render it once through rt_render / rt_render_device first");
  std::lock_guard<std::mutex> lk2(g_staticMu);
  CU(cudaSetDevice(c->device));
  unsigned long long* scratch = nullptr;
  if ((rc = staticScratch(c->device, &scratch))) return rc;
  double* fb = reinterpret_cast<double*>(scratch + kStaticCtrWords);
  cudaStream_t st = static_cast<cudaStream_t>(cuda_stream);
  const size_t rowElems = (size_t)q.width * 3;
  // the array keeps the last call's contents: the counter block and the rows this call accumulates into start at zero
  CU(cudaMemsetAsync(scratch, 0, kStaticCtrWords * sizeof(unsigned long long), st));
  CU(cudaMemsetAsync(fb + q.row0 * rowElems, 0, (size_t)(q.row1 - q.row0) * rowElems * sizeof(double), st));
  if (q.sample_count) {
    const HostScene& hs = e->hs;
    RenderArgs a{};  // the same launch parameters as renderLaunch's renderKernel branch (results never depend on the scheduler knobs)
    a.W = q.width; a.H = q.height; a.s0 = q.sample_offset; a.ns = q.sample_count; a.maxDepth = q.max_depth;
    a.row0 = q.row0; a.row1 = q.row1; a.col0 = q.col0; a.col1 = q.col1; a.rowStep = 1;
    a.seed = q.seed;
    a.total_paths = (unsigned long long)(q.row1 - q.row0) * (q.col1 - q.col0) * q.sample_count;
    const unsigned long long warps = (unsigned long long)c->numSMs * 16ull;
    unsigned long long chunk = 1024;
    while (chunk > 4 && a.total_paths / chunk < warps * 8ull) chunk /= 2;
    a.chunk = (uint32_t)chunk;
    a.chunkMin = (uint32_t)std::min<unsigned long long>(chunk, 32);
    a.chunkDiv = (uint32_t)std::max<unsigned long long>(1, warps * 8);
    a.shadeMin = 16; a.rectMin = hs.n_cuboids > hs.n_rects ? 6u : 3u; a.rareMin = hs.n_media > 0 ? 4u : 8u; a.keepNode = hs.n_nodes <= 64 ? 8 : 12; a.keepLeaf = hs.n_nodes <= 64 ? 6 : 8;
    a.lightCap = hs.lightCap;
    a.counter = scratch; a.ctr = nullptr; a.out = fb;
    cudaError_t ce;
    if (q.flags & RT_FLAG_FAST_MATH) ce = launchRender<MATH_FAST>((int)q.precision, 0, dsc, a, hs.stackDepth, c->numSMs, st, nullptr);
    else if (q.flags & RT_FLAG_STRICT_MATH) ce = launchRender<MATH_STRICT>((int)q.precision, 0, dsc, a, hs.stackDepth, c->numSMs, st, nullptr);
    else ce = launchRender<MATH_LITERAL>((int)q.precision, 0, dsc, a, hs.stackDepth, c->numSMs, st, nullptr);
    if (ce != cudaSuccess) return fail(RT_ERR_CUDA, std::string("render launch: ") + cudaGetErrorString(ce));
  }
  // rt_render's contract on a device buffer: the rendered region is OVERWRITTEN with this call's sums, other pixels stay untouched
  const size_t off = q.row0 * rowElems + (size_t)q.col0 * 3;
  CU(cudaMemcpy2DAsync(d_out + off, rowElems * sizeof(double), fb + off, rowElems * sizeof(double), (size_t)(q.col1 - q.col0) * 3 * sizeof(double), q.row1 - q.row0,
                       cudaMemcpyDeviceToDevice, st));
  return RT_OK;
}

int rt_trace(rt_scene_t scene, uint32_t precision, uint32_t flags, uint32_t n, const double* rays, double t_min, double t_max, uint64_t seed,
             uint32_t log_stride, int32_t* hit, double* rec, uint32_t* prim, uint32_t* mat, uint8_t* front, uint64_t* nvisit, uint64_t* vhash, uint32_t* vlog) {
  const std::shared_ptr<SceneEntry> ep = lookup(scene);
  if (!ep) return fail(RT_ERR_HANDLE, "unknown scene handle");
  SceneEntry* e = ep.get();
  std::lock_guard<std::mutex> lk(e->mu);
  if (!rays && n) return fail(RT_ERR_ARG, "rays is null");
  if (n == 0) return RT_OK;
  if ((flags & RT_FLAG_FAST_MATH) && (flags & RT_FLAG_STRICT_MATH)) return fail(RT_ERR_ARG, "RT_FLAG_FAST_MATH and RT_FLAG_STRICT_MATH exclude each other");
  CU(cudaSetDevice(e->main.device));
  const void* dsc = nullptr;
  int rc = deviceScene(e->hs, &e->main, precision, &dsc);
  if (rc) return rc;
  const uint32_t stride = log_stride ? log_stride : 1;
  struct Buf { void* p = nullptr; ~Buf() { if (p) cudaFree(p); } };
  Buf bRays, bHit, bRec, bPrim, bMat, bFront, bNv, bVh, bLog;
  CU(cudaMalloc(&bRays.p, (size_t)n * 7 * sizeof(double)));
  CU(cudaMalloc(&bHit.p, (size_t)n * 4)); CU(cudaMalloc(&bRec.p, (size_t)n * 9 * sizeof(double))); CU(cudaMalloc(&bPrim.p, (size_t)n * 4));
  CU(cudaMalloc(&bMat.p, (size_t)n * 4)); CU(cudaMalloc(&bFront.p, (size_t)n)); CU(cudaMalloc(&bNv.p, (size_t)n * 8)); CU(cudaMalloc(&bVh.p, (size_t)n * 8));
  CU(cudaMalloc(&bLog.p, (size_t)n * stride * 4));
  CU(cudaMemcpy(bRays.p, rays, (size_t)n * 7 * sizeof(double), cudaMemcpyHostToDevice));
  TraceArgs a{};
  a.n = n; a.logStride = log_stride; a.tmin = t_min; a.tmax = t_max; a.seed = seed; a.rays = (const double*)bRays.p;
  a.hit = (int32_t*)bHit.p; a.rec = (double*)bRec.p; a.prim = (uint32_t*)bPrim.p; a.mat = (uint32_t*)bMat.p; a.front = (uint8_t*)bFront.p;
  a.nvisit = (unsigned long long*)bNv.p; a.vhash = (unsigned long long*)bVh.p; a.vlog = (uint32_t*)bLog.p;
  if ((flags & RT_FLAG_COOP) && !e->hs.coopEligible) return fail(RT_ERR_UNSUPPORTED, "RT_FLAG_COOP: the scene is not eligible for the scan-stream traversal");
  const bool coop = useCoop(e->hs, flags);
  cudaError_t ce = coop ? ((flags & RT_FLAG_STRICT_MATH) ? launchTraceCoop<MATH_STRICT>((int)precision, dsc, a, 0) : launchTraceCoop<MATH_LITERAL>((int)precision, dsc, a, 0))
                   : (flags & RT_FLAG_FAST_MATH) ? launchTrace<MATH_FAST>((int)precision, dsc, a, e->hs.stackDepth, 0)
                   : (flags & RT_FLAG_STRICT_MATH) ? launchTrace<MATH_STRICT>((int)precision, dsc, a, e->hs.stackDepth, 0)
                                                   : launchTrace<MATH_LITERAL>((int)precision, dsc, a, e->hs.stackDepth, 0);
  if (ce != cudaSuccess) return fail(RT_ERR_CUDA, std::string("trace launch: ") + cudaGetErrorString(ce));
  CU(cudaDeviceSynchronize());
  if (hit) CU(cudaMemcpy(hit, bHit.p, (size_t)n * 4, cudaMemcpyDeviceToHost));
  if (rec) CU(cudaMemcpy(rec, bRec.p, (size_t)n * 9 * sizeof(double), cudaMemcpyDeviceToHost));
  if (prim) CU(cudaMemcpy(prim, bPrim.p, (size_t)n * 4, cudaMemcpyDeviceToHost));
  if (mat) CU(cudaMemcpy(mat, bMat.p, (size_t)n * 4, cudaMemcpyDeviceToHost));
  if (front) CU(cudaMemcpy(front, bFront.p, (size_t)n, cudaMemcpyDeviceToHost));
  if (nvisit) CU(cudaMemcpy(nvisit, bNv.p, (size_t)n * 8, cudaMemcpyDeviceToHost));
  if (vhash) CU(cudaMemcpy(vhash, bVh.p, (size_t)n * 8, cudaMemcpyDeviceToHost));
  if (vlog && log_stride) CU(cudaMemcpy(vlog, bLog.p, (size_t)n * log_stride * 4, cudaMemcpyDeviceToHost));
  return RT_OK;
}

int rt_texture_value(rt_scene_t scene, uint32_t tex, uint32_t precision, uint32_t flags, uint32_t n, const double* q, double* out) {
  const std::shared_ptr<SceneEntry> ep = lookup(scene);
  if (!ep) return fail(RT_ERR_HANDLE, "unknown scene handle");
  SceneEntry* e = ep.get();
  std::lock_guard<std::mutex> lk(e->mu);
  if (tex >= e->hs.n_textures) return fail(RT_ERR_ARG, "texture index out of range");
  if (n == 0) return RT_OK;
  if (!q || !out) return fail(RT_ERR_ARG, "q/out is null");
  CU(cudaSetDevice(e->main.device));
  const void* dsc = nullptr;
  int rc = deviceScene(e->hs, &e->main, precision, &dsc);
  if (rc) return rc;
  double *dq = nullptr, *dout = nullptr;
  CU(cudaMalloc(&dq, (size_t)n * 5 * sizeof(double))); CU(cudaMalloc(&dout, (size_t)n * 3 * sizeof(double)));
  CU(cudaMemcpy(dq, q, (size_t)n * 5 * sizeof(double), cudaMemcpyHostToDevice));
  cudaError_t ce = (flags & RT_FLAG_STRICT_MATH) ? launchTexture<MATH_STRICT>((int)precision, dsc, tex, n, dq, dout, 0) : launchTexture<MATH_LITERAL>((int)precision, dsc, tex, n, dq, dout, 0);
  if (ce == cudaSuccess) ce = cudaMemcpy(out, dout, (size_t)n * 3 * sizeof(double), cudaMemcpyDeviceToHost);
  cudaFree(dq); cudaFree(dout);
  if (ce != cudaSuccess) return fail(RT_ERR_CUDA, std::string("texture probe: ") + cudaGetErrorString(ce));
  return RT_OK;
}

int rt_tonemap(const double* rgb_sum, uint32_t n, uint32_t total_samples, uint8_t* out_rgb8) {
  if (n == 0) return RT_OK;
  if (!rgb_sum || !out_rgb8) return fail(RT_ERR_ARG, "null buffer");
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) { cudaGetLastError(); return fail(RT_ERR_CUDA, "no CUDA device available (this library has no CPU path)"); }
  double* din = nullptr; uint8_t* dout = nullptr;
  CU(cudaMalloc(&din, (size_t)n * sizeof(double))); CU(cudaMalloc(&dout, n));
  CU(cudaMemcpy(din, rgb_sum, (size_t)n * sizeof(double), cudaMemcpyHostToDevice));
  cudaError_t ce = launchTonemap(din, n, (double)total_samples, dout, 0);
  if (ce == cudaSuccess) ce = cudaMemcpy(out_rgb8, dout, n, cudaMemcpyDeviceToHost);
  cudaFree(din); cudaFree(dout);
  if (ce != cudaSuccess) return fail(RT_ERR_CUDA, std::string("tonemap: ") + cudaGetErrorString(ce));
  return RT_OK;
}

int rt_selftest_division(uint64_t n, uint64_t seed, uint64_t* checked, uint64_t* mismatches) {
  if (!checked || !mismatches) return fail(RT_ERR_ARG, "null output");
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) { cudaGetLastError(); return fail(RT_ERR_CUDA, "no CUDA device available (this library has no CPU path)"); }
  unsigned long long* d = nullptr;
  CU(cudaMalloc(&d, 2 * sizeof(unsigned long long)));
  CU(cudaMemset(d, 0, 2 * sizeof(unsigned long long)));
  cudaError_t ce = launchDivSelftest(n, seed, d, 0);
  unsigned long long h[2] = {0, 0};
  if (ce == cudaSuccess) ce = cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
  cudaFree(d);
  if (ce != cudaSuccess) return fail(RT_ERR_CUDA, std::string("division selftest: ") + cudaGetErrorString(ce));
  *checked = h[0]; *mismatches = h[1];
  return RT_OK;
}

int rt_selftest_box_filter(uint64_t n, uint64_t seed, uint64_t out3[3]) {
  if (!out3) return fail(RT_ERR_ARG, "null output");
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) { cudaGetLastError(); return fail(RT_ERR_CUDA, "no CUDA device available (this library has no CPU path)"); }
  unsigned long long* d = nullptr;
  CU(cudaMalloc(&d, 3 * sizeof(unsigned long long)));
  CU(cudaMemset(d, 0, 3 * sizeof(unsigned long long)));
  cudaError_t ce = launchBoxFilterSelftest(n, seed, d, 0);
  unsigned long long h[3] = {0, 0, 0};
  if (ce == cudaSuccess) ce = cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
  cudaFree(d);
  if (ce != cudaSuccess) return fail(RT_ERR_CUDA, std::string("box filter selftest: ") + cudaGetErrorString(ce));
  out3[0] = h[0]; out3[1] = h[1]; out3[2] = h[2];
  return RT_OK;
}

int rt_strict_math_eval(uint32_t fn, uint32_t n, const double* x, const double* y, double* out, int as_float) {
  if (n == 0) return RT_OK;
  if (!x || !y || !out || fn > 4) return fail(RT_ERR_ARG, "bad argument");
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) { cudaGetLastError(); return fail(RT_ERR_CUDA, "no CUDA device available (this library has no CPU path)"); }
  struct Buf { void* p = nullptr; ~Buf() { if (p) cudaFree(p); } } dx, dy, dout;
  CU(cudaMalloc(&dx.p, (size_t)n * 8)); CU(cudaMalloc(&dy.p, (size_t)n * 8)); CU(cudaMalloc(&dout.p, (size_t)n * 8));
  CU(cudaMemcpy(dx.p, x, (size_t)n * 8, cudaMemcpyHostToDevice)); CU(cudaMemcpy(dy.p, y, (size_t)n * 8, cudaMemcpyHostToDevice));
  cudaError_t ce = launchStrictMathEval(fn, n, (const double*)dx.p, (const double*)dy.p, (double*)dout.p, as_float, 0);
  if (ce == cudaSuccess) ce = cudaMemcpy(out, dout.p, (size_t)n * 8, cudaMemcpyDeviceToHost);
  if (ce != cudaSuccess) return fail(RT_ERR_CUDA, std::string("strict math eval: ") + cudaGetErrorString(ce));
  return RT_OK;
}

void rt_shutdown(void) {
  {
    std::lock_guard<std::mutex> lk(g_mu);
    g_scenes.clear();
  }
  g_scratch.clear();
  std::lock_guard<std::mutex> lk(g_multi.mu);
  for (void* c : g_multi.comms) if (c && g_multi.commDestroy) g_multi.commDestroy(c);
  g_multi.comms.clear(); g_multi.wantNccl = false; g_multi.defaultGpus = 1;
}

}  // extern "C"
